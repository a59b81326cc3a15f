"""Guard bands around every buffer the kernels write.

compute-sanitizer is closed on this GPU pool (profiles/r02_sanitizer_unavailable.log), so
out-of-bounds WRITES are looked for the direct way: every written tensor (outputs, KV pool, score
pool, split-KV workspace, compacted tables, victims, pruned K/V) is a window in the middle of a
larger allocation filled with a byte pattern; after the op the bands on both sides must be intact.
Writes inside the windows are already compared bit for bit against the oracle by the kernel tests.
Shapes are picked to hit the delicate paths: split groups (tickets + partials), teams, the fused
cache write into a fresh block, partial last tiles, the element-wise cache-write path."""
import numpy as np
import pytest
import torch

from oracle import paged_ops, policy
from oracle.policy import Policy
from tests.util import rnd

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
BAND = 1 << 16          # bytes on each side
PAT = 0x5A


class Guarded:
    """A CUDA tensor that is a window of a larger patterned allocation."""

    def __init__(self, shape, dtype, fill=None):
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        self.n = n
        self.raw = torch.full((BAND + n + BAND,), PAT, dtype=torch.uint8, device=DEV)
        self.t = self.raw[BAND:BAND + n].view(dtype).view(*shape)
        if fill is not None:
            self.t.copy_(fill if isinstance(fill, torch.Tensor) else torch.full_like(self.t, fill))

    def intact(self):
        lo, hi = self.raw[:BAND], self.raw[BAND + self.n:]
        return bool((lo == PAT).all()) and bool((hi == PAT).all())


def _ops():
    from vllm_paged_compression_b200 import ops
    return ops


@pytest.mark.parametrize("B,max_len,Hq,Hkv,D,bs", [(3, 2100, 32, 8, 128, 16),      # few groups: split KV, tickets, partials
                                                    (64, 600, 32, 8, 128, 16),      # the bench shape: teams of 2
                                                    (130, 300, 8, 1, 128, 16),      # TP-8 per-rank shape
                                                    (9, 700, 8, 2, 64, 8),          # block 8
                                                    (5, 900, 16, 2, 128, 32)])
def test_decode_layer_writes_stay_inside(B, max_len, Hq, Hkv, D, bs):
    ops = _ops()
    rs = np.random.RandomState(B + max_len)
    kv_lens = rs.randint(max(1, max_len // 2), max_len + 1, size=B).astype(np.int32)
    kv_lens[0] = (max_len // bs) * bs                 # a block that fills this step (scored)
    kv_lens[1 % B] = (max_len // bs - 1) * bs + 1     # a token that opens a fresh block
    n = [policy.cdiv(int(x), bs) for x in kv_lens]
    M = max(n)
    nb = sum(n) + 3
    perm = rs.permutation(nb)
    tables = np.zeros((B, M), np.int32)
    o = 0
    for r in range(B):
        tables[r, :n[r]] = perm[o:o + n[r]]
        o += n[r]
    slots = np.array([int(tables[r, n[r] - 1]) * bs + (int(kv_lens[r]) - 1) % bs for r in range(B)], np.int64)
    scale = D ** -0.5
    kv = Guarded((2, nb, bs * Hkv * D), torch.bfloat16,
                 fill=(rnd(7, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16).to(DEV))
    pool = Guarded((nb, Hkv), torch.float32, fill=0.0)
    out = Guarded((B, Hq, D), torch.bfloat16, fill=0.0)
    need = ops.paged_decode_workspace_bytes(B, Hq, Hkv, D, bs, int(kv_lens.max()))
    ws = Guarded((need,), torch.uint8, fill=0)
    assert ws.t.data_ptr() % 256 == 0
    q = (rnd(8, B, Hq, D).float() * 0.5).to(torch.bfloat16).to(DEV)
    key = (rnd(9, B, Hkv, D).float() * 0.5).to(torch.bfloat16).to(DEV)
    val = rnd(10, B, Hkv, D).to(DEV)
    kc, vc = paged_ops.split_kv_cache(kv.t, Hkv, D)
    st = ops.make_decode_step(torch.from_numpy(slots).to(DEV), torch.from_numpy(tables).to(DEV),
                              torch.from_numpy(kv_lens).to(DEV), torch.from_numpy(kv_lens).to(DEV), B, Hq, Hkv, D,
                              bs, int(kv_lens.max()), scale, q.stride(0), key.stride(0), val.stride(0),
                              out.t.stride(0), vc.stride(0), fuse_cache_write=True, workspace=ws.t)
    for _ in range(3):        # the workspace must come back clean: three launches in a row
        ops.decode_layer(st, out.t, q, key, val, kc.data_ptr(), vc.data_ptr(), ops.make_score_pool(pool.t, 0), None)
    torch.cuda.synchronize()
    for name, g in (("kv pool", kv), ("score pool", pool), ("out", out), ("workspace", ws)):
        assert g.intact(), f"write outside the {name}"
    assert torch.isfinite(out.t.float()).all()
    assert int(ws.t[:256 + (1 << 22)].to(torch.int32).sum()) == 0, "counters / tickets not re-armed"


@pytest.mark.parametrize("Hkv,D,bs", [(8, 128, 16), (1, 128, 16), (4, 64, 32), (8, 128, 8)])
def test_prompt_time_writes_stay_inside(Hkv, D, bs):
    """prune_prompt_and_cache (streaming / block / element-wise paths), the prefill kernel and the
    token-level prune."""
    ops = _ops()
    Hq = Hkv * 4
    pol = Policy(block_size=bs, cache_prune_type="budget", cache_budget=8 * bs)
    lens = [5, 8 * bs, 8 * bs + 1, 20 * bs + 3, 33 * bs]
    T = sum(lens)
    k = (rnd(21, T, Hkv, D).float() * 0.5).to(torch.bfloat16).to(DEV)
    v = rnd(22, T, Hkv, D).to(DEV)
    q = (rnd(23, T, Hq, D).float() * 0.5).to(torch.bfloat16).to(DEV)
    m5, m6, toff, slots = [], [], [0], []
    src = dst = blk = 0
    for L in lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        nblk = policy.cdiv(kept, bs)
        m5.append((src, kept, first, pruned, dst))
        m6.append((src, L, src, kept, first, pruned))
        toff.append(toff[-1] + nblk)
        slots.extend(range(blk * bs, blk * bs + kept))
        src, dst, blk = src + L, dst + kept, blk + nblk
    nb = blk
    kv = Guarded((2, nb, bs * Hkv * D), torch.bfloat16, fill=0.0)
    pool = Guarded((nb, Hkv), torch.float32, fill=0.0)
    out = Guarded((T, Hq, D), torch.bfloat16, fill=0.0)
    kc, vc = paged_ops.split_kv_cache(kv.t, Hkv, D)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    ops.prune_prompt_and_cache(k, v, kc, vc, torch.tensor(slots, dtype=torch.int64, device=DEV), i32(m5), i32(toff),
                               toff[-1], ops.make_score_pool(pool.t, 0))
    m6h = np.asarray(m6, np.int32)
    ops.prefill_attn_topleft_causal(out.t, q, k, v, torch.from_numpy(m6h).to(DEV), max(lens), D ** -0.5,
                                    req_meta_host=m6h)
    # token-level prune of the same batch into guarded outputs
    from oracle import token_prune
    tp6, moff, qoff = [], [0], [0]
    src = dst = 0
    for L in lens:
        first, mid, kd = token_prune.prune_layout(L, pol.cache_budget, bs)
        tp6.append((src, L, first, mid, kd, dst))
        moff.append(moff[-1] + mid)
        qoff.append(qoff[-1] + L)
        src, dst = src + L, dst + L - kd
    pk = Guarded((dst, Hkv, D), torch.bfloat16, fill=0.0)
    pv = Guarded((dst, Hkv, D), torch.bfloat16, fill=0.0)
    ops.prompt_token_prune(k, v, pk.t, pv.t, i32(tp6), i32(moff), i32(qoff), qoff[-1], moff[-1], "value_l2")
    torch.cuda.synchronize()
    for name, g in (("kv pool", kv), ("score pool", pool), ("prefill out", out), ("pruned K", pk), ("pruned V", pv)):
        assert g.intact(), f"write outside the {name}"
    assert torch.isfinite(out.t.float()).all()


def test_selection_and_zeroing_writes_stay_inside():
    ops = _ops()
    from vllm_paged_compression_b200.config import PagedEvictConfig
    B, L, M, bs, Hkv, D, nb = 37, 3, 23, 16, 2, 64, 900
    rs = np.random.RandomState(5)
    nblk = rs.randint(3, M + 1, size=B).astype(np.int32)
    tables = np.zeros((B, M), np.int32)
    for r in range(B):
        tables[r, :nblk[r]] = rs.permutation(nb)[:nblk[r]]
    seq = (nblk.astype(np.int64) * bs + 4 * bs).astype(np.int32)      # over budget, on a block boundary
    pe = PagedEvictConfig(cache_prune_type="budget", evict_method="global", cache_budget=128)
    pol = ops.make_policy(pe, bs)
    pools = [torch.rand(nb, Hkv, device=DEV) for _ in range(L)]
    ptrs = torch.tensor([p.data_ptr() for p in pools], dtype=torch.int64, device=DEV)
    out_tab = Guarded((L, B, M), torch.int32, fill=0)
    out_vic = Guarded((L, B), torch.int32, fill=-1)
    ops.evict_select_compact_layers(torch.from_numpy(tables).to(DEV), torch.from_numpy(nblk).to(DEV),
                                    torch.from_numpy(seq).to(DEV), pol, ptrs, Hkv, out_tab.t, out_vic.t)
    kvs = [Guarded((2, nb, bs * Hkv * D), torch.bfloat16, fill=1.0) for _ in range(L)]
    kptr = torch.tensor([g.t[0].data_ptr() for g in kvs], dtype=torch.int64, device=DEV)
    vptr = torch.tensor([g.t[1].data_ptr() for g in kvs], dtype=torch.int64, device=DEV)
    zs = torch.from_numpy(rs.randint(-1, nb * bs, size=(L, B)).astype(np.int64)).to(DEV)
    zs[0, 0], zs[1, 1] = 0, nb * bs - 1                               # first and last slot of the pool
    ops.zero_cache_slots_layers(kptr, vptr, zs, Hkv, D, bs, bs * Hkv * D)
    torch.cuda.synchronize()
    assert out_tab.intact() and out_vic.intact()
    assert int((out_vic.t >= 1).sum()) == L * int((seq > 128).sum())      # every over-budget row of every layer
    for l, g in enumerate(kvs):
        assert g.intact(), f"zeroing wrote outside pool {l}"
        zeroed = int((g.t[0].view(nb * bs, -1, 8)[..., 0] == 0).sum())       # loose: at least one slot zeroed
        assert zeroed > 0
