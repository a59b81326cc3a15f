"""GPU parity tests: every kernel through the C ABI (vllm_paged_compression_b200.ops)
against the CPU oracle on the same seeded inputs.  Integer/byte/index results are
bit-exact; attention outputs within the bf16 tolerance of BASELINE.json (atol 2e-2; the
tests use the tighter bars noted inline where the arithmetic allows).

Mirrors the reference's own kernel tests: tests/kernels/test_cache.py:121-205
(reshape_and_cache, exact) and tests/kernels/test_attention.py:122-330 (paged attention
vs ref_single_query_cached_kv_attention).
"""
import numpy as np
import pytest
import torch

from oracle import paged_ops, policy, scoring
from oracle.policy import Policy
from tests.util import load_npz, rnd

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _ops():
    from vllm_paged_compression_b200 import ops
    return ops


def _pool(seed, nb, bs, Hkv, D, scale=0.5):
    kv = (rnd(seed, 2, nb, bs * Hkv * D).float() * scale).to(torch.bfloat16)
    return kv


def _bits(t):
    return t.detach().cpu().contiguous().view(torch.int16).numpy()


# --------------------------------------------------------------------------- W1
@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("bs", [8, 16, 32])
@pytest.mark.parametrize("n_tok", [1, 7, 200, 1000])
def test_reshape_and_cache_exact(D, bs, n_tok):
    ops = _ops()
    Hkv, nb = 4, 96 if bs >= 16 else 192
    rs = np.random.RandomState(n_tok * 7 + D + bs)
    slots = rs.permutation(nb * bs)[:n_tok].astype(np.int64)
    slots[rs.rand(n_tok) < 0.1] = -1                      # PAD_SLOT_ID rows are skipped
    key, value = rnd(1, n_tok, Hkv, D), rnd(2, n_tok, Hkv, D)
    kv_ref = _pool(3, nb, bs, Hkv, D)
    kv_dev = kv_ref.clone().to(DEV)
    kc, vc = paged_ops.split_kv_cache(kv_ref, Hkv, D)
    paged_ops.reshape_and_cache(key, value, kc, vc, torch.from_numpy(slots))
    kcd, vcd = paged_ops.split_kv_cache(kv_dev, Hkv, D)
    ops.reshape_and_cache(key.to(DEV), value.to(DEV), kcd, vcd, torch.from_numpy(slots).to(DEV))
    torch.cuda.synchronize()
    assert np.array_equal(_bits(kv_dev), _bits(kv_ref))


def test_reshape_and_cache_contiguous_block_fast_path_and_strided_rows():
    """Whole-block tiles take the vectorised V path; key/value are column slices of a
    fused qkv activation (row stride > Hkv*D) as in llama.py:195-207."""
    ops = _ops()
    Hkv, D, bs, nb = 8, 128, 16, 64
    n_tok = 5 * bs
    qkv = rnd(11, n_tok, (32 + 2 * Hkv) * D)
    key = qkv[:, 32 * D:(32 + Hkv) * D]
    value = qkv[:, (32 + Hkv) * D:]
    blocks = [9, 3, 40, 41, 7]
    slots = torch.tensor([b * bs + i for b in blocks for i in range(bs)], dtype=torch.int64)
    kv_ref = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16)
    kv_dev = kv_ref.clone().to(DEV)
    kc, vc = paged_ops.split_kv_cache(kv_ref, Hkv, D)
    paged_ops.reshape_and_cache(key.reshape(n_tok, Hkv, D), value.reshape(n_tok, Hkv, D), kc, vc, slots)
    qkv_d = qkv.to(DEV)
    kcd, vcd = paged_ops.split_kv_cache(kv_dev, Hkv, D)
    ops.reshape_and_cache(qkv_d[:, 32 * D:(32 + Hkv) * D], qkv_d[:, (32 + Hkv) * D:], kcd, vcd,
                          slots.to(DEV))
    torch.cuda.synchronize()
    assert np.array_equal(_bits(kv_dev), _bits(kv_ref))


def test_reshape_and_cache_empty_and_errors():
    ops = _ops()
    Hkv, D, bs, nb = 2, 64, 16, 4
    kv = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16, device=DEV)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    e = torch.empty(0, Hkv, D, dtype=torch.bfloat16, device=DEV)
    ops.reshape_and_cache(e, e, kc, vc, torch.empty(0, dtype=torch.int64, device=DEV))   # no-op
    with pytest.raises(RuntimeError):      # fp8 / scales rejected like torch_cuda_sdpa.py:769-772
        ops.reshape_and_cache(e, e, kc, vc, torch.empty(0, dtype=torch.int64, device=DEV), "fp8")
    kv96 = torch.zeros(2, nb, bs * Hkv * 96, dtype=torch.bfloat16, device=DEV)
    kc96, vc96 = paged_ops.split_kv_cache(kv96, Hkv, 96)
    k96 = torch.zeros(1, Hkv, 96, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size"):
        ops.reshape_and_cache(k96, k96, kc96, vc96, torch.zeros(1, dtype=torch.int64, device=DEV))


def test_reshape_and_cache_golden_from_reference_cpu_kernel():
    """tests/golden/ref_ops.npz was written by the reference's csrc/cpu/cache.cpp."""
    ops = _ops()
    g = load_npz("ref_ops.npz")
    nb, bs, Hkv, D = 48, 16, 2, 64
    kv = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16, device=DEV)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    ops.reshape_and_cache(rnd(2, 200, Hkv, D).to(DEV), rnd(3, 200, Hkv, D).to(DEV), kc, vc,
                          torch.from_numpy(g["rc_slots"]).to(DEV))
    torch.cuda.synchronize()
    assert np.array_equal(_bits(kv), g["rc_kv_cache"])


# --------------------------------------------------------------------- R1 + W1 + S3
def _prompt_batch(prompt_lens, pol, nb, rs):
    """Host-side metadata the backend builds for a prefill batch."""
    bs = pol.block_size
    free = list(rs.permutation(nb))
    tables, slots, meta, tile_off = [], [], [], [0]
    src = dst = 0
    for L in prompt_lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        n = policy.cdiv(kept, bs)
        tab = [int(free.pop()) for _ in range(n)]
        tables.append(tab)
        s = policy.prefill_slot_mapping(tab, L, bs)
        assert len(s) == kept
        slots.extend(s)
        meta.append([src, kept, first, pruned, dst])
        tile_off.append(tile_off[-1] + n)
        src += L
        dst += kept
    return tables, slots, meta, tile_off


@pytest.mark.parametrize("D,bs,Hkv", [(64, 16, 4), (128, 16, 4), (128, 32, 4), (128, 16, 8), (128, 16, 1),
                                      (128, 32, 8), (64, 32, 6), (128, 16, 3), (128, 8, 8), (64, 8, 4), (128, 8, 1)])
@pytest.mark.parametrize("mode", ["budget", "percentage"])
def test_prune_prompt_and_cache(D, bs, Hkv, mode):
    """R1 + W1 + S3.  Hkv 8 / 4 / 6 / 3 / 1 walk the block kernel's heads-per-CTA variants
    (8, 4, 2, 1); prompt lengths cover whole blocks, the partial last block (element-wise
    path) and prompts shorter than a block."""
    ops = _ops()
    nb = 256
    pol = Policy(block_size=bs, cache_prune_type=mode, cache_budget=8 * bs, evict_freq=2)
    prompt_lens = [5, bs, 8 * bs, 8 * bs + 1, 13 * bs + 5, 20 * bs]
    rs = np.random.RandomState(D + bs)
    tables, slots, meta, tile_off = _prompt_batch(prompt_lens, pol, nb, rs)
    T = sum(prompt_lens)
    key, value = rnd(21, T, Hkv, D), rnd(22, T, Hkv, D)
    # oracle: gather -> reshape_and_cache -> per-block scores
    kv_ref = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16)
    kc, vc = paged_ops.split_kv_cache(kv_ref, Hkv, D)
    gidx, s = [], 0
    for L in prompt_lens:
        gidx.append(torch.from_numpy(policy.prompt_gather_index(L, pol)) + s)
        s += L
    gidx = torch.cat(gidx)
    paged_ops.reshape_and_cache(key[gidx], value[gidx], kc, vc, torch.tensor(slots))
    # score pool: one row per PHYSICAL block, two extra head columns on each side to check that a
    # TP rank only touches its own columns [head_offset, head_offset + Hkv)
    HT, hoff = Hkv + 4, 2
    ref_scores = np.full((nb, HT), -1.0, dtype=np.float32)
    vp = value[gidx].float().numpy()
    for r, m in enumerate(meta):
        for j in range(m[1] // bs):
            blk = slots[m[4] + j * bs] // bs
            ref_scores[blk, hoff:hoff + Hkv] = scoring.per_head_block_score(vp[m[4] + j * bs: m[4] + (j + 1) * bs])
    # device
    kv_dev = torch.zeros_like(kv_ref, device=DEV)
    kcd, vcd = paged_ops.split_kv_cache(kv_dev, Hkv, D)
    score = torch.full((nb, HT), -1.0, dtype=torch.float32, device=DEV)
    ops.prune_prompt_and_cache(
        key.to(DEV), value.to(DEV), kcd, vcd, torch.tensor(slots, dtype=torch.int64, device=DEV),
        torch.tensor(meta, dtype=torch.int32, device=DEV),
        torch.tensor(tile_off, dtype=torch.int32, device=DEV), tile_off[-1],
        ops.make_score_pool(score, head_offset=hoff))
    torch.cuda.synchronize()
    assert np.array_equal(_bits(kv_dev), _bits(kv_ref))
    assert np.array_equal(score.cpu().numpy().view(np.int32), ref_scores.view(np.int32))   # bit-exact


# ------------------------------------------------------------------ P3 + S4 + C1 + C2
def _evict_case(B, M, pol, rs, tie=False):
    bs = pol.block_size
    n_blocks = rs.randint(1, M + 1, size=B).astype(np.int32)
    n_blocks[0] = M
    tables = np.zeros((B, M), dtype=np.int32)
    ids, o = rs.permutation(100000), 0
    for r in range(B):
        tables[r, :n_blocks[r]] = ids[o:o + n_blocks[r]] + 1          # block ids unique over the batch
        o += int(n_blocks[r])
    # logical lengths: mix of trigger / non-trigger steps
    seq_lens = np.zeros(B, dtype=np.int32)
    for r in range(B):
        base = int(n_blocks[r]) * bs + int(rs.randint(0, 40)) * bs
        seq_lens[r] = base if rs.rand() < 0.6 else base - int(rs.randint(1, bs))
    Hkv = 4
    scores = rs.rand(B, M, Hkv).astype(np.float32) * 50
    if tie:
        scores[:, 2:, :] = scores[:, 2:3, :]           # all candidates >= 2 tie; lowest index must win
    # the device keeps scores with the physical block: pool[block_id][head]
    pool = np.zeros((100002, Hkv), dtype=np.float32)
    for r in range(B):
        pool[tables[r, :n_blocks[r]]] = scores[r, :n_blocks[r]]
    return tables, n_blocks, seq_lens, scores, Hkv, pool


@pytest.mark.parametrize("method,topk", [("streamingLLM", -1), ("value_l2", -1), ("value_l2", 3)])
@pytest.mark.parametrize("mode", ["budget", "percentage"])
@pytest.mark.parametrize("tie", [False, True])
def test_evict_select_compact_bit_exact(method, topk, mode, tie):
    ops = _ops()
    from vllm_paged_compression_b200._lib import PcEvictPolicy
    bs, B, M = 16, 37, 70
    pol = Policy(block_size=bs, cache_prune_type=mode, cache_budget=128, evict_freq=2,
                 decode_evict_method=method, topk_blocks=topk)
    rs = np.random.RandomState(5 + (13 if tie else 0))
    tables, n_blocks, seq_lens, scores, Hkv, pool = _evict_case(B, M, pol, rs, tie)
    kv_lens = (n_blocks * bs - rs.randint(0, bs, size=B)).astype(np.int32)
    # oracle
    victims = {}
    for r in range(B):
        n = int(n_blocks[r])
        if n >= 3 and policy.decode_trigger(int(seq_lens[r]), n * bs, pol):
            lo, hi = policy.victim_candidates(n, pol)
            if method == "streamingLLM":
                victims[r] = 1
            else:
                victims[r] = policy.argmin_lowest_index(scoring.combine_heads(scores[r, :n - 1]), lo, hi)
    ref_tables = policy.compact_block_tables(tables, n_blocks, victims)
    ref_victim = np.array([victims.get(r, -1) for r in range(B)], dtype=np.int32)
    ref_kv = np.array([kv_lens[r] - (bs if r in victims else 0) for r in range(B)], dtype=np.int32)
    # device
    cpol = PcEvictPolicy(0 if mode == "budget" else 1, bs, 128, 2, 1, 2,
                         0 if method == "streamingLLM" else 1, topk)
    d = lambda a: torch.from_numpy(a).to(DEV)
    out_t = torch.full((B, M), -7, dtype=torch.int32, device=DEV)
    out_kv = torch.zeros(B, dtype=torch.int32, device=DEV)
    out_v = torch.zeros(B, dtype=torch.int32, device=DEV)
    sc = d(pool.copy())
    ops.evict_select_compact(d(tables), d(n_blocks), d(seq_lens), d(kv_lens), cpol, Hkv, out_t, out_kv,
                             out_v, ops.make_score_pool(sc) if method == "value_l2" else None,
                             kv_lens_pre_adjusted=False)
    torch.cuda.synchronize()
    assert np.array_equal(out_v.cpu().numpy(), ref_victim)
    assert np.array_equal(out_t.cpu().numpy(), ref_tables)
    assert np.array_equal(out_kv.cpu().numpy(), ref_kv)
    assert np.array_equal(sc.cpu().numpy(), pool)      # scores stay with their blocks: nothing to compact
    assert len(victims) > 3


@pytest.mark.parametrize("method", ["value_l2", "streamingLLM"])
def test_evict_select_compact_layers_matches_per_layer(method):
    """One launch for every layer of a step (grid batch x layers) == one launch per layer."""
    ops = _ops()
    from vllm_paged_compression_b200._lib import PcEvictPolicy
    bs, B, M, L = 16, 29, 50, 5
    pol = Policy(block_size=bs, cache_prune_type="percentage", evict_freq=2, decode_evict_method=method,
                 topk_blocks=-1)
    rs = np.random.RandomState(77)
    tables, n_blocks, seq_lens, scores, Hkv, pool = _evict_case(B, M, pol, rs, False)
    cpol = PcEvictPolicy(1, bs, 128, 2, 1, 2, 0 if method == "streamingLLM" else 1, -1)
    d = lambda a: torch.from_numpy(a).to(DEV)
    dt, dn, ds = d(tables), d(n_blocks), d(seq_lens)
    Ls = L if method == "value_l2" else 1
    pools = [d((pool * (1.0 + 0.5 * np.sin(np.arange(pool.size).reshape(pool.shape) * (l + 1)))).astype(np.float32))
             for l in range(Ls)]
    ref_t, ref_v = [], []
    for l in range(Ls):
        out_t = torch.full((B, M), -7, dtype=torch.int32, device=DEV)
        out_v = torch.zeros(B, dtype=torch.int32, device=DEV)
        ops.evict_select_compact(dt, dn, ds, None, cpol, Hkv, out_t, None, out_v,
                                 ops.make_score_pool(pools[l]) if method == "value_l2" else None)
        ref_t.append(out_t)
        ref_v.append(out_v)
    got_t = torch.full((Ls, B, M), -9, dtype=torch.int32, device=DEV)
    got_v = torch.zeros(Ls, B, dtype=torch.int32, device=DEV)
    ptrs = torch.tensor([p_.data_ptr() for p_ in pools], dtype=torch.int64, device=DEV) if method == "value_l2" else None
    ops.evict_select_compact_layers(dt, dn, ds, cpol, ptrs, Hkv if method == "value_l2" else 0, got_t, got_v)
    torch.cuda.synchronize()
    assert torch.equal(got_t, torch.stack(ref_t))
    assert torch.equal(got_v, torch.stack(ref_v))
    assert int((got_v >= 0).sum()) > 3
    if method == "value_l2":
        assert not torch.equal(got_v[0], got_v[1])      # the layers' pools differ, so do their victims


def test_compaction_known_answer_from_survey():
    """SURVEY.md 8c KAT: [[10..14],[20..24],[30..34]], victim 1 in every row ->
    [[10,12,13,14],[20,22,23,24],[30,32,33,34]] (v1 semantics; HEAD rows>=1 are buggy)."""
    ops = _ops()
    from vllm_paged_compression_b200._lib import PcEvictPolicy
    t = torch.tensor([[10, 11, 12, 13, 14], [20, 21, 22, 23, 24], [30, 31, 32, 33, 34]],
                     dtype=torch.int32, device=DEV)
    n = torch.full((3,), 5, dtype=torch.int32, device=DEV)
    seq = torch.full((3,), 160, dtype=torch.int32, device=DEV)
    out_t = torch.zeros(3, 5, dtype=torch.int32, device=DEV)
    out_v = torch.zeros(3, dtype=torch.int32, device=DEV)
    ops.evict_select_compact(t, n, seq, None, PcEvictPolicy(0, 16, 128, 2, 1, 2, 0, -1), 1, out_t, None, out_v)
    assert out_t.cpu().tolist() == [[10, 12, 13, 14, 0], [20, 22, 23, 24, 0], [30, 32, 33, 34, 0]]
    assert out_v.cpu().tolist() == [1, 1, 1]


# ------------------------------------------------------------------------- A2 + S2
def _decode_case(B, Hq, Hkv, D, bs, max_len, seed, ragged=True, nb=None):
    rs = np.random.RandomState(seed)
    kv_lens = (rs.randint(1, max_len + 1, size=B) if ragged else np.full(B, max_len)).astype(np.int32)
    kv_lens[0] = max_len
    M = policy.cdiv(max_len, bs)
    need = int(sum(policy.cdiv(int(l), bs) for l in kv_lens))
    nb = nb or need + 8
    perm = rs.permutation(nb)
    tables = np.zeros((B, M), dtype=np.int32)
    o = 0
    for r in range(B):
        n = policy.cdiv(int(kv_lens[r]), bs)
        tables[r, :n] = perm[o:o + n]
        o += n
    scale = D ** -0.5
    kv = (rnd(seed + 1, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16)
    q = (rnd(seed + 2, B, Hq, D).float() * 0.5).to(torch.bfloat16)
    return q, kv, tables, kv_lens, scale


@pytest.mark.parametrize("Hq,Hkv,D,bs", [(32, 8, 128, 16), (32, 8, 64, 16), (8, 2, 128, 32), (16, 2, 64, 32),
                                          (8, 8, 128, 16), (32, 2, 128, 16), (40, 8, 128, 16),
                                          (32, 8, 128, 8), (8, 2, 64, 8)])       # block 8: scripts/diff_page_sizes.sh:16
@pytest.mark.parametrize("B,max_len", [(1, 40), (3, 700), (16, 2100)])
def test_paged_decode_attention(Hq, Hkv, D, bs, B, max_len):
    ops = _ops()
    q, kv, tables, kv_lens, scale = _decode_case(B, Hq, Hkv, D, bs, max_len, seed=B * 31 + D + bs + Hq)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    ref = paged_ops.paged_attention(q, kc, vc, tables.tolist(), kv_lens.tolist(), scale).float()
    kvd = kv.to(DEV)
    kcd, vcd = paged_ops.split_kv_cache(kvd, Hkv, D)
    out = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.paged_decode_attn_score(out, q.to(DEV), kcd, vcd, torch.from_numpy(tables).to(DEV),
                                torch.from_numpy(kv_lens).to(DEV), scale, max_len)
    torch.cuda.synchronize()
    got = out.float().cpu()
    assert torch.isfinite(got).all()
    # reference kernel test bar is atol 1e-3 on values ~U(-s, s); here outputs are bf16-rounded
    # averages of N(0, 0.5) values: one bf16 ulp at |x| < 1 is 2^-8 = 3.9e-3
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=8e-3, rtol=0)


def test_paged_decode_golden_from_reference_cpu_kernel():
    """pa_out in ref_ops.npz came from the reference's csrc/cpu/attention.cpp paged_attention_v1."""
    ops = _ops()
    g = load_npz("ref_ops.npz")
    nb, bs, Hkv, Hq, D = 48, 16, 2, 8, 64
    q = (rnd(4, 4, Hq, D).float() * 0.3).to(torch.bfloat16)
    kv2 = (rnd(5, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16).to(DEV)
    kc2, vc2 = paged_ops.split_kv_cache(kv2, Hkv, D)
    out = torch.empty(4, Hq, D, dtype=torch.bfloat16, device=DEV)
    bt = torch.from_numpy(g["pa_block_tables"].astype(np.int32)).to(DEV)
    kl = torch.from_numpy(g["pa_kv_lens"].astype(np.int32)).to(DEV)
    ops.paged_attention_v1(out, q.to(DEV), kc2, vc2, Hkv, D ** -0.5, bt, kl, bs, int(g["pa_kv_lens"].max()))
    np.testing.assert_allclose(out.float().cpu().numpy(), g["pa_out"], atol=4e-3, rtol=0)
    out2 = torch.empty_like(out)
    ops.paged_attention_v2(out2, None, None, None, q.to(DEV), kc2, vc2, Hkv, D ** -0.5, bt, kl, bs,
                           int(g["pa_kv_lens"].max()))
    assert torch.equal(out, out2)


def test_paged_decode_empty_sequence_and_determinism():
    ops = _ops()
    B, Hq, Hkv, D, bs = 4, 32, 8, 128, 16
    q, kv, tables, kv_lens, scale = _decode_case(B, Hq, Hkv, D, bs, 4096, seed=77)
    kv_lens[2] = 0
    kvd = kv.to(DEV)
    kcd, vcd = paged_ops.split_kv_cache(kvd, Hkv, D)
    outs = []
    for _ in range(3):
        out = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
        ops.paged_decode_attn_score(out, q.to(DEV), kcd, vcd, torch.from_numpy(tables).to(DEV),
                                    torch.from_numpy(kv_lens).to(DEV), scale, 4096)
        outs.append(out.cpu())
    assert (outs[0][2] == 0).all()
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])   # split-KV combine is ordered


@pytest.mark.parametrize("B,max_len,zeros", [(130, 300, (5, 31, 32, 129)), (2, 20000, ()), (257, 50, (0, 256)),
                                             (7, 33, (3,))])
def test_paged_decode_streamk_schedule(B, max_len, zeros):
    """Stream-K ranges crossing many row / head boundaries, rows of zero length inside the
    batch, more rows than one scan chunk, and a group split over > 100 warps; run twice on
    the same workspace (arrival tickets must come back to zero)."""
    ops = _ops()
    Hq, Hkv, D, bs = 32, 8, 128, 16
    q, kv, tables, kv_lens, scale = _decode_case(B, Hq, Hkv, D, bs, max_len, seed=B + max_len)
    for z in zeros:
        kv_lens[z] = 0
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    ref = paged_ops.paged_attention(q, kc, vc, tables.tolist(), kv_lens.tolist(), scale).float()
    kvd = kv.to(DEV)
    kcd, vcd = paged_ops.split_kv_cache(kvd, Hkv, D)
    outs = []
    for _ in range(2):
        out = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
        ops.paged_decode_attn_score(out, q.to(DEV), kcd, vcd, torch.from_numpy(tables).to(DEV),
                                    torch.from_numpy(kv_lens).to(DEV), scale, max_len)
        torch.cuda.synchronize()
        outs.append(out.float().cpu())
    assert torch.isfinite(outs[0]).all()
    np.testing.assert_allclose(outs[0].numpy(), ref.numpy(), atol=8e-3, rtol=0)
    assert torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("Hkv,D,bs", [(8, 128, 16), (8, 64, 16), (2, 128, 32), (4, 64, 32), (8, 128, 8), (2, 64, 8)])
def test_fused_last_block_score_bit_exact(Hkv, D, bs):
    """Rows with seq_len % bs == 0 get pool[last block id][head_offset + h] = per-head value_l2 of
    their last (just filled) block, bit-identical to oracle/scoring.py; written to every pool
    listed in the pc_score_pool (the TP peer stores, emulated here with two local pools)."""
    ops = _ops()
    B, Hq, M = 9, Hkv * 4, 40
    rs = np.random.RandomState(D + bs)
    nblk = rs.randint(1, M + 1, size=B)
    nblk[0] = M
    full = rs.rand(B) < 0.6
    full[0] = True
    kv_lens = np.array([n * bs if f else n * bs - int(rs.randint(1, bs)) for n, f in zip(nblk, full)],
                       dtype=np.int32)
    seq_lens = (kv_lens + rs.randint(0, 5, size=B) * bs).astype(np.int32)     # logical >= physical
    nb = int(nblk.sum()) + 4
    perm = rs.permutation(nb)
    tables = np.zeros((B, M), dtype=np.int32)
    o = 0
    for r in range(B):
        tables[r, :nblk[r]] = perm[o:o + nblk[r]]
        o += nblk[r]
    kv = rnd(9, 2, nb, bs * Hkv * D)
    q = rnd(10, B, Hq, D)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    HT, hoff = 2 * Hkv, Hkv          # this rank computes the upper half of the heads
    ref = np.full((nb, HT), -1.0, dtype=np.float32)
    for r in range(B):
        if seq_lens[r] % bs == 0 and kv_lens[r] % bs == 0:
            blk = int(tables[r, nblk[r] - 1])
            vb = paged_ops.depage_v_block(vc, blk).float().numpy()
            ref[blk, hoff:hoff + Hkv] = scoring.per_head_block_score(vb)
    kvd = kv.to(DEV)
    kcd, vcd = paged_ops.split_kv_cache(kvd, Hkv, D)
    score = torch.full((nb, HT), -1.0, dtype=torch.float32, device=DEV)
    peer = torch.full((nb, HT), -1.0, dtype=torch.float32, device=DEV)
    sp = ops.make_score_pool(score, head_offset=hoff, peer_ptrs=[peer.data_ptr(), score.data_ptr()], local=1)
    out = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=DEV)
    ops.paged_decode_attn_score(out, q.to(DEV), kcd, vcd, torch.from_numpy(tables).to(DEV),
                                torch.from_numpy(kv_lens).to(DEV), D ** -0.5, M * bs,
                                torch.from_numpy(seq_lens).to(DEV), sp)
    torch.cuda.synchronize()
    assert np.array_equal(score.cpu().numpy().view(np.int32), ref.view(np.int32))
    assert torch.equal(score, peer)


# ------------------------------------------------------------------------------ A1
@pytest.mark.parametrize("Hq,Hkv,D", [(8, 2, 64), (8, 2, 128), (4, 4, 128)])
@pytest.mark.parametrize("mode", ["budget", "percentage", "none"])
def test_prefill_attention_topleft_causal(Hq, Hkv, D, mode):
    ops = _ops()
    bs = 16
    pol = Policy(block_size=bs, cache_prune_type="budget" if mode == "none" else mode,
                 cache_budget=(1 << 20) if mode == "none" else 128, evict_freq=2)
    prompt_lens = [1, 17, 128, 200, 333]
    T = sum(prompt_lens)
    q = (rnd(31, T, Hq, D).float() * 0.5).to(torch.bfloat16)
    k = (rnd(32, T, Hkv, D).float() * 0.5).to(torch.bfloat16)
    v = rnd(33, T, Hkv, D)
    meta, gidx, kept_l, s = [], [], [], 0
    for L in prompt_lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        meta.append([s, L, s, kept, first, pruned])
        gidx.append(torch.from_numpy(policy.prompt_gather_index(L, pol)) + s)
        kept_l.append(kept)
        s += L
    gidx = torch.cat(gidx)
    ref = paged_ops.prefill_attention(q, k[gidx], v[gidx], prompt_lens, kept_l, D ** -0.5).float()
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.prefill_attn_topleft_causal(out, q.to(DEV), k.to(DEV), v.to(DEV),
                                    torch.tensor(meta, dtype=torch.int32, device=DEV), max(prompt_lens),
                                    D ** -0.5)
    torch.cuda.synchronize()
    got = out.float().cpu()
    assert torch.isfinite(got).all()
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=2e-2, rtol=0)


@pytest.mark.parametrize("case", [
    # (Hq, Hkv, D, block, mode, evict_freq/budget, prompt_lens): tile edges of the tcgen05 kernel
    # (128-query tiles, 64-wide kv tiles), multi-tile rows, block 32, both prune modes
    (8, 2, 128, 16, "percentage", 2, [127, 128, 129, 255, 257, 700]),
    (8, 2, 128, 16, "percentage", 4, [1024, 63, 65, 1500]),
    (4, 1, 128, 32, "budget", 256, [96, 640, 1111]),
    (8, 8, 64, 16, "budget", 128, [513, 64, 390]),
    (16, 2, 128, 16, "none", 0, [300, 1, 900]),
    # kernel dispatch of the product library (prefill_attn_tc.cu: prefill_attn): head size 128 with an even
    # number of q heads per kv head and prompts <= 6144 tokens runs the dual-query-tile kernel
    # (prefill_attn_dq.cu, the cases above); these two reach the one-tile kernel at head size 128 --
    # a prompt beyond 6144 tokens (one CTA per query tile) and three q heads per kv head (persistent CTAs)
    (4, 2, 128, 16, "percentage", 2, [6272, 129, 1]),
    (6, 2, 128, 16, "percentage", 2, [127, 640, 129, 1100]),
    # and the dual-query-tile kernel at its longest prompt, ragged next to a one-token prompt
    (4, 2, 128, 16, "percentage", 4, [6144, 1, 1999]),
    (8, 4, 128, 16, "none", 0, [130, 1, 64, 1000, 2, 257]),     # short items next to empty ones (query tiles past a prompt's end)
])
def test_prefill_attention_tile_edges_and_strided_rows(case):
    """q/k/v are column slices of one fused [T, (Hq + 2 Hkv) * D] buffer (row stride != H * D),
    which is how the qkv projection hands them over (and exercises the tensor-map strides)."""
    Hq, Hkv, D, bs, mode, knob, prompt_lens = case
    ops = _ops()
    pol = Policy(block_size=bs, cache_prune_type="budget" if mode == "none" else mode,
                 cache_budget=(1 << 20) if mode == "none" else (knob if mode == "budget" else 128),
                 evict_freq=knob if mode == "percentage" else 2)
    T = sum(prompt_lens)
    qkv = (rnd(51, T, (Hq + 2 * Hkv) * D).float() * 0.7).to(torch.bfloat16)
    q = qkv[:, :Hq * D].view(T, Hq, D)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    meta, gidx, kept_l, s = [], [], [], 0
    for L in prompt_lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        meta.append([s, L, s, kept, first, pruned])
        gidx.append(torch.from_numpy(policy.prompt_gather_index(L, pol)) + s)
        kept_l.append(kept)
        s += L
    gidx = torch.cat(gidx)
    ref = paged_ops.prefill_attention(q.contiguous(), k[gidx].contiguous(), v[gidx].contiguous(),
                                      prompt_lens, kept_l, D ** -0.5).float()
    dqkv = qkv.to(DEV)
    dq = dqkv[:, :Hq * D].view(T, Hq, D)
    dk = dqkv[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D)
    dv = dqkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.prefill_attn_topleft_causal(out, dq, dk, dv, torch.tensor(meta, dtype=torch.int32, device=DEV),
                                    max(prompt_lens), D ** -0.5)
    torch.cuda.synchronize()
    got = out.float().cpu()
    assert torch.isfinite(got).all()
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=2e-2, rtol=0)


@pytest.mark.timeout(120)
@pytest.mark.parametrize("Hq,Hkv", [(16, 4), (12, 4)])    # dual-query-tile kernel / one-tile kernel (odd group)
def test_prefill_attention_many_short_items_persistent_walk(Hq, Hkv):
    """More (tile, head, request) items than resident CTAs (2 x 148), so every CTA of the
    persistent launch walks several items back to back: phases of all mbarriers keep counting
    across items, Q is re-loaded behind the previous item's last S MMA, O / P buffers are reused."""
    ops = _ops()
    D, bs = 128, 16
    pol = Policy(block_size=bs, cache_prune_type="percentage", cache_budget=128, evict_freq=2)
    g = torch.Generator().manual_seed(77)
    prompt_lens = [int(x) for x in torch.randint(1, 420, (40,), generator=g)] + [512, 384, 129]
    T = sum(prompt_lens)
    q = (rnd(71, T, Hq, D).float() * 0.6).to(torch.bfloat16)
    k = (rnd(72, T, Hkv, D).float() * 0.6).to(torch.bfloat16)
    v = rnd(73, T, Hkv, D)
    meta, gidx, kept_l, s = [], [], [], 0
    for L in prompt_lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        meta.append([s, L, s, kept, first, pruned])
        gidx.append(torch.from_numpy(policy.prompt_gather_index(L, pol)) + s)
        kept_l.append(kept)
        s += L
    gidx = torch.cat(gidx)
    ref = paged_ops.prefill_attention(q, k[gidx], v[gidx], prompt_lens, kept_l, D ** -0.5).float()
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    dq, dk, dv = q.to(DEV), k.to(DEV), v.to(DEV)
    dm = torch.tensor(meta, dtype=torch.int32, device=DEV)
    for _ in range(3):                       # repeated launches: no state may leak between them
        ops.prefill_attn_topleft_causal(out, dq, dk, dv, dm, max(prompt_lens), D ** -0.5)
    torch.cuda.synchronize()
    got = out.float().cpu()
    assert torch.isfinite(got).all()
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=2e-2, rtol=0)


def test_prefill_attention_large_logits_rescale_path():
    """Scores that grow along the row force the lazy O rescale in TMEM (max jumps > 2^8)."""
    ops = _ops()
    Hq, Hkv, D, L = 2, 1, 128, 640
    q = torch.ones(L, Hq, D, dtype=torch.bfloat16) * 0.5
    ramp = torch.linspace(0.0, 6.0, L).view(L, 1, 1)
    k = (ramp * torch.ones(L, Hkv, D)).to(torch.bfloat16)          # logit_j ~ 0.5 * 6 j / L * D * scale
    v = rnd(61, L, Hkv, D)
    meta = torch.tensor([[0, L, 0, L, L, 0]], dtype=torch.int32, device=DEV)
    ref = paged_ops.prefill_attention(q, k, v, [L], [L], D ** -0.5).float()
    out = torch.full((L, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.prefill_attn_topleft_causal(out, q.to(DEV), k.to(DEV), v.to(DEV), meta, L, D ** -0.5)
    torch.cuda.synchronize()
    np.testing.assert_allclose(out.float().cpu().numpy(), ref.numpy(), atol=2e-2, rtol=0)


# ------------------------------------------------------------------ block maintenance
def test_copy_and_swap_blocks():
    ops = _ops()
    Hkv, D, bs, nb = 2, 64, 16, 32
    kv = rnd(41, 2, nb, bs * Hkv * D).to(DEV)
    ref = kv.clone()
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    pairs = torch.tensor([[0, 5], [0, 6], [9, 1]], dtype=torch.int64)
    ops.copy_blocks(kc, vc, pairs.to(DEV))
    for s, d in pairs.tolist():
        ref[:, d] = ref[:, s]
    assert torch.equal(kv, ref)
    host = torch.zeros(nb, bs * Hkv * D, dtype=torch.bfloat16).pin_memory()
    ops.swap_blocks(kv[0], host, torch.tensor([[3, 7], [4, 0]]))
    torch.cuda.synchronize()
    assert torch.equal(host[7], kv[0, 3].cpu()) and torch.equal(host[0], kv[0, 4].cpu())
    # runs of consecutive (src, dst) pairs travel as one copy each: runs, a broken run, a lone pair
    host.zero_()
    m = torch.tensor([[10, 20], [11, 21], [12, 22], [13, 24], [14, 25], [2, 3], [30, 29], [31, 30]])
    ops.swap_blocks(kv[0], host, m)
    torch.cuda.synchronize()
    want = torch.zeros_like(host)
    for s_, d_ in m.tolist():
        want[d_] = kv[0, s_].cpu()
    assert torch.equal(host, want)
    back = torch.zeros_like(kv[0])
    ops.swap_blocks(host, back, torch.tensor([[20, 0], [21, 1], [22, 2], [24, 9]]))      # host -> device
    torch.cuda.synchronize()
    assert torch.equal(back[:3].cpu(), want[20:23]) and torch.equal(back[9].cpu(), want[24])


# --------------------------------------------------------------------------- W1 fused into A2
@pytest.mark.parametrize("Hq,Hkv,D,bs", [(32, 8, 128, 16), (8, 2, 64, 16), (16, 2, 128, 32), (32, 2, 64, 32),
                                          (32, 8, 128, 8), (8, 2, 64, 8)])
@pytest.mark.parametrize("B,max_len", [(5, 90), (33, 1300)])
def test_decode_layer_fused_cache_write(Hq, Hkv, D, bs, B, max_len):
    """pc_decode_layer with fuse_cache_write: the attention kernel itself stores the step's
    K/V rows.  Cache bytes must equal the oracle's reshape_and_cache (and the un-fused
    path's) bit for bit, outputs must equal the un-fused path's bit for bit (same arithmetic
    on the same bytes), scores of blocks that fill this step likewise.  kv lengths are
    chosen so that new tokens land at every offset, including 0 (fresh block) and bs-1."""
    ops = _ops()
    seed = B * 7 + D + bs + Hq
    q, kv, tables, kv_lens, scale = _decode_case(B, Hq, Hkv, D, bs, max_len, seed=seed)
    for r in range(min(B, bs)):                 # force every in-block offset to occur
        kv_lens[r] = max(1, (int(kv_lens[r]) // bs) * bs - r) if kv_lens[r] >= bs else r + 1
    n = [policy.cdiv(int(x), bs) for x in kv_lens]
    slots = np.array([int(tables[r, n[r] - 1]) * bs + (int(kv_lens[r]) - 1) % bs for r in range(B)], np.int64)
    key = (rnd(seed + 5, B, Hkv, D).float() * 0.5).to(torch.bfloat16)
    val = rnd(seed + 6, B, Hkv, D)
    # oracle: write, then attend
    kv_ref = kv.clone()
    kc, vc = paged_ops.split_kv_cache(kv_ref, Hkv, D)
    paged_ops.reshape_and_cache(key, val, kc, vc, torch.from_numpy(slots))
    ref = paged_ops.paged_attention(q, kc, vc, tables.tolist(), kv_lens.tolist(), scale).float()
    seq_lens = kv_lens.copy()                   # rows with seq_len % bs == 0 are scored
    results = []
    for fused in (False, True):
        kvd = kv.clone().to(DEV)
        kcd, vcd = paged_ops.split_kv_cache(kvd, Hkv, D)
        nb = kvd.shape[1]
        pool = torch.zeros(nb, Hkv, dtype=torch.float32, device=DEV)
        sp = ops.make_score_pool(pool, 0)
        out = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
        qd, kd, vd = q.to(DEV), key.to(DEV), val.to(DEV)
        st = ops.make_decode_step(torch.from_numpy(slots).to(DEV), torch.from_numpy(tables).to(DEV),
                                  torch.from_numpy(kv_lens).to(DEV), torch.from_numpy(seq_lens).to(DEV),
                                  B, Hq, Hkv, D, bs, int(kv_lens.max()), scale, qd.stride(0), kd.stride(0),
                                  vd.stride(0), out.stride(0), vcd.stride(0), fuse_cache_write=fused)
        assert st.fuse_cache_write == int(fused)
        ops.decode_layer(st, out, qd, kd, vd, kcd.data_ptr(), vcd.data_ptr(), sp, None)
        torch.cuda.synchronize()
        results.append((out.cpu(), kvd.cpu(), pool.cpu()))
    (o0, c0, s0), (o1, c1, s1) = results
    assert np.array_equal(_bits(c0), _bits(kv_ref)), "un-fused cache bytes differ from the oracle"
    assert np.array_equal(_bits(c1), _bits(kv_ref)), "fused cache bytes differ from the oracle"
    assert torch.equal(o0, o1)
    assert torch.equal(s0, s1)
    assert any(int(x) % bs == 0 for x in kv_lens) and s1.abs().sum() > 0
    np.testing.assert_allclose(o1.float().numpy(), ref.numpy(), atol=8e-3, rtol=0)


@pytest.mark.parametrize("repack", [True, False])
@pytest.mark.parametrize("kind,D", [("value_l2", 64), ("inverse_key_l2", 128)])
def test_prompt_token_prune_bit_exact(kind, D, repack):
    """f3: pc_prompt_token_prune against oracle/token_prune.py (itself pinned to the reference's
    prune_prompt_inplace): a batch of prompts below / at / above the budget, with duplicated
    tokens so that equal scores straddle the cut (ties -> lowest token index goes first).  Pure
    gather: bit-exact."""
    from oracle import token_prune
    from tests.util import token_prune_data
    bs, budget, H = 16, 128, 4
    lens = [300, 100, 131, 128, 1030, 161]
    ks, vs = [], []
    for i, L in enumerate(lens):
        k, v = token_prune_data(500 + i, L, H, D)
        if L > 200:      # ties: copy a few middle tokens over others (all heads / one head)
            k[40:48], v[40:48] = k[60:68].clone(), v[60:68].clone()
            k[90, 1], v[90, 1] = k[33, 1].clone(), v[33, 1].clone()
        ks.append(k.to(torch.bfloat16))
        vs.append(v.to(torch.bfloat16))
    key, value = torch.cat(ks).to(DEV), torch.cat(vs).to(DEV)
    tp6, moff, qoff = [], [0], [0]
    src = dst = 0
    want_k, want_v = [], []
    for L, k, v in zip(lens, ks, vs):
        first, mid, kd = token_prune.prune_layout(L, budget, bs)
        tp6.append((src, L, first, mid, kd, dst))
        moff.append(moff[-1] + mid)
        qoff.append(qoff[-1] + L)
        pk, pv = token_prune.prune_prompt(k.float().numpy(), v.float().numpy(), budget, bs, kind, repack=repack)
        assert pk.shape[0] == L - kd
        want_k.append(torch.from_numpy(pk).to(torch.bfloat16))
        want_v.append(torch.from_numpy(pv).to(torch.bfloat16))
        src += L
        dst += L - kd
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    out_k = torch.full_like(key, 7.0)
    out_v = torch.full_like(value, 7.0)
    _ops().prompt_token_prune(key, value, out_k, out_v, i32(tp6), i32(moff), i32(qoff), qoff[-1], moff[-1], kind,
                           head_mixing_repack=repack)
    torch.cuda.synchronize()
    assert torch.equal(out_k[:dst].cpu(), torch.cat(want_k)) and torch.equal(out_v[:dst].cpu(), torch.cat(want_v))
    assert bool((out_k[dst:] == 7.0).all()) and bool((out_v[dst:] == 7.0).all())     # nothing written past the end


def test_torch_library_ops_match_ctypes_path():
    """torch.ops.pagedevict_b200.* (and the fork's names after alias_vllm_ops) give the same bytes
    as the ctypes wrappers: reshape_and_cache, paged_attention_v1/v2, zero_cache_slots."""
    from vllm_paged_compression_b200 import torch_ops
    ns = torch_ops.load()
    torch_ops.alias_vllm_ops()
    ops = _ops()
    nb, bs, Hkv, Hq, D, B = 40, 16, 2, 8, 128, 5
    s = D ** -0.5
    kv = (rnd(71, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16).to(DEV)
    kv2 = kv.clone()
    key, value = rnd(72, 30, Hkv, D).to(DEV), rnd(73, 30, Hkv, D).to(DEV)
    slots = torch.from_numpy(np.random.RandomState(3).permutation(nb * bs)[:30].astype(np.int64)).to(DEV)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    kc2, vc2 = paged_ops.split_kv_cache(kv2, Hkv, D)
    ops.reshape_and_cache(key, value, kc, vc, slots)
    torch.ops._C_cache_ops.reshape_and_cache(key, value, kc2, vc2, slots, "auto", 1.0, 1.0)
    assert torch.equal(kv, kv2)
    q = (rnd(74, B, Hq, D).float() * 0.3).to(torch.bfloat16).to(DEV)
    lens = torch.tensor([1, 16, 37, 160, 99], dtype=torch.int32, device=DEV)
    bt = torch.from_numpy(np.stack([np.random.RandomState(10 + i).permutation(nb)[:10] for i in range(B)])
                          .astype(np.int32)).to(DEV)
    o1, o2, o3 = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
    ops.paged_attention_v1(o1, q, kc, vc, Hkv, s, bt, lens, bs, 160)
    ns.paged_attention_v1(o2, q, kc, vc, Hkv, s, bt, lens, bs, 160, None, "auto", 1.0, 1.0, 0, 0, 0, 64, 0)
    dummy = torch.empty(1, device=DEV)
    torch.ops._C.paged_attention_v2(o3, dummy, dummy, dummy, q, kc, vc, Hkv, s, bt, lens, bs, 160, None, "auto",
                                    1.0, 1.0, 0, 0, 0, 64, 0)
    assert torch.equal(o1, o2) and torch.equal(o1, o3)
    z = torch.tensor([5, -1, 77], dtype=torch.int64, device=DEV)
    ops.zero_cache_slots(kc, vc, z)
    ns.zero_cache_slots(kc2, vc2, z)
    assert torch.equal(kv, kv2)
    with pytest.raises(RuntimeError):
        ns.paged_attention_v1(o2, q, kc, vc, Hkv, s, bt, lens, bs, 160, None, "fp8", 1.0, 1.0, 0, 0, 0, 64, 0)


@pytest.mark.parametrize("D", [64, 128])
def test_prefill_attention_block_size_8_gather(D):
    """Block size 8 (scripts/diff_page_sizes.sh:16): the sink block is 8 tokens and the pruned run a
    multiple of 8, so the streamingLLM gather runs on 8-row TMA boxes (picked from the host copy of
    req_meta); without the host copy such a request is refused by the launch-time check only when
    misaligned to 8."""
    ops = _ops()
    Hq, Hkv, bs = 8, 2, 8
    pol = Policy(block_size=bs, cache_prune_type="budget", cache_budget=128)
    prompt_lens = [5, 128, 137, 300, 1001]
    T = sum(prompt_lens)
    q = (rnd(41, T, Hq, D).float() * 0.5).to(torch.bfloat16)
    k = (rnd(42, T, Hkv, D).float() * 0.5).to(torch.bfloat16)
    v = rnd(43, T, Hkv, D)
    meta, gidx, kept_l, s = [], [], [], 0
    for L in prompt_lens:
        kept, pruned, first = policy.prompt_prune_layout(L, pol)
        meta.append([s, L, s, kept, first, pruned])
        gidx.append(torch.from_numpy(policy.prompt_gather_index(L, pol)) + s)
        kept_l.append(kept)
        s += L
    assert any((m[4] | m[5]) & 15 for m in meta if m[5]), "the case must need 8-row boxes"
    gidx = torch.cat(gidx)
    ref = paged_ops.prefill_attention(q, k[gidx], v[gidx], prompt_lens, kept_l, D ** -0.5).float()
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    m = np.asarray(meta, np.int32)
    ops.prefill_attn_topleft_causal(out, q.to(DEV), k.to(DEV), v.to(DEV), torch.from_numpy(m).to(DEV),
                                    max(prompt_lens), D ** -0.5, req_meta_host=m)
    torch.cuda.synchronize()
    got = out.float().cpu()
    assert torch.isfinite(got).all()
    np.testing.assert_allclose(got.numpy(), ref.numpy(), atol=2e-2, rtol=0)
    bad = m.copy()
    bad[3, 5] += 4          # pruned run not a multiple of 8: refused before the launch
    with pytest.raises(RuntimeError):
        ops.prefill_attn_topleft_causal(out, q.to(DEV), k.to(DEV), v.to(DEV), torch.from_numpy(bad).to(DEV),
                                        max(prompt_lens), D ** -0.5, req_meta_host=bad)
