"""pc_paged_decode_attn_score_static: the decode attention entry point that allocates nothing.

Its split-KV workspace is a file-scope device array of the library (PC_DECODE_STATIC_SCRATCH_BYTES,
include/pagedevict_b200.h).  Called here straight through the C ABI (ctypes, raw pointers) and
compared with pc_paged_decode_attn_score on the same inputs, twice in a row (the array keeps the last
call's contents).  This file sorts last on purpose.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import paged_ops, policy
from tests.util import rnd

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATOL = 8e-3          # the bar of the decode kernel tests (tests/test_gpu_kernels.py)


def _case(B, Hq, Hkv, D, bs, max_len, seed):
    rs = np.random.RandomState(seed)
    kv_lens = rs.randint(1, max_len + 1, size=B).astype(np.int32)
    kv_lens[0] = max_len
    kv_lens[B - 1] = (max_len // bs) * bs            # a row whose last block is full (scored)
    M = policy.cdiv(max_len, bs)
    n = [policy.cdiv(int(x), bs) for x in kv_lens]
    nb = sum(n) + 8
    perm = rs.permutation(nb)
    tables = np.zeros((B, M), np.int32)
    o = 0
    for r in range(B):
        tables[r, :n[r]] = perm[o:o + n[r]]
        o += n[r]
    kv = (rnd(seed + 1, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16).to(DEV)
    q = (rnd(seed + 2, B, Hq, D).float() * 0.5).to(torch.bfloat16).to(DEV)
    return q, kv, torch.from_numpy(tables).to(DEV), torch.from_numpy(kv_lens).to(DEV), nb


def _call_static(lib, ops, out, q, kc, vc, tables, kv_lens, Hkv, D, bs, max_len, scale, seq_lens, score):
    B, Hq, _ = q.shape
    return lib.pc_paged_decode_attn_score_static(
        out.data_ptr(), q.data_ptr(), kc.data_ptr(), vc.data_ptr(), tables.data_ptr(), tables.stride(0),
        kv_lens.data_ptr(), B, Hq, Hkv, D, bs, int(max_len), float(scale), q.stride(0), out.stride(0),
        kc.stride(0), seq_lens.data_ptr() if score is not None else None,
        C.byref(score) if score is not None else None, ops.PC_BF16,
        torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("B,max_len,Hq,Hkv,D,bs", [(3, 2100, 32, 8, 128, 16),    # few rows: split KV (tickets + partials)
                                                    (64, 600, 32, 8, 128, 16),    # the bench shape
                                                    (9, 700, 8, 2, 64, 32)])
def test_static_scratch_entry_matches_workspace_entry(B, max_len, Hq, Hkv, D, bs):
    from vllm_paged_compression_b200 import _lib, ops
    lib = _lib.load()
    q, kv, tables, kv_lens, nb = _case(B, Hq, Hkv, D, bs, max_len, seed=B + max_len + D)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    scale = D ** -0.5
    seq_lens = kv_lens.clone()
    pool_a = torch.zeros(nb, Hkv, dtype=torch.float32, device=DEV)
    pool_b = torch.zeros_like(pool_a)
    want = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.paged_decode_attn_score(want, q, kc, vc, tables, kv_lens, scale, max_len, seq_lens=seq_lens,
                                score=ops.make_score_pool(pool_a, 0))
    torch.cuda.synchronize()
    assert torch.isfinite(want.float()).all()
    sp = ops.make_score_pool(pool_b, 0)
    for i in range(2):                                           # two calls in a row on the shared scratch
        got = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
        pool_b.zero_()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = _call_static(lib, ops, got, q, kc, vc, tables, kv_lens, Hkv, D, bs, max_len, scale, seq_lens, sp)
        assert rc == 0, lib.pc_last_error().decode()
        torch.cuda.synchronize()
        if i == 1:                                               # (the first call may load the kernel lazily)
            assert torch.cuda.mem_get_info()[0] == free0, "the call changed the device's free memory"
        np.testing.assert_allclose(got.float().cpu().numpy(), want.float().cpu().numpy(), atol=ATOL, rtol=0)
        assert torch.equal(got, want)                            # same kernel, same plan, ordered combine
        assert torch.equal(pool_b, pool_a)                       # fused last-block scores


def test_static_scratch_entry_rejects_oversized_launch():
    """Every (row, kv head group) takes a partial slot: 512 rows x 8 kv heads at head size 128 need
    ~17 MB + the ticket table > PC_DECODE_STATIC_SCRATCH_BYTES.  Rejected before any launch."""
    from vllm_paged_compression_b200 import _lib, ops
    lib = _lib.load()
    B, Hq, Hkv, D, bs, max_len = 512, 32, 8, 128, 16, 32
    q, kv, tables, kv_lens, nb = _case(B, Hq, Hkv, D, bs, max_len, seed=5)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    out = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=DEV)
    rc = _call_static(lib, ops, out, q, kc, vc, tables, kv_lens, Hkv, D, bs, max_len, D ** -0.5, None, None)
    assert rc == -1         # PC_ERR_INVALID
    msg = lib.pc_last_error().decode()
    assert "PC_DECODE_STATIC_SCRATCH_BYTES" in msg and "pc_paged_decode_attn_score" in msg, msg
    torch.cuda.synchronize()
    assert int(out.abs().sum()) == 0                             # nothing ran
    ops.paged_decode_attn_score(out, q, kc, vc, tables, kv_lens, D ** -0.5, max_len)   # the workspace entry takes it
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all() and float(out.float().abs().sum()) > 0


def test_static_scratch_entry_under_graph_capture():
    """The case it exists for: captured into a CUDA graph (no allocation may happen there), replayed."""
    from vllm_paged_compression_b200 import _lib, ops
    lib = _lib.load()
    B, Hq, Hkv, D, bs, max_len = 5, 32, 8, 128, 16, 1500
    q, kv, tables, kv_lens, nb = _case(B, Hq, Hkv, D, bs, max_len, seed=11)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    scale = D ** -0.5
    want = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=DEV)
    ops.paged_decode_attn_score(want, q, kc, vc, tables, kv_lens, scale, max_len)
    got = torch.zeros_like(want)
    rc = _call_static(lib, ops, got, q, kc, vc, tables, kv_lens, Hkv, D, bs, max_len, scale, None, None)   # warm
    assert rc == 0, lib.pc_last_error().decode()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        rc = _call_static(lib, ops, got, q, kc, vc, tables, kv_lens, Hkv, D, bs, max_len, scale, None, None)
    assert rc == 0, lib.pc_last_error().decode()
    for _ in range(2):
        got.fill_(float("nan"))
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(got, want)
