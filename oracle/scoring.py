"""value_l2 block scoring oracle (fp32, FIXED reduction order).  TEST INFRASTRUCTURE ONLY.

Reference: KVCachePruner.get_block_score, sub_evict_method "value_l2"
(vllm/attention/kvcache_prunner.py:51-56):
    torch.norm(V_blk[bs,Hkv,D], p=2, dim=-1).mean(dim=1).sum(dim=0)
  = (1/Hkv) * sum_t sum_h ||V[t,h,:]||_2
called from torch_cuda_sdpa.py:1278-1288 (prompt blocks) and 1518-1526 (the block
that just filled at decode).

The reference evaluates this in the cache dtype with torch's own reduction order and
on a mis-viewed V block.  North-star semantics: fp32, deterministic
order.  The order below is THE definition shared by the CUDA kernels
(csrc/score_common.cuh) and this oracle, so evicted block ids are bit-exact:

  lane_j  = sum_{i=0..D/8-1} v[8i+j]^2      sequential in i, fmaf(v, v, acc), acc0 = 0
  sumsq   = ((l0+l1)+(l2+l3)) + ((l4+l5)+(l6+l7))
  norm_t  = sqrt_rn(sumsq)
  head_h  = sum_{t=0..bs-1} norm_t          sequential in t, acc0 = 0
  block   = (sum_{h ascending} head_h) / Hkv        (fp32 add, fp32 divide, RN)

For bf16 / fp16 inputs v*v is exact in fp32, so fmaf(v,v,acc) == fl(v*v + acc) and
numpy float32 arithmetic reproduces the GPU bits.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def per_head_block_score(v_blk: np.ndarray) -> np.ndarray:
    """v_blk: [bs, Hkv, D] float32 holding exactly-representable bf16/fp16 values.
    -> [Hkv] float32 per-head partial  sum_t ||V[t,h,:]||."""
    v = np.ascontiguousarray(v_blk, dtype=F32)
    bs, H, D = v.shape
    assert D % 8 == 0
    x = v.reshape(bs, H, D // 8, 8)
    acc = np.zeros((bs, H, 8), dtype=F32)
    for i in range(D // 8):
        xi = x[:, :, i, :]
        acc = (xi * xi + acc).astype(F32)
    s01 = (acc[..., 0] + acc[..., 1]).astype(F32)
    s23 = (acc[..., 2] + acc[..., 3]).astype(F32)
    s45 = (acc[..., 4] + acc[..., 5]).astype(F32)
    s67 = (acc[..., 6] + acc[..., 7]).astype(F32)
    sumsq = ((s01 + s23).astype(F32) + (s45 + s67).astype(F32)).astype(F32)
    norm = np.sqrt(sumsq, dtype=F32)                     # [bs, H]
    head = np.zeros((H,), dtype=F32)
    for t in range(bs):
        head = (head + norm[t]).astype(F32)
    return head


def combine_heads(per_head: np.ndarray) -> np.ndarray:
    """per_head [..., Hkv] fp32 -> block score [...]: heads ascending, then / Hkv."""
    ph = np.asarray(per_head, dtype=F32)
    H = ph.shape[-1]
    acc = np.zeros(ph.shape[:-1], dtype=F32)
    for h in range(H):
        acc = (acc + ph[..., h]).astype(F32)
    return (acc / F32(H)).astype(F32)


def block_score(v_blk: np.ndarray) -> np.float32:
    return combine_heads(per_head_block_score(v_blk))


def reference_formula_fp32(v_blk: np.ndarray) -> float:
    """get_block_score evaluated in float64-accumulated numpy, for tolerance checks of
    the fixed-order definition against the reference formula."""
    v = np.asarray(v_blk, dtype=np.float64)
    return float(np.sqrt((v * v).sum(-1)).mean(axis=1).sum(axis=0))
