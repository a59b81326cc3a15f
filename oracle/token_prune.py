"""Token-level prompt prune oracle (HEAD's local / global / inverse_key_l2 prompt path).
TEST INFRASTRUCTURE ONLY.

Restates the second branch of KVCachePruner.prune_prompt_inplace
(vllm/attention/kvcache_prunner.py:220-279): keep the first block and the last slice, and of
the middle slice drop, per kv head, the k lowest-scoring tokens (`torch.topk(scores, k,
largest=False, dim=0)` on scores [T_mid, H]); the survivors are re-packed with
`middle[mask].view(-1, H, D)` -- boolean-mask order (token-major, head-minor) re-chunked H at a
time, which mixes source heads inside a column.

Pinned by tests/golden/ref_token_prune.npz: the reference's own function executed on CPU in
fp32 on inputs whose k-th / (k+1)-th scores are well separated (generator
tests/golden/make_golden_token_prune.py).  Declared deviation (like D1 for block scores):
scores here are fp32 in the fixed order of oracle/scoring.py, ties go to the lowest token index.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

F32 = np.float32


def token_norms(x: np.ndarray) -> np.ndarray:
    """x [T, H, D] float32 (bf16-representable) -> [T, H] fp32 L2 norms, reduction order of
    csrc/score_common.cuh: 8 strided partial sums (fma), butterfly, sqrt."""
    x = np.ascontiguousarray(x, dtype=F32)
    T, H, D = x.shape
    xr = x.reshape(T, H, D // 8, 8)
    acc = np.zeros((T, H, 8), dtype=F32)
    for i in range(D // 8):
        xi = xr[:, :, i, :]
        acc = (xi * xi + acc).astype(F32)
    s01 = (acc[..., 0] + acc[..., 1]).astype(F32)
    s23 = (acc[..., 2] + acc[..., 3]).astype(F32)
    s45 = (acc[..., 4] + acc[..., 5]).astype(F32)
    s67 = (acc[..., 6] + acc[..., 7]).astype(F32)
    sumsq = ((s01 + s23).astype(F32) + (s45 + s67).astype(F32)).astype(F32)
    return np.sqrt(sumsq, dtype=F32)


def token_scores(k_mid: np.ndarray, v_mid: np.ndarray, kind: str) -> np.ndarray:
    """KVCachePruner.get_token_score (kvcache_prunner.py:28-49): value_l2 for local / global,
    1 / (||K|| + 1e-8) for inverse_key_l2."""
    if kind == "value_l2":
        return token_norms(v_mid)
    if kind == "inverse_key_l2":
        return (F32(1.0) / (token_norms(k_mid) + F32(1e-8)).astype(F32)).astype(F32)
    raise ValueError(kind)


def prune_layout(q_len: int, budget: int, bs: int) -> Tuple[int, int, int]:
    """-> (first, mid, k_drop) (kvcache_prunner.py:221-236)."""
    if q_len <= budget:
        return q_len, 0, 0
    rem = q_len % bs
    first = bs
    end_mid = q_len - bs - rem
    mid = end_mid - first
    k = max(mid - (budget - 2 * bs), 0)
    return first, mid, k


def keep_mask(scores: np.ndarray, k: int) -> np.ndarray:
    """[T, H] bool: per head the k smallest (score, token index) pairs are dropped."""
    T, H = scores.shape
    keep = np.ones((T, H), dtype=bool)
    if k <= 0:
        return keep
    for h in range(H):
        order = np.lexsort((np.arange(T), scores[:, h]))      # by score, then by index
        keep[order[:k], h] = False
    return keep


def prune_prompt(key: np.ndarray, value: np.ndarray, budget: int, bs: int, kind: str,
                 repack: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """key / value [q_len, H, D] float32 -> pruned (key, value) [q_len - k, H, D]."""
    q_len, H, D = key.shape
    first, mid, k = prune_layout(q_len, budget, bs)
    if k == 0:
        return key.copy(), value.copy()
    km, vm = key[first:first + mid], value[first:first + mid]
    keep = keep_mask(token_scores(km, vm, kind), k)
    if repack:      # middle[mask].view(-1, H, D)
        pk, pv = km[keep].reshape(-1, H, D), vm[keep].reshape(-1, H, D)
    else:
        pk = np.stack([km[keep[:, h], h] for h in range(H)], axis=1)
        pv = np.stack([vm[keep[:, h], h] for h in range(H)], axis=1)
    return (np.concatenate([key[:first], pk, key[first + mid:]]),
            np.concatenate([value[:first], pv, value[first + mid:]]))
