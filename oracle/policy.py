"""Integer-domain oracle: eviction policy arithmetic.  TEST INFRASTRUCTURE ONLY.

Plain Python ints / numpy int arrays; every function cites the reference lines it
restates (paths relative to /root/reference).  Bit-exact domain.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np


def cdiv(a: int, b: int) -> int:
    return -(a // -b)


@dataclass
class Policy:
    """Union of HEAD's PagedEvictConfig (vllm/config.py:933-991) and the README-era
    percentage fields (benchmarks/llama3.1_8b_test.txt:1, kvcache_prunner_old.py)."""
    block_size: int = 16
    cache_prune_type: str = "budget"          # "budget" | "percentage"
    cache_budget: int = 1024                  # budget mode
    evict_freq: int = 2                       # percentage mode
    initial_blocks: int = 1
    num_block_merge: int = 2                  # percentage mode only
    prompt_evict_method: str = "streamingLLM"
    # "streamingLLM" | "value_l2" | "random" (HEAD as shipped for local/global)
    # | "streamingLLM-1" | "inverse_key_l2" (slot-zeroing variants, W2)
    decode_evict_method: str = "value_l2"
    topk_blocks: int = -1                     # -1 = global, >1 = local window
    disable_evict_prefill: bool = False


# --------------------------------------------------------------------------- P1/P2
def prompt_prune_layout(q_len: int, p: Policy) -> Tuple[int, int, int]:
    """-> (kept, pruned, first): the pruned prompt keeps source tokens
    [0, first) U [first + pruned, q_len).

    budget:     KVCachePruner.get_pruned_length / prune_prompt_inplace streamingLLM
                (vllm/attention/kvcache_prunner.py:284-308, 183-218)
    percentage: kvcache_prunner_old.py:141-200 + core/page_evict_kv_util_bak.py:6-31
    """
    bs = p.block_size
    if p.cache_prune_type == "budget":
        if q_len <= p.cache_budget:                       # kvcache_prunner.py:289
            return q_len, 0, min(q_len, bs)
        rem = q_len % bs
        first = bs                                        # :296 (initial block)
        end_mid = q_len - bs - rem                        # :299
        mid = end_mid - first
        mid_keep = max(p.cache_budget - first - bs, 0)    # :303-305
        pruned = max(mid - mid_keep, 0)                   # :198-200
        return q_len - pruned, pruned, first
    elif p.cache_prune_type == "percentage":
        if q_len <= (p.initial_blocks + p.num_block_merge) * bs:   # _old.py:146
            return q_len, 0, min(q_len, p.initial_blocks * bs)
        rem = q_len % bs
        first = p.initial_blocks * bs                     # _old.py:155
        end_mid = q_len - p.num_block_merge * bs - rem    # _old.py:158
        mid = end_mid - first
        pruned = int(mid / p.evict_freq)                  # _old.py:171
        pruned = bs * math.floor(pruned / bs)             # _old.py:173
        return q_len - pruned, pruned, first
    raise ValueError(p.cache_prune_type)


def prompt_gather_index(q_len: int, p: Policy) -> np.ndarray:
    """Source row of every kept prompt token, in cache order (R1):
    cat(K[0:first], K[first+pruned:end_mid], K[end_mid:q_len])
    (kvcache_prunner.py:203-215)."""
    kept, pruned, first = prompt_prune_layout(q_len, p)
    j = np.arange(kept, dtype=np.int64)
    return np.where(j < first, j, j + pruned)


def num_required_blocks_after_prune(q_len: int, p: Policy) -> Tuple[int, int]:
    """Block-manager twin (vllm/core/page_evict_kv_util.py:6-36; note `<` at :10).
    -> (num_blocks, kept_tokens)."""
    bs = p.block_size
    if p.cache_prune_type == "budget" and q_len < p.cache_budget:
        return cdiv(q_len, bs), q_len
    kept, _, _ = prompt_prune_layout(q_len, p)
    return cdiv(kept, bs), kept


# ------------------------------------------------------------------------------ C3
def prefill_slot_mapping(block_table: Sequence[int], query_len: int, bs: int) -> List[int]:
    """TorchCUDASDPAMetadataBuilder._compute_slot_mapping prompt branch
    (vllm/attention/backends/torch_cuda_sdpa.py:577-589).  `block_table` is the
    already-short (post-prune) table; remainder uses the ORIGINAL query_len."""
    out: List[int] = []
    rem = query_len % bs
    if rem == 0:
        for b in block_table:
            out.extend(range(b * bs, b * bs + bs))
    else:
        for b in block_table[:-1]:
            out.extend(range(b * bs, b * bs + bs))
        out.extend(range(block_table[-1] * bs, block_table[-1] * bs + rem))
    return out


def decode_slot_mapping(block_table: Sequence[int], context_len: int, bs: int) -> int:
    """Decode branch (torch_cuda_sdpa.py:591-598): last table entry, logical
    context_len % bs."""
    return block_table[-1] * bs + (context_len % bs)


# --------------------------------------------------------------------------- P3/C2
def decode_trigger(seq_len: int, kv_len_before: int, p: Policy) -> bool:
    """seq_len = logical length incl. the token being decoded; kv_len_before =
    physical KV length incl. that token, before the -= block_size adjust.

    budget:     torch_cuda_sdpa.py:1390-1407,1491-1500; model_runner.py:540-543
    percentage: torch_cuda_sdpa.py:1310 (commented) + page_evict_kv_util_bak.py:69
    """
    bs = p.block_size
    if p.decode_evict_method == "inverse_key_l2":        # model_runner.py:545-547: never frees
        return False
    if p.cache_prune_type == "budget":
        return seq_len > p.cache_budget and seq_len % bs == 0
    return (seq_len % (bs * p.evict_freq) == 0
            and kv_len_before > (p.initial_blocks + p.num_block_merge) * bs)


def score_trigger(seq_len: int, p: Policy) -> bool:
    """The last block of the row becomes full with this step's token, so it gets its
    value_l2 score now (intended semantics of torch_cuda_sdpa.py:1518-1526; the build
    scores every block when it fills, see DESIGN.md deviation D3)."""
    return seq_len % p.block_size == 0


def victim_candidates(n_blocks_before: int, p: Policy) -> Tuple[int, int]:
    """[lo, hi) table indices eligible for eviction (torch_cuda_sdpa.py:1404-1405,
    1528-1534): streamingLLM -> {1}; global -> [1, n-1); local -> [1, 1+topk)
    (clamped to n-1: the just-filled last block is never a candidate)."""
    if p.decode_evict_method in ("streamingLLM", "streamingLLM-1"):
        return 1, 2
    if p.topk_blocks is None or p.topk_blocks < 0:
        return 1, n_blocks_before - 1
    return 1, min(1 + p.topk_blocks, n_blocks_before - 1)


def random_victim(rng, n_blocks_before: int, p: Policy) -> int:
    """HEAD as shipped (`enable_random_evict = True`, torch_cuda_sdpa.py:802): one
    `random.randint` per triggered row from the process-wide Python RNG every layer re-seeds
    with 1234 in its constructor (:796); window :1504-1513 -- local [1, topk] (NOT clamped
    to the table), global [1, n-2]."""
    if p.topk_blocks is not None and p.topk_blocks > 0:
        return rng.randint(1, 1 + p.topk_blocks - 1)
    return rng.randint(1, n_blocks_before - 1 - 1)


def zero_slot(rng, seq_len: int, table: Sequence[int], p: Policy) -> Optional[int]:
    """W2: the cache slot whose K and V rows are overwritten with zeros on this step, or None.
    streamingLLM-1 (torch_cuda_sdpa.py:1390-1401): over budget, not a block boundary ->
      slot (seq_len % bs - 1) of the table's 2nd block (the next victim is zeroed token by token).
    inverse_key_l2 (:1444-1454): over budget -> a random slot of a random middle block, two
      randint draws per row (block first)."""
    bs = p.block_size
    if seq_len <= p.cache_budget:
        return None
    if p.decode_evict_method == "streamingLLM-1":
        if seq_len % bs == 0:
            return None
        return table[1] * bs + (seq_len % bs - 1)
    if p.decode_evict_method == "inverse_key_l2":
        blk = rng.randint(1, len(table) - 2)
        off = rng.randint(0, bs - 1)
        return table[blk] * bs + off
    return None


def argmin_lowest_index(scores: np.ndarray, lo: int, hi: int) -> int:
    """Linear scan with strict `<` (torch_cuda_sdpa.py:1536-1541): ties -> lowest
    index.  `scores` fp32 1-D."""
    rmv = lo
    for i in range(lo, hi):
        if scores[i] < scores[rmv]:
            rmv = i
    return rmv


# ------------------------------------------------------------------------------ C1
def compact_block_tables(tables: np.ndarray, n_blocks: Sequence[int],
                         victims: Dict[int, int]) -> np.ndarray:
    """Delete column victims[row] from each triggered row, shift left, 0-pad.
    Semantics = torch_cuda_sdpa_v1.py:1485-1551 (lists + make_tensor_with_pad) =
    HEAD row 0 (torch_cuda_sdpa.py:1562-1586); HEAD rows >= 1 are buggy
    (SURVEY.md 0.9).  Returned width stays M (device buffers are fixed-width); the
    reference-shaped view is `narrow_compacted`."""
    B, M = tables.shape
    out = np.zeros_like(tables)
    for r in range(B):
        n = int(n_blocks[r])
        row = list(tables[r, :n])
        if r in victims:
            del row[victims[r]]
        out[r, :len(row)] = row
    return out


def narrow_compacted(out: np.ndarray, n_after: Sequence[int]) -> np.ndarray:
    """make_tensor_with_pad width = longest live row (vllm/utils.py make_tensor_with_pad)."""
    w = max(int(x) for x in n_after) if len(n_after) else 0
    return out[:, :w]
