"""Host-side integer arithmetic of the eviction policy (product code; the checker's
independent restatement lives in oracle/policy.py and the golden vectors in
tests/golden/ref_policy.json pin both).

  budget mode      vllm/core/page_evict_kv_util.py:6-36,
                   vllm/attention/kvcache_prunner.py:284-326 (HEAD flags)
  percentage mode  vllm/core/page_evict_kv_util_bak.py:6-31,69,
                   vllm/attention/kvcache_prunner_old.py:141-200 (README flags)
"""
from __future__ import annotations

from typing import Tuple


def cdiv(a: int, b: int) -> int:
    return -(a // -b)


def prompt_prune_layout(q_len: int, cfg, block_size: int) -> Tuple[int, int, int]:
    """-> (kept, pruned, first).  The streamingLLM prompt prune keeps source tokens
    [0, first) and [first + pruned, q_len): the sink block, then the newest middle
    tokens, then the recent window (kvcache_prunner.py:183-218)."""
    bs = block_size
    if cfg is None:
        return q_len, 0, q_len
    if cfg.cache_prune_type == "budget":
        if q_len <= cfg.cache_budget:
            return q_len, 0, min(q_len, bs)
        tail = bs + q_len % bs                       # recent window: last full block + remainder
        middle = q_len - bs - tail                   # tokens between sink block and window
        keep_mid = max(cfg.cache_budget - 2 * bs, 0)
        pruned = max(middle - keep_mid, 0)
        return q_len - pruned, pruned, bs
    sink = cfg.initial_blocks * bs
    if q_len <= (cfg.initial_blocks + cfg.num_block_merge) * bs:
        return q_len, 0, min(q_len, sink)
    middle = q_len - cfg.num_block_merge * bs - q_len % bs - sink
    pruned = (int(middle / cfg.evict_freq) // bs) * bs
    return q_len - pruned, pruned, sink


def num_required_blocks_after_prune(q_len: int, cfg, block_size: int) -> Tuple[int, int]:
    """Block-manager view: (blocks to allocate, tokens cached) for a new prompt.  Note the
    strict `<` of page_evict_kv_util.py:10 (q_len == budget takes the prune branch and
    still keeps everything)."""
    if cfg is None or (cfg.cache_prune_type == "budget" and q_len < cfg.cache_budget):
        return cdiv(q_len, block_size), q_len
    kept = prompt_prune_layout(q_len, cfg, block_size)[0]
    return cdiv(kept, block_size), kept


def decode_trigger(seq_len: int, kv_len_before: int, cfg, block_size: int) -> bool:
    """Does the decode step that brings the sequence to logical length `seq_len` evict
    one block?  kv_len_before = physical KV length including the new token.
    budget: torch_cuda_sdpa.py:1390-1407, model_runner.py:540-543.
    percentage: torch_cuda_sdpa.py:1310 + page_evict_kv_util_bak.py:69."""
    if cfg is None or cfg.decode_evict_method == "inverse_key_l2":   # model_runner.py:545-547: never frees
        return False
    n_blocks = cdiv(kv_len_before, block_size)
    if n_blocks < 3:                                 # sink + victim + just-filled block
        return False
    if cfg.cache_prune_type == "budget":
        return seq_len > cfg.cache_budget and seq_len % block_size == 0
    return (seq_len % (block_size * cfg.evict_freq) == 0
            and n_blocks * block_size > (cfg.initial_blocks + cfg.num_block_merge) * block_size)


def random_victim(rng, n_blocks: int, cfg) -> int:
    """HEAD as shipped (`enable_random_evict = True`, torch_cuda_sdpa.py:802): the victim of a
    triggered row is one `randint` from the process-wide Python RNG (:796 seeds it with 1234 in
    every layer's constructor) -- local window [1, topk] (:1504-1509, not clamped to the table),
    global [1, n-2] (:1510-1513)."""
    if cfg.topk_blocks is not None and cfg.topk_blocks > 0:
        return rng.randint(1, cfg.topk_blocks)
    return rng.randint(1, n_blocks - 2)


def zero_slot(rng, seq_len: int, table, cfg, block_size: int) -> int:
    """W2: the cache slot whose K and V rows are zeroed on this decode step, or -1.
    streamingLLM-1 (torch_cuda_sdpa.py:1390-1401): over budget and off a block boundary -> slot
      (seq_len % bs - 1) of the table's 2nd block.
    inverse_key_l2 (:1444-1454): over budget -> a random slot of a random middle block, two draws
      per row (block, then slot)."""
    if seq_len <= cfg.cache_budget:
        return -1
    m = cfg.decode_evict_method
    if m == "streamingLLM-1":
        if seq_len % block_size == 0:
            return -1
        return int(table[1]) * block_size + (seq_len % block_size - 1)
    if m == "inverse_key_l2":
        blk = rng.randint(1, len(table) - 2)
        off = rng.randint(0, block_size - 1)
        return int(table[blk]) * block_size + off
    return -1
