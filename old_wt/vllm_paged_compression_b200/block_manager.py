"""Host block bookkeeping of the paged-eviction path: physical (`_num_cached_slots`) vs
logical (`_num_full_slots`) slot accounting, prune-aware prompt allocation, and freeing the
block a decode step evicted.

Restates only the slices of the fork's block manager that the path touches:
  vllm/core/block/block_table.py:59-66,94-127,136-196,317-323,456-466  (BlockTable)
  vllm/core/block/common.py:293-296                                    (BlockList.remove)
  vllm/core/block_manager.py:113-156,537-543   (can_allocate / get_seq_kv_len / free_prunned_blocks)
  vllm/core/scheduler.py:1879-1890             (Scheduler.update_block_tables)
Prefix caching, copy-on-write, swapping and sliding windows are the stock vLLM features
around it and are out of scope.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, Dict, List, Optional

from .config import PagedEvictConfig
from .policy import cdiv, num_required_blocks_after_prune


class NaiveBlockAllocator:
    """Free-list allocator with the reuse order of vllm/core/block/naive_block.py:131-145
    (allocate = popleft, free = appendleft: the block freed last is handed out first), so
    block ids match the reference traces."""

    def __init__(self, num_blocks: int) -> None:
        self.num_blocks = num_blocks
        self._free: Deque[int] = deque(range(num_blocks))

    def get_num_free_blocks(self) -> int:
        return len(self._free)

    def allocate(self) -> int:
        if not self._free:
            raise RuntimeError("No free blocks available")      # BlockAllocator.NoFreeBlocksError
        return self._free.popleft()

    def free(self, block_id: int) -> None:
        self._free.appendleft(block_id)


class BlockTable:
    def __init__(self, block_size: int, allocator: NaiveBlockAllocator,
                 paged_evict_config: Optional[PagedEvictConfig] = None) -> None:
        self._block_size = block_size
        self._allocator = allocator
        self.paged_evict_config = paged_evict_config
        self._blocks: List[int] = []
        self._num_full_slots = 0        # logical tokens
        self._num_cached_slots = 0      # tokens physically in the KV cache

    # block_table.py:94-127 + 317-323
    def allocate(self, num_tokens: int) -> None:
        assert not self._blocks and num_tokens > 0
        bs = self._block_size
        if self.paged_evict_config is None:
            n_blocks, cached = cdiv(num_tokens, bs), num_tokens
        else:
            n_blocks, _ = num_required_blocks_after_prune(num_tokens, self.paged_evict_config, bs)
            rem = num_tokens % bs
            cached = n_blocks * bs if rem == 0 else (n_blocks - 1) * bs + rem
        self._blocks = [self._allocator.allocate() for _ in range(n_blocks)]
        self._num_full_slots = num_tokens
        self._num_cached_slots = cached

    # block_table.py:136-196 (one decode token; ensure_num_empty_slots + append)
    def append_slots(self, num_tokens: int = 1) -> None:
        assert self._blocks, "no blocks have been allocated"
        need = cdiv(self._num_cached_slots + num_tokens, self._block_size) - len(self._blocks)
        for _ in range(max(need, 0)):
            self._blocks.append(self._allocator.allocate())
        self._num_full_slots += num_tokens
        self._num_cached_slots += num_tokens

    # block_table.py:456-466 + common.py:293-296
    def get_seq_kv_len(self) -> int:
        return self._num_cached_slots

    def free_prunned_blocks(self, rmv_block_idx: int) -> None:
        block = self._blocks[rmv_block_idx]
        self._allocator.free(block)
        self._num_cached_slots -= self._block_size     # evicted blocks are always full
        del self._blocks[rmv_block_idx]

    def free(self) -> None:
        for b in self._blocks:
            self._allocator.free(b)
        self._blocks = []

    @property
    def physical_block_ids(self) -> List[int]:
        return self._blocks

    @property
    def num_full_slots(self) -> int:
        return self._num_full_slots


class SelfAttnBlockSpaceManager:
    """The calls the scheduler/engine make on this path (block_manager.py)."""

    def __init__(self, block_size: int, num_gpu_blocks: int,
                 paged_evict_config: Optional[PagedEvictConfig] = None) -> None:
        self.block_size = block_size
        self.paged_evict_config = paged_evict_config
        self.allocator = NaiveBlockAllocator(num_gpu_blocks)
        self.block_tables: Dict[int, BlockTable] = {}

    def can_allocate(self, num_prompt_tokens: int) -> bool:
        """block_manager.py:113-156: the post-prune block count is what must fit."""
        n, _ = num_required_blocks_after_prune(num_prompt_tokens, self.paged_evict_config,
                                               self.block_size)
        return n <= self.allocator.get_num_free_blocks()

    def allocate(self, seq_id: int, num_prompt_tokens: int) -> None:
        bt = BlockTable(self.block_size, self.allocator, self.paged_evict_config)
        bt.allocate(num_prompt_tokens)
        self.block_tables[seq_id] = bt

    def can_append_slots(self, num_seqs: int) -> bool:
        return num_seqs <= self.allocator.get_num_free_blocks()

    def append_slots(self, seq_id: int) -> None:
        self.block_tables[seq_id].append_slots(1)

    def get_block_table(self, seq_id: int) -> List[int]:
        return self.block_tables[seq_id].physical_block_ids

    def get_seq_kv_len(self, seq_id: int) -> int:               # block_manager.py:537-538
        return self.block_tables[seq_id].get_seq_kv_len()

    def free_prunned_blocks(self, seq_id: int, rmv_block_idx: int) -> None:   # :540-543
        self.block_tables[seq_id].free_prunned_blocks(rmv_block_idx)

    def free(self, seq_id: int) -> None:
        self.block_tables.pop(seq_id).free()
