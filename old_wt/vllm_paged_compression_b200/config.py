"""PagedEvictConfig / CacheConfig / CLI flags of the paged-eviction path.

Mirrors vllm/config.py:933-991 (PagedEvictConfig), the `paged_evict_config` field of
CacheConfig (vllm/config.py:1040,1051) and the argparse block of
vllm/engine/arg_utils.py:202-212,971-1036,1132-1156.  Both flag sets are accepted:

  HEAD  : --enable-paged-eviction --disable-evict-prefill --cache-prune-type budget
          --evict-method {streamingLLM,local,global} --cache-budget N --topk-blocks K
  README: --cache-prune-type percentage --prompt-evict-method streamingLLM
          --decode-evict-method {streamingLLM,value_l2} --evict-freq N
          (README.md:38-45; semantics from kvcache_prunner_old.py / page_evict_kv_util_bak.py,
          benchmarks/llama3.1_8b_test.txt:1 -> initial_blocks=1, num_blocks_merge=2)

`--evict-method` is an alias: streamingLLM -> prompt streamingLLM + decode streamingLLM;
global/local -> prompt streamingLLM slice + decode value_l2 over the global / local
window (kvcache_prunner.py:13 sub_evict_method = "value_l2").  The token-level prompt
scoring of HEAD's global/local and the slot-zeroing methods (streamingLLM-1,
inverse_key_l2) are outside the hot path and rejected loudly.
"""
from __future__ import annotations

import argparse
import logging
from dataclasses import dataclass, field
from typing import Optional

logger = logging.getLogger(__name__)

_HEAD_METHODS = ["streamingLLM", "streamingLLM-1", "inverse_key_l2", "local", "global"]


class PagedEvictConfig:
    """Field names follow vllm/config.py:933-991; README-era fields added alongside."""

    def __init__(self,
                 disable_evict_prefill: bool = False,
                 cache_prune_type: str = "budget",
                 evict_method: Optional[str] = "streamingLLM",
                 cache_budget: int = 1024,
                 topk_blocks: int = 3,
                 *,
                 prompt_evict_method: Optional[str] = None,
                 decode_evict_method: Optional[str] = None,
                 evict_freq: int = 2,
                 initial_blocks: int = 1,
                 num_block_merge: int = 2) -> None:
        self.disable_evict_prefill = disable_evict_prefill
        self.cache_prune_type = cache_prune_type
        self.evict_method = evict_method
        self.cache_budget = cache_budget
        self.topk_blocks = topk_blocks
        self.initial_blocks = initial_blocks       # vllm/config.py:967: hard-coded 1
        self.evict_freq = evict_freq
        self.num_block_merge = num_block_merge
        self.prompt_evict_method = prompt_evict_method
        self.decode_evict_method = decode_evict_method
        self._verify_args()

    def _verify_args(self) -> None:
        if self.cache_prune_type not in ["percentage", "budget"]:
            raise ValueError("Cache prune type must be either 'percentage' or 'budget'. Got "
                             f"{self.cache_prune_type}.")
        # resolve the --evict-method alias into (prompt, decode) methods
        if self.decode_evict_method is None or self.prompt_evict_method is None:
            em = self.evict_method or "streamingLLM"
            if em not in _HEAD_METHODS:
                raise ValueError(f"Unknown evict_method {em!r}")
            if em in ("streamingLLM-1", "inverse_key_l2"):
                raise NotImplementedError(
                    f"evict_method={em!r} (slot-zeroing variants) is outside the B200 hot path; "
                    "supported: streamingLLM, local, global")
            self.prompt_evict_method = self.prompt_evict_method or "streamingLLM"
            self.decode_evict_method = self.decode_evict_method or (
                "streamingLLM" if em == "streamingLLM" else "value_l2")
        if self.prompt_evict_method != "streamingLLM":
            raise NotImplementedError("prompt_evict_method must be 'streamingLLM' (token-level prompt "
                                      "scoring is outside the hot path)")
        if self.decode_evict_method not in ("streamingLLM", "value_l2"):
            raise ValueError(f"decode_evict_method must be streamingLLM or value_l2, got "
                             f"{self.decode_evict_method!r}")
        if self.initial_blocks != 1:
            raise ValueError("initial_blocks must be 1 (vllm/config.py:967)")
        if self.cache_prune_type == "budget":
            if self.cache_budget < 0:
                raise ValueError(f"Cache budget must be positive. Got {self.cache_budget}.")
            if self.cache_budget < 128:          # vllm/config.py:977-980
                logger.warning("Cache budget is set to less than 128 tokens. "
                               "Setting cache budget to 128 tokens.")
                self.cache_budget = 128
        else:
            if self.evict_freq < 1:
                raise ValueError(f"evict_freq must be >= 1, got {self.evict_freq}")
            if self.num_block_merge < 1:
                raise ValueError("Number of blocks to merge must be positive. Got "
                                 f"{self.num_block_merge}.")
        if self.evict_method == "local" and self.topk_blocks <= 1:   # vllm/config.py:982-984
            raise ValueError("Topk blocks must be > 1 when using 'local' evict method. "
                             f"Got {self.topk_blocks}.")
        if self.evict_method == "global" or (self.evict_method not in ("local",)
                                              and self.decode_evict_method == "value_l2"
                                              and self.evict_method != "local"):
            self.topk_blocks = -1                                     # vllm/config.py:986-988

    def verify_with_block_size(self, block_size: int) -> None:
        """Slot mapping and whole-block eviction need budget % block_size == 0
        (torch_cuda_sdpa.py:595-597 relies on it) and, like KVCachePruner.__init__
        (kvcache_prunner.py:15), budget >= 3 blocks."""
        if self.cache_prune_type == "budget":
            if self.cache_budget % block_size != 0:
                raise ValueError(f"cache_budget ({self.cache_budget}) must be a multiple of block_size "
                                 f"({block_size})")
            assert self.cache_budget >= 3 * block_size, \
                "Cache budget must be at least 3 times the original block size"

    def __repr__(self) -> str:
        return (f"PagedEvictConfig(prune={self.cache_prune_type}, prompt={self.prompt_evict_method}, "
                f"decode={self.decode_evict_method}, budget={self.cache_budget}, "
                f"evict_freq={self.evict_freq}, topk={self.topk_blocks}, "
                f"disable_evict_prefill={self.disable_evict_prefill})")


@dataclass
class CacheConfig:
    """The slice of vllm/config.py:993-1059 the attention path reads."""
    block_size: int = 16
    cache_dtype: str = "auto"
    paged_evict_config: Optional[PagedEvictConfig] = None
    num_gpu_blocks: Optional[int] = None
    max_num_seqs: int = 256

    def __post_init__(self):
        if self.paged_evict_config is not None:
            self.paged_evict_config.verify_with_block_size(self.block_size)


def add_cli_args(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    """The paged-eviction argparse block (vllm/engine/arg_utils.py:971-1036) plus the
    README spellings it commented out (:1003-1008)."""
    parser.add_argument("--enable-paged-eviction", action="store_true", default=False,
                        help="Enable paged eviction algorithm to prune the KV cache.")
    parser.add_argument("--disable-evict-prefill", action="store_true", default=False,
                        help="Compute prefill attention first and prune the prompt KV afterwards.")
    parser.add_argument("--cache-prune-type", type=str, choices=["percentage", "budget"],
                        default="budget", help="The KV Cache prune type")
    parser.add_argument("--evict-method", type=str, choices=_HEAD_METHODS, default=None,
                        help="The paged eviction method to use (HEAD spelling).")
    parser.add_argument("--prompt-evict-method", type=str, default=None,
                        help="README spelling: prompt-time method (streamingLLM).")
    parser.add_argument("--decode-evict-method", type=str, default=None,
                        help="README spelling: decode-time method (streamingLLM | value_l2).")
    parser.add_argument("--evict-freq", type=int, default=2,
                        help="How often to trigger eviction in terms of number of blocks")
    parser.add_argument("--cache-budget", type=int, default=1024,
                        help="The number of tokens to keep in the KV cache for each request after pruning")
    parser.add_argument("--topk-blocks", type=int, default=3,
                        help="The number of top-k blocks to compare when pruning")
    parser.add_argument("--block-size", type=int, default=16, choices=[16, 32],
                        help="Token block size")
    return parser


def paged_evict_config_from_args(args: argparse.Namespace) -> Optional[PagedEvictConfig]:
    """EngineArgs.create_engine_config slice (vllm/engine/arg_utils.py:1132-1145)."""
    if not args.enable_paged_eviction:
        return None
    evict_method = args.evict_method
    if evict_method is None and args.prompt_evict_method is None and args.decode_evict_method is None:
        evict_method = "streamingLLM"
    topk = args.topk_blocks
    if evict_method is None and (args.decode_evict_method == "value_l2"):
        topk = -1       # README flags have no window -> global
    return PagedEvictConfig(
        disable_evict_prefill=args.disable_evict_prefill,
        cache_prune_type=args.cache_prune_type,
        evict_method=evict_method,
        cache_budget=args.cache_budget,
        topk_blocks=topk,
        prompt_evict_method=args.prompt_evict_method,
        decode_evict_method=args.decode_evict_method,
        evict_freq=args.evict_freq)
