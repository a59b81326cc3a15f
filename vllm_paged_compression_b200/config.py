"""PagedEvictConfig / CacheConfig / CLI flags of the paged-eviction path.

Mirrors vllm/config.py:933-991 (PagedEvictConfig), the `paged_evict_config` field of
CacheConfig (vllm/config.py:1040,1051) and the argparse block of
vllm/engine/arg_utils.py:202-212,971-1036,1132-1156.  Both flag sets are accepted:

  HEAD  : --enable-paged-eviction --disable-evict-prefill --cache-prune-type budget
          --evict-method {streamingLLM,streamingLLM-1,inverse_key_l2,local,global}
          --cache-budget N --topk-blocks K
  README: --cache-prune-type percentage --prompt-evict-method streamingLLM
          --decode-evict-method {streamingLLM,value_l2} --evict-freq N
          (README.md:38-45; semantics from kvcache_prunner_old.py / page_evict_kv_util_bak.py,
          benchmarks/llama3.1_8b_test.txt:1 -> initial_blocks=1, num_blocks_merge=2)

`--evict-method` resolves into a (prompt, decode) method pair:

  streamingLLM    prompt streamingLLM slice, decode: evict table entry 1
  streamingLLM-1  the same, plus W2: while over budget and off a block boundary, slot
                  (seq_len % bs - 1) of the table's 2nd block is zeroed (torch_cuda_sdpa.py:1396-1401)
  inverse_key_l2  prompt: token-level 1/||K|| top-k drop; decode: W2 only -- a random slot of a random
                  middle block is zeroed every step over budget, nothing is ever freed
                  (torch_cuda_sdpa.py:1444-1454, model_runner.py:545-547)
  global / local  decode value_l2 argmin over the global / local window (kvcache_prunner.py:13
                  sub_evict_method = "value_l2"); prompt = streamingLLM slice by default.
                  `head_random_evict=True` (--head-random-evict, env PC_HEAD_RANDOM_EVICT=1)
                  reproduces HEAD as shipped: `enable_random_evict = True` (torch_cuda_sdpa.py:802)
                  evicts a random window block drawn from Python's `random` (seeded 1234), and the
                  prompt is pruned token by token (value_l2 top-k drop, kvcache_prunner.py:220-279).
"""
from __future__ import annotations

import argparse
import logging
import os
from dataclasses import dataclass, field
from typing import Optional

logger = logging.getLogger(__name__)

_HEAD_METHODS = ["streamingLLM", "streamingLLM-1", "inverse_key_l2", "local", "global"]
# prompt: the streamingLLM slice, or the per-head token top-k drop of kvcache_prunner.py:220-279
_PROMPT_METHODS = ["streamingLLM", "token_value_l2", "token_inverse_key_l2"]
# decode: "random" = HEAD's enable_random_evict for local / global; the last two are the W2 variants
_DECODE_METHODS = ["streamingLLM", "value_l2", "random", "streamingLLM-1", "inverse_key_l2"]


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
                 num_block_merge: int = 2,
                 head_random_evict: Optional[bool] = None) -> None:
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
        if head_random_evict is None:
            head_random_evict = os.environ.get("PC_HEAD_RANDOM_EVICT", "0") == "1"
        self.head_random_evict = bool(head_random_evict)
        self._verify_args()

    def _verify_args(self) -> None:
        if self.cache_prune_type not in ["percentage", "budget"]:
            raise ValueError("Cache prune type must be either 'percentage' or 'budget'. Got "
                             f"{self.cache_prune_type}.")
        # resolve --evict-method into (prompt, decode) methods
        if self.decode_evict_method is None or self.prompt_evict_method is None:
            em = self.evict_method or "streamingLLM"
            if em not in _HEAD_METHODS:
                raise ValueError(f"Unknown evict_method {em!r}")
            if em in ("streamingLLM", "streamingLLM-1"):
                prompt, decode = "streamingLLM", em
            elif em == "inverse_key_l2":
                prompt, decode = "token_inverse_key_l2", "inverse_key_l2"
            elif self.head_random_evict:
                prompt, decode = "token_value_l2", "random"
            else:
                prompt, decode = "streamingLLM", "value_l2"
            self.prompt_evict_method = self.prompt_evict_method or prompt
            self.decode_evict_method = self.decode_evict_method or decode
        if self.prompt_evict_method not in _PROMPT_METHODS:
            raise ValueError(f"prompt_evict_method must be one of {_PROMPT_METHODS}, got "
                             f"{self.prompt_evict_method!r}")
        if self.decode_evict_method not in _DECODE_METHODS:
            raise ValueError(f"decode_evict_method must be one of {_DECODE_METHODS}, got "
                             f"{self.decode_evict_method!r}")
        if (self.decode_evict_method in ("random", "streamingLLM-1", "inverse_key_l2")
                or self.prompt_evict_method != "streamingLLM") and self.cache_prune_type != "budget":
            raise ValueError(f"{self.decode_evict_method!r} / {self.prompt_evict_method!r} exist in budget "
                             "mode only (kvcache_prunner.py:11)")
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
        if self.evict_method == "global" or (self.evict_method != "local"
                                              and self.decode_evict_method in ("value_l2", "random")):
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
    parser.add_argument("--head-random-evict", action="store_true", default=False,
                        help="local / global as HEAD ships them (enable_random_evict = True, "
                             "torch_cuda_sdpa.py:802): random window block at decode time, token-level "
                             "value_l2 prompt prune.")
    parser.add_argument("--block-size", type=int, default=16, choices=[8, 16, 32],
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
        evict_freq=args.evict_freq,
        head_random_evict=True if getattr(args, "head_random_evict", False) else None)
