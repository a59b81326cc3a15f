"""TORCH_SDPA paged-eviction attention backend, B200-native.

Host-side mirror of vllm/attention/backends/torch_cuda_sdpa.py (class names, constructor
and forward signatures, metadata fields, `get_rmv_block_idx` contract) and of
vllm/attention/ops/paged_attn.py (`PagedAttention`).  The per-request Python loops,
torch.cat pruning, host argmin and per-row table rebuilding of the reference are replaced
by at most three kernel launches per layer and step, all through the C ABI (ops.py):

  prefill : prune_prompt_and_cache (R1+W1+S3)  ->  prefill_attn_topleft_causal (A1)
  decode  : reshape_and_cache (W1) -> [evict_select_compact (P3+S4+C1+C2) on trigger
            steps] -> paged_decode_attn_score (A2+S2)

Declared deviations from HEAD (DESIGN.md "Deviations"): value_l2 instead of HEAD's
hard-coded random eviction, fp32 fixed-order scores on the true V layout
(0.4), v1 compaction semantics for every row (0.9), scores kept for every live request
instead of being dropped whenever the batch composition changes
(block_l2norm_manager.py:34-62).
"""
from __future__ import annotations

import os
import random
from dataclasses import dataclass, field
from itertools import accumulate
from typing import Any, Dict, List, Optional, Tuple, Type

import numpy as np
import torch

from . import ops
from .abstract import (PAD_SLOT_ID, AttentionBackend, AttentionImpl, AttentionMetadata,
                       AttentionMetadataBuilder, AttentionType, CommonAttentionState)
from .config import CacheConfig, PagedEvictConfig
from .policy import decode_trigger, prompt_prune_layout, random_victim, zero_slot


# =============================================================================== ops
class PagedAttention:
    """vllm/attention/ops/paged_attn.py:28-254, same static surface."""

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [64, 128]          # built instantiations (reference list: paged_attn.py:36-37)

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)     # paged_attn.py:40-46

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int,
                       head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        x = 16 // kv_cache.element_size()                                 # paged_attn.py:49-62
        num_blocks = kv_cache.shape[1]
        key_cache = kv_cache[0].view(num_blocks, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1].view(num_blocks, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype="auto",
                             k_scale=1.0, v_scale=1.0) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(),
                              kv_cache_dtype, k_scale, v_scale)           # paged_attn.py:65-84

    @staticmethod
    def forward_decode(query, key_cache, value_cache, block_tables, seq_lens, max_seq_len,
                       kv_cache_dtype="auto", num_kv_heads=None, scale=1.0, alibi_slopes=None,
                       k_scale=1.0, v_scale=1.0, tp_rank=0, blocksparse_local_blocks=0,
                       blocksparse_vert_stride=0, blocksparse_block_size=64,
                       blocksparse_head_sliding_step=0) -> torch.Tensor:
        """paged_attn.py:87-191.  The v1/v2 heuristic disappears: one split-KV kernel."""
        output = torch.empty_like(query)
        ops.paged_attention_v1(output, query, key_cache, value_cache, num_kv_heads, scale,
                               block_tables, seq_lens, value_cache.shape[3], max_seq_len,
                               alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
                               blocksparse_local_blocks, blocksparse_vert_stride,
                               blocksparse_block_size, blocksparse_head_sliding_step)
        return output

    @staticmethod
    def swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst) -> None:
        ops.swap_blocks(src_kv_cache[0], dst_kv_cache[0], src_to_dst)     # paged_attn.py:233-243
        ops.swap_blocks(src_kv_cache[1], dst_kv_cache[1], src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        for kv in kv_caches:                                              # paged_attn.py:246-254
            # a block is one contiguous run in each of the K and V pools
            ops.copy_blocks(kv[0], kv[1], src_to_dists)


# ========================================================================== backend
class TorchCUDASDPABackend(AttentionBackend):
    """torch_cuda_sdpa.py:43-88."""

    @staticmethod
    def get_name() -> str:
        return "TORCH_SDPA"

    @staticmethod
    def get_impl_cls() -> Type["TorchCUDASDPABackendImpl"]:
        return TorchCUDASDPABackendImpl

    @staticmethod
    def get_metadata_cls() -> Type["TorchCUDASDPAMetadata"]:
        return TorchCUDASDPAMetadata

    @staticmethod
    def get_state_cls() -> Type[CommonAttentionState]:
        return CommonAttentionState

    @staticmethod
    def get_builder_cls() -> Type["TorchCUDASDPAMetadataBuilder"]:
        return TorchCUDASDPAMetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)


# ========================================================================= metadata
@dataclass
class TorchCUDASDPAMetadata(AttentionMetadata):
    """torch_cuda_sdpa.py:91-458: base fields + seq/kv length tensors + the fork's
    seq_kv_lens / seq_kv_lens_tensor / reqid_mapping_blocktables."""
    seq_lens_tensor: Optional[torch.Tensor]
    max_prefill_seq_len: int
    max_decode_seq_len: int
    use_cuda_graph: bool
    seq_lens: Optional[List[int]] = None
    seq_start_loc: Optional[torch.Tensor] = None
    context_lens_tensor: Optional[torch.Tensor] = None
    max_query_len: Optional[int] = None
    max_decode_query_len: Optional[int] = None
    query_start_loc: Optional[torch.Tensor] = None
    block_tables: Optional[torch.Tensor] = None
    _cached_prefill_metadata: Optional["TorchCUDASDPAMetadata"] = None
    _cached_decode_metadata: Optional["TorchCUDASDPAMetadata"] = None
    encoder_seq_lens: Optional[List[int]] = None
    encoder_seq_lens_tensor: Optional[torch.Tensor] = None
    max_encoder_seq_len: Optional[int] = None
    num_encoder_tokens: Optional[int] = None
    cross_slot_mapping: Optional[torch.Tensor] = None
    cross_block_tables: Optional[torch.Tensor] = None
    seq_kv_lens: Optional[List[int]] = None
    seq_kv_lens_tensor: Optional[torch.Tensor] = None
    reqid_mapping_blocktables: Optional[Dict[str, Dict[int, List[int]]]] = None

    def __post_init__(self):
        # per-step scratch shared by all layers (filled by the first impl that sees it)
        self._pc_step: Optional[Dict[str, Any]] = None

    @property
    def prefill_metadata(self) -> Optional["TorchCUDASDPAMetadata"]:
        if self.num_prefills == 0:
            return None
        if self._cached_prefill_metadata is None:
            n = self.num_prefills
            sl = lambda t, e: None if t is None else t[:e]
            self._cached_prefill_metadata = TorchCUDASDPAMetadata(
                num_prefills=n, num_prefill_tokens=self.num_prefill_tokens, num_decode_tokens=0,
                slot_mapping=sl(self.slot_mapping, self.num_prefill_tokens),
                multi_modal_placeholder_index_maps=self.multi_modal_placeholder_index_maps,
                seq_lens=sl(self.seq_lens, n), seq_lens_tensor=sl(self.seq_lens_tensor, n),
                max_query_len=self.max_query_len, max_prefill_seq_len=self.max_prefill_seq_len,
                max_decode_seq_len=0, query_start_loc=sl(self.query_start_loc, n + 1),
                context_lens_tensor=sl(self.context_lens_tensor, n),
                block_tables=sl(self.block_tables, n), use_cuda_graph=False,
                seq_kv_lens=sl(self.seq_kv_lens, n), seq_kv_lens_tensor=sl(self.seq_kv_lens_tensor, n))
        return self._cached_prefill_metadata

    @property
    def decode_metadata(self) -> Optional["TorchCUDASDPAMetadata"]:
        if self.num_decode_tokens == 0:
            return None
        if self._cached_decode_metadata is None:
            n = self.num_prefills
            sl = lambda t, s: None if t is None else t[s:]
            self._cached_decode_metadata = TorchCUDASDPAMetadata(
                num_prefills=0, num_prefill_tokens=0, num_decode_tokens=self.num_decode_tokens,
                slot_mapping=sl(self.slot_mapping, self.num_prefill_tokens),
                multi_modal_placeholder_index_maps=None,
                seq_lens_tensor=sl(self.seq_lens_tensor, n), max_prefill_seq_len=0,
                max_decode_seq_len=self.max_decode_seq_len, block_tables=sl(self.block_tables, n),
                use_cuda_graph=self.use_cuda_graph, seq_kv_lens=sl(self.seq_kv_lens, n),
                seq_kv_lens_tensor=sl(self.seq_kv_lens_tensor, n))
        return self._cached_decode_metadata

    def get_seq_lens(self, attn_type: str):
        """torch_cuda_sdpa.py:322-349 (decoder self-attention only)."""
        if attn_type in (AttentionType.DECODER, AttentionType.ENCODER_ONLY):
            return self.seq_lens, (self.seq_lens if self.seq_kv_lens is None else self.seq_kv_lens)
        raise AttributeError(f"Invalid attention type {str(attn_type)}")


def _h2d(arr: np.ndarray, device, pin_memory: bool) -> torch.Tensor:
    """async_tensor_h2d (vllm/utils.py) for a numpy array."""
    t = torch.from_numpy(arr)
    if pin_memory and torch.cuda.is_available():
        t = t.pin_memory()
    return t.to(device, non_blocking=True)


def prefill_slots_np(block_table: List[int], query_len: int, bs: int) -> np.ndarray:
    """Prompt branch of _compute_slot_mapping (torch_cuda_sdpa.py:577-589), vectorised:
    every block of the (already short) table contributes bs slots, the last one only
    query_len % bs when that is non-zero."""
    if not block_table:
        return np.empty(0, dtype=np.int64)
    t = np.asarray(block_table, dtype=np.int64)
    s = (t[:, None] * bs + np.arange(bs, dtype=np.int64)[None, :]).reshape(-1)
    rem = query_len % bs
    return s if rem == 0 else s[:len(s) - (bs - rem)]


class TorchCUDASDPAMetadataBuilder(AttentionMetadataBuilder[TorchCUDASDPAMetadata]):
    """torch_cuda_sdpa.py:461-697.  Reads the same attributes of `input_builder` /
    `runner` / `inter_data_list` the fork's builder reads (SURVEY.md 8c last bullet)."""

    def __init__(self, input_builder) -> None:
        self.slot_mapping: List[np.ndarray] = []
        self.prefill_seq_lens: List[int] = []
        self.context_lens: List[int] = []
        self.block_tables: List[List[int]] = []
        self.curr_seq_lens: List[int] = []
        self.num_prefills = 0
        self.num_prefill_tokens = 0
        self.num_decode_tokens = 0
        self.input_builder = input_builder
        self.runner = input_builder.runner
        self.sliding_window = input_builder.sliding_window
        self.block_size = input_builder.block_size
        self.paged_evict_config = self.runner.cache_config.paged_evict_config
        self.seq_kv_lens: List[int] = []
        self.reqid_mapping_blocktables: Dict[str, Dict[int, List[int]]] = {}

    def _add_seq_group(self, inter_data, chunked_prefill_enabled: bool) -> None:
        is_prompt = inter_data.is_prompt
        block_tables = inter_data.block_tables
        self.reqid_mapping_blocktables[inter_data.request_id] = block_tables
        for (seq_id, token_len, seq_len, curr_seq_len, query_len, context_len) in zip(
                inter_data.seq_ids, [len(t) for t in inter_data.input_tokens],
                inter_data.orig_seq_lens, inter_data.seq_lens, inter_data.query_lens,
                inter_data.context_lens):
            self.context_lens.append(context_len)
            if is_prompt:
                self.num_prefills += 1
                self.num_prefill_tokens += token_len
                self.prefill_seq_lens.append(seq_len)
            else:
                assert query_len == 1, "seq_len: {}, context_len: {}, query_len: {}".format(
                    seq_len, context_len, query_len)
                self.num_decode_tokens += query_len
                self.curr_seq_lens.append(curr_seq_len)
            block_table: List[int] = []
            if getattr(inter_data, "prefix_cache_hit", False):
                block_table = block_tables[seq_id]
            elif (chunked_prefill_enabled or not is_prompt) and block_tables is not None:
                block_table = block_tables[seq_id]
            self.block_tables.append(block_table)
            is_profile_run = block_tables is None or all(not v for v in block_tables.values())
            self._compute_slot_mapping(is_profile_run, is_prompt, seq_id, query_len, context_len,
                                       block_tables)
        if getattr(inter_data, "seq_kv_lens", None) is not None:
            self.seq_kv_lens.extend(inter_data.seq_kv_lens.values())

    def _compute_slot_mapping(self, is_profile_run: bool, is_prompt: bool, seq_id: int,
                              query_len: int, context_len: int, block_tables) -> None:
        """torch_cuda_sdpa.py:565-598 (eviction on) / backends/utils.py compute_slot_mapping
        (eviction off, no sliding window)."""
        assert self.sliding_window is None, \
            "Sliding window is not supported when paged_evict is enabled"
        bs = self.block_size
        if is_profile_run:
            self.slot_mapping.append(np.full(query_len, PAD_SLOT_ID, dtype=np.int64))
            return
        table = block_tables[seq_id]
        if self.paged_evict_config is None:
            pos = np.arange(context_len, context_len + query_len, dtype=np.int64)
            t = np.asarray(table, dtype=np.int64)
            self.slot_mapping.append(t[pos // bs] * bs + pos % bs)
        elif is_prompt:
            self.slot_mapping.append(prefill_slots_np(table, query_len, bs))
        else:
            assert query_len == 1, "decode request should have query_len=1 when paged_evict is enabled"
            self.slot_mapping.append(
                np.asarray([table[-1] * bs + (context_len % bs)], dtype=np.int64))

    def build(self, seq_lens: List[int], query_lens: List[int], cuda_graph_pad_size: int,
              batch_size: int) -> TorchCUDASDPAMetadata:
        for inter_data in self.input_builder.inter_data_list:
            self._add_seq_group(inter_data, self.input_builder.chunked_prefill_enabled)
        if cuda_graph_pad_size != -1:
            raise NotImplementedError("CUDA-graph padded batches are not supported with paged "
                                      "eviction (every documented run is --enforce-eager)")
        device = self.runner.device
        pin = bool(getattr(self.runner, "pin_memory", False))
        max_query_len = max(query_lens)
        assert max_query_len > 0, "query_lens: {}".format(query_lens)
        width = max((len(t) for t in self.block_tables), default=0)
        bt = np.zeros((len(self.block_tables), width), dtype=np.int32)     # make_tensor_with_pad
        for i, t in enumerate(self.block_tables):
            bt[i, :len(t)] = t
        slot = (np.concatenate(self.slot_mapping) if self.slot_mapping
                else np.empty(0, dtype=np.int64))
        i32 = lambda x: _h2d(np.asarray(x, dtype=np.int32), device, pin)
        return TorchCUDASDPAMetadata(
            num_prefills=self.num_prefills, slot_mapping=_h2d(slot, device, pin),
            multi_modal_placeholder_index_maps={}, num_prefill_tokens=self.num_prefill_tokens,
            num_decode_tokens=self.num_decode_tokens, seq_lens=seq_lens,
            seq_lens_tensor=i32(seq_lens), max_query_len=max_query_len,
            max_prefill_seq_len=max(self.prefill_seq_lens, default=0),
            max_decode_seq_len=max(self.curr_seq_lens, default=0),
            query_start_loc=i32(list(accumulate(query_lens, initial=0))),
            seq_start_loc=i32(list(accumulate(seq_lens, initial=0))),
            context_lens_tensor=i32(self.context_lens), block_tables=_h2d(bt, device, pin),
            use_cuda_graph=False, seq_kv_lens=self.seq_kv_lens,
            seq_kv_lens_tensor=i32(self.seq_kv_lens),
            reqid_mapping_blocktables=self.reqid_mapping_blocktables)


# ==================================================================== per-device state
def _flat_rows(reqid_mapping_blocktables) -> List[Tuple[str, int]]:
    """Batch row -> (request_id, seq_id), the order the builder appended them in
    (= BlockL2NormManager.flat_sequence_map, block_l2norm_manager.py:56-62)."""
    return [(rid, sid) for rid, tabs in reqid_mapping_blocktables.items() for sid in tabs]


# ============================================================================== impl
class TorchCUDASDPABackendImpl(AttentionImpl[TorchCUDASDPAMetadata]):
    """torch_cuda_sdpa.py:728-1586.

    Extra keyword arguments (all optional; the reference passes cache_config / prefix,
    vllm/attention/layer.py:102-105):
      tp_rank, tp_size   tensor-parallel position; num_kv_heads is the LOCAL head count
                         (vllm/model_executor/models/llama.py:117-130)
      score_exchange     parallel.ScoreExchange: allocates the per-layer score pools so that
                         every rank can store into every rank's pool (None = single GPU)
      rng                source of the host-side draws of the `random` / `inverse_key_l2` decode
                         methods: an object with randint(a, b).  Default = Python's process-wide
                         `random` module, re-seeded with 1234 by every layer's constructor exactly
                         like HEAD (torch_cuda_sdpa.py:796), so a model built from these impls draws
                         the same sequence as the fork: layer order, rows in row order.
    """

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", blocksparse_params: Optional[Dict[str, Any]] = None,
                 logits_soft_cap: Optional[float] = None, **kwargs) -> None:
        if blocksparse_params is not None:
            raise ValueError("Torch SPDA does not support block-sparse attention.")
        if alibi_slopes is not None or sliding_window is not None:
            raise NotImplementedError("alibi / sliding window are not supported by the "
                                      "paged-eviction backend (torch_cuda_sdpa.py:567)")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = None
        self.sliding_window = None
        self.kv_cache_dtype = kv_cache_dtype
        assert self.num_heads % self.num_kv_heads == 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        supported = PagedAttention.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. "
                             f"Supported head sizes are: {supported}.")
        if kv_cache_dtype != "auto":
            raise NotImplementedError("Torch SDPA backend does not support FP8 KV cache. "
                                      "Please use xFormers backend instead.")
        self.cache_config: Optional[CacheConfig] = kwargs.get("cache_config", None)
        self.enabled_paged_eviction = False
        self.paged_evict_config: Optional[PagedEvictConfig] = None
        if self.cache_config is not None and self.cache_config.paged_evict_config is not None:
            self.enabled_paged_eviction = True
            self.paged_evict_config = self.cache_config.paged_evict_config
        self.prefix: str = kwargs.get("prefix", "")
        self.layer_idx = _extract_layer_index(self.prefix)
        self.tp_rank = int(kwargs.get("tp_rank", 0))
        self.tp_size = int(kwargs.get("tp_size", 1))
        self.score_exchange = kwargs.get("score_exchange", None)
        if self.tp_size > 1 and self._value_l2 and self.score_exchange is None:
            raise ValueError("value_l2 eviction under tensor parallelism needs a score_exchange "
                             "(every rank must see every head's block scores)")
        # HEAD/v1 compute the compacted table on streamingLLM trigger steps but hand the
        # un-compacted one to the kernel (is_pruned_decode vs is_prune_decode,
        # torch_cuda_sdpa.py:903,910 vs 858,1046).  Default = the intended behaviour;
        # the flag reproduces HEAD for the golden traces.
        self.head_compat_uncompacted = bool(kwargs.get(
            "head_compat_uncompacted", os.environ.get("PC_HEAD_COMPAT_UNCOMPACTED", "0") == "1"))
        # f3: True = the reference's `middle[mask].view(-1, H, D)` re-pack, which mixes source heads
        # inside a column; False keeps each head's surviving tokens in its own column
        self.head_mixing_repack = bool(kwargs.get(
            "head_mixing_repack", os.environ.get("PC_HEAD_MIXING_REPACK", "1") == "1"))
        self.rng = kwargs.get("rng", None)
        if self.rng is None and self.enabled_paged_eviction and \
                self.paged_evict_config.decode_evict_method in ("random", "inverse_key_l2"):
            random.seed(1234)              # torch_cuda_sdpa.py:796
            self.rng = random
        self.reqids_mapping_rmv_blockIdx: Dict[int, Tuple[str, int]] = {}
        self._pending_victims = None      # (device int32 [B], rows) of the last value_l2 step
        self._score_pool: Optional[torch.Tensor] = None   # fp32 [num_blocks, Hkv * tp_size]
        self._score = None                                 # ops.PcScorePool over it
        self._c_policy = None
        self._cache_ptrs = None           # (kv_cache data_ptr, key ptr, value ptr, views)

    # ------------------------------------------------------------------ state helpers
    @property
    def _value_l2(self) -> bool:
        return self.enabled_paged_eviction and self.paged_evict_config.decode_evict_method == "value_l2"

    @property
    def _decode_method(self) -> Optional[str]:
        return self.paged_evict_config.decode_evict_method if self.enabled_paged_eviction else None

    def _policy_struct(self):
        if self._c_policy is None:
            self._c_policy = ops.make_policy(self.paged_evict_config, self.cache_config.block_size)
        return self._c_policy

    def _scores(self, device, num_blocks: int):
        """Score pool of this layer (one fp32 row per physical block, every TP rank's
        heads), allocated on first use next to the KV pool it shadows."""
        t = self._score_pool
        if t is None or t.device != device or t.shape[0] < num_blocks:
            heads_total = self.num_kv_heads * self.tp_size
            if self.score_exchange is not None:
                t, peers = self.score_exchange.alloc_pool(num_blocks, heads_total, device)
                self._score = ops.make_score_pool(t, self.tp_rank * self.num_kv_heads, peers, self.tp_rank)
            else:
                t = torch.zeros(num_blocks, heads_total, dtype=torch.float32, device=device)
                self._score = ops.make_score_pool(t, 0)
            self._score_pool = t
        return self._score

    def _caches(self, kv_cache: torch.Tensor):
        """key/value cache views + raw pointers, cached per kv_cache tensor."""
        c = self._cache_ptrs
        if c is None or c[0] != kv_cache.data_ptr() or c[4] != kv_cache.shape[1]:
            kc, vc = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads, self.head_size)
            ops._cache_dims(kc, vc)       # layout check once
            c = (kv_cache.data_ptr(), kc, vc, (kc.data_ptr(), vc.data_ptr(), vc.stride(0)), kv_cache.shape[1])
            self._cache_ptrs = c
        return c[1], c[2], c[3]

    def _prepare_step(self, md: TorchCUDASDPAMetadata, device, bs: int) -> Dict[str, Any]:
        """Host work done ONCE per engine step (by the first layer that sees `md`): policy
        arithmetic and one packed H2D of the int32 side tables.  The reference repeats its
        equivalent loops in every layer."""
        if md._pc_step is not None:
            return md._pc_step
        cfg = self.paged_evict_config
        st: Dict[str, Any] = {}
        rows = _flat_rows(md.reqid_mapping_blocktables) if md.reqid_mapping_blocktables else []
        st["rows"] = rows
        pack: List[np.ndarray] = []
        names: List[str] = []
        if md.num_prefills > 0:
            seq_lens = md.seq_lens[:md.num_prefills]
            m5, m6, m6_full, toff = [], [], [], [0]
            # f3: token-level prompt methods prune into a packed tensor first (rows dst ..), the
            # cache write and the attention then read that tensor with an identity gather
            tok = (cfg is not None and cfg.prompt_evict_method != "streamingLLM"
                   and any(prompt_prune_layout(L, cfg, bs)[1] > 0 for L in seq_lens))
            tp6, moff, qoff = [], [0], [0]
            src = dst = 0
            for L in seq_lens:
                kept, pruned, first = prompt_prune_layout(L, cfg, bs)
                if tok:
                    mid = (L - first - (bs + L % bs)) if pruned > 0 else 0
                    tp6.append((src, L, first, mid, pruned, dst))
                    moff.append(moff[-1] + mid)
                    qoff.append(qoff[-1] + L)
                    m5.append((dst, kept, kept, 0, dst))
                    m6.append((src, L, dst, kept, kept, 0))
                else:
                    m5.append((src, kept, first, pruned, dst))
                    m6.append((src, L, src, kept, first, pruned))
                m6_full.append((src, L, src, L, L, 0))
                toff.append(toff[-1] + -(-kept // bs))
                src += L
                dst += kept
            st["total_tiles"] = toff[-1]
            st["kept_tokens"] = dst
            use = m6_full if (cfg is None or cfg.disable_evict_prefill) else m6
            st["m6_host"] = np.ascontiguousarray(np.asarray(use, np.int32).reshape(-1, 6))
            pack = [np.asarray(m5, np.int32).reshape(-1), np.asarray(toff, np.int32),
                    st["m6_host"].reshape(-1)]
            names = ["m5", "toff", "m6"]
            st["token_prune"] = tok
            if tok:
                st["total_mid"], st["total_tokens"] = moff[-1], qoff[-1]
                pack += [np.asarray(tp6, np.int32).reshape(-1), np.asarray(moff, np.int32),
                         np.asarray(qoff, np.int32)]
                names += ["tp6", "moff", "qoff"]
        else:
            kv = md.seq_kv_lens if md.seq_kv_lens else md.seq_lens
            if md.reqid_mapping_blocktables:
                # ground truth: entries each row owns in the host block table
                nblk = [len(t) for tabs in md.reqid_mapping_blocktables.values() for t in tabs.values()]
            else:
                nblk = [-(-k // bs) for k in kv]
            trig = [decode_trigger(s, n * bs, cfg, bs) for s, n in zip(md.seq_lens, nblk)]
            # on trigger rows kv lens arrive already reduced by block_size (model_runner.py:540-543)
            st["trigger_rows"] = [i for i, t in enumerate(trig) if t]
            st["max_kv_len"] = max(kv, default=0)
            st["nblk_host"] = nblk
            if cfg is not None and cfg.decode_evict_method in ("streamingLLM-1", "inverse_key_l2"):
                st["tables_host"] = [t for tabs in md.reqid_mapping_blocktables.values() for t in tabs.values()]
                if cfg.decode_evict_method == "streamingLLM-1":      # W2: the same slots in every layer
                    zs = [zero_slot(None, s, t, cfg, bs) for s, t in zip(md.seq_lens, st["tables_host"])]
                    st["zero_slots"] = (_h2d(np.asarray(zs, np.int64), device, True)
                                        if any(z >= 0 for z in zs) else None)
            pack = [np.asarray(nblk, np.int32)]
            names = ["nblk"]
        flat = np.concatenate(pack) if pack else np.empty(0, np.int32)
        dev = _h2d(flat, device, True)
        o = 0
        for n, a in zip(names, pack):
            st[n] = dev[o:o + a.size]
            o += a.size
        if "m5" in st:
            st["m5"] = st["m5"].view(-1, 5)
            st["m6"] = st["m6"].view(-1, 6)
        if "tp6" in st:
            st["tp6"] = st["tp6"].view(-1, 6)
        md._pc_step = st
        return st

    # ------------------------------------------------------------------------ forward
    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: torch.Tensor, attn_metadata: TorchCUDASDPAMetadata, k_scale: float = 1.0,
                v_scale: float = 1.0, attn_type: str = AttentionType.DECODER,
                output: Optional[torch.Tensor] = None) -> torch.Tensor:
        """query [T, Hq*D], key/value [T, Hkv*D], kv_cache [2, nb, bs*Hkv*D] (numel 0 on
        the profile run) -> [T, Hq*D]   (torch_cuda_sdpa.py:812-1064)."""
        assert k_scale == 1.0 and v_scale == 1.0
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("Paged eviction is only enabled for decoder attention")
        md = attn_metadata
        self.reqids_mapping_rmv_blockIdx = {}      # reset the mapping for each forward pass
        self._pending_victims = None
        q = query.view(-1, self.num_heads, self.head_size)
        k = key.view(-1, self.num_kv_heads, self.head_size)
        v = value.view(-1, self.num_kv_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size) if output is not None else torch.empty_like(q)
        is_profile_run = kv_cache.numel() == 0
        assert md.num_prefills == 0 or md.num_decode_tokens == 0, \
            "paged eviction does not support chunked prefill"
        dev = query.device
        if self.cache_config is not None:
            bs = self.cache_config.block_size
        else:
            bs = 16 if is_profile_run else kv_cache.shape[2] // (self.num_kv_heads * self.head_size)
        st = self._prepare_step(md, dev, bs)
        if md.num_prefills > 0:
            self._forward_prefill(q, k, v, out, md, st, None if is_profile_run else kv_cache)
        elif md.num_decode_tokens > 0:
            self._forward_decode(q, k, v, out, md, st, kv_cache, bs)
        return out.view(-1, self.num_heads * self.head_size)

    def _forward_prefill(self, q, k, v, out, md, st, kv_cache) -> None:
        kc_src, vc_src = k, v
        if st.get("token_prune"):
            # f3 (kvcache_prunner.py:220-279): per head the lowest-scoring middle tokens go, the
            # survivors are re-packed the reference's way; T rows so q / k / v share one TMA bound
            kc_src, vc_src = torch.empty_like(k), torch.empty_like(v)
            ops.prompt_token_prune(k, v, kc_src, vc_src, st["tp6"], st["moff"], st["qoff"], st["total_tokens"],
                                   st["total_mid"], self.paged_evict_config.prompt_evict_method,
                                   head_mixing_repack=self.head_mixing_repack)
        if kv_cache is not None:
            key_cache, value_cache, _ = self._caches(kv_cache)
            score = self._scores(q.device, kv_cache.shape[1]) if self._value_l2 else None
            if self.enabled_paged_eviction:
                # R1 + W1 + S3 in one pass; slot_mapping has exactly `kept` entries per request
                ops.prune_prompt_and_cache(kc_src, vc_src, key_cache, value_cache, md.slot_mapping, st["m5"],
                                           st["toff"], st["total_tiles"], score)
            else:
                ops.reshape_and_cache(k, v, key_cache, value_cache, md.slot_mapping)
        # A1: top-left aligned causal mask over the kept rows (or all rows when
        # disable_evict_prefill / eviction off) -- torch_cuda_sdpa.py:1069-1127
        if self.enabled_paged_eviction and self.paged_evict_config.disable_evict_prefill:
            kc_src, vc_src = k, v
        ops.prefill_attn_topleft_causal(out, q, kc_src, vc_src, st["m6"], md.max_query_len, self.scale,
                                        req_meta_host=st["m6_host"])

    def _decode_step_struct(self, q, k, v, out, md, st, blk_stride: int, bs: int):
        """struct pc_decode_step shared by every layer of this engine step."""
        ds = st.get("c_step")
        if ds is not None and (ds.q_stride, ds.key_stride, ds.value_stride, ds.out_stride,
                               ds.kv_block_stride) == (q.stride(0), k.stride(0), v.stride(0),
                                                       out.stride(0), blk_stride):
            return ds
        for t in (q, k, v, out):
            if t.stride()[1:] != (self.head_size, 1) or t.dtype != torch.bfloat16:
                raise RuntimeError("query/key/value/out rows must be dense bf16 [heads, head_size]")
        tables = md.block_tables
        kv_lens = md.seq_kv_lens_tensor if self.enabled_paged_eviction else md.seq_lens_tensor
        trig = st.get("trigger_rows", [])
        compact = None
        if trig:
            compact = st.get("compact")
            if compact is None:
                compact = torch.empty_like(tables)
                st["compact"] = compact
        attend_unc = bool(trig) and self.head_compat_uncompacted and \
            self._decode_method in ("streamingLLM", "streamingLLM-1")
        ds = ops.make_decode_step(
            md.slot_mapping, tables, kv_lens, md.seq_lens_tensor, q.shape[0], self.num_heads,
            self.num_kv_heads, self.head_size, bs, st["max_kv_len"], self.scale, q.stride(0),
            k.stride(0), v.stride(0), out.stride(0), blk_stride,
            policy=self._policy_struct() if self.enabled_paged_eviction else None,
            n_blocks=st["nblk"], compact_tables=compact, any_trigger=bool(trig),
            attend_uncompacted=attend_unc)
        st["c_step"] = ds
        return ds

    def _forward_decode(self, q, k, v, out, md, st, kv_cache, bs: int) -> None:
        """W1 -> [P3 + S4 + C1 on trigger steps] -> A2 + S2, one native call."""
        _, _, (kptr, vptr, blk_stride) = self._caches(kv_cache)
        ds = self._decode_step_struct(q, k, v, out, md, st, blk_stride, bs)
        score = self._scores(q.device, kv_cache.shape[1]) if self._value_l2 else None
        trig = st.get("trigger_rows", [])
        method = self._decode_method
        # W2: this step's zeroed slots land before the attention reads the pool
        zs = st.get("zero_slots")
        if method == "inverse_key_l2":        # two draws per over-budget row, every layer its own
            cfg = self.paged_evict_config
            host = [zero_slot(self.rng, s, t, cfg, bs) for s, t in zip(md.seq_lens, st["tables_host"])]
            zs = _h2d(np.asarray(host, np.int64), q.device, True) if any(z >= 0 for z in host) else None
        if zs is not None:
            kc, vc, _ = self._caches(kv_cache)
            ops.zero_cache_slots(kc, vc, zs)
        victims = None
        host_victims = None
        if trig:
            if method == "random":            # HEAD as shipped: one draw per triggered row, row order
                host_victims = {r: random_victim(self.rng, st["nblk_host"][r], self.paged_evict_config)
                                for r in trig}
                hv = np.full(q.shape[0], -1, np.int32)
                for r, i in host_victims.items():
                    if not 1 <= i < st["nblk_host"][r]:
                        raise RuntimeError(f"random victim {i} outside the row's {st['nblk_host'][r]} blocks")
                    hv[r] = i
                victims = _h2d(hv, q.device, True)
            else:
                victims = torch.empty(q.shape[0], dtype=torch.int32, device=q.device)
        ops.decode_layer(ds, out, q, k, v, kptr, vptr, score, victims)
        if trig:
            rows = st["rows"]
            if self._value_l2:
                self._pending_victims = (victims, trig, rows)
            elif host_victims is not None:
                self.reqids_mapping_rmv_blockIdx = {r: (rows[r][0], i) for r, i in host_victims.items()}
            else:                          # streamingLLM(-1): the victim is always table index 1
                self.reqids_mapping_rmv_blockIdx = {r: (rows[r][0], 1) for r in trig}

    def get_rmv_block_idx(self) -> Dict[int, Tuple[str, int]]:
        """{batch_row: (request_id, table_index)} of THIS step (torch_cuda_sdpa.py:1066-1067).
        Only the caller (layer 0, model_runner.py:1797-1802) pays the one int32[B] D2H."""
        if self._pending_victims is not None:
            victims, trig, rows = self._pending_victims
            host = victims.cpu().tolist()
            self.reqids_mapping_rmv_blockIdx = {r: (rows[r][0], host[r]) for r in trig if host[r] >= 0}
            self._pending_victims = None
        return self.reqids_mapping_rmv_blockIdx


def _extract_layer_index(prefix: str) -> int:
    """vllm/utils.py extract_layer_index: the single integer component of a dotted name."""
    ints = [int(p) for p in prefix.split(".") if p.isdigit()]
    return ints[0] if len(ints) == 1 else -1
