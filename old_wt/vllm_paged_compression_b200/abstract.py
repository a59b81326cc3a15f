"""The attention plugin surface the backend implements, restated so the package imports
without the fork installed (vllm/attention/backends/abstract.py:17-258 and
vllm/attention/backends/utils.py PAD_SLOT_ID / CommonAttentionState).  Method names,
argument order and defaults are the fork's; INTEGRATION.md shows how the classes in
backend.py register against a live checkout (they only need these names).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Any, Dict, Generic, List, Optional, Tuple, Type, TypeVar

import torch

PAD_SLOT_ID = -1          # vllm/attention/backends/utils.py:25


class AttentionType:      # abstract.py:17-29
    DECODER = "decoder"
    ENCODER = "encoder"
    ENCODER_ONLY = "encoder_only"
    ENCODER_DECODER = "encoder_decoder"


class AttentionBackend(ABC):   # abstract.py:32-100
    @staticmethod
    @abstractmethod
    def get_name() -> str: ...

    @staticmethod
    @abstractmethod
    def get_impl_cls() -> Type["AttentionImpl"]: ...

    @staticmethod
    @abstractmethod
    def get_metadata_cls() -> Type["AttentionMetadata"]: ...

    @staticmethod
    @abstractmethod
    def get_state_cls(): ...

    @classmethod
    def make_metadata(cls, *args, **kwargs) -> "AttentionMetadata":
        return cls.get_metadata_cls()(*args, **kwargs)

    @staticmethod
    @abstractmethod
    def get_builder_cls() -> Type["AttentionMetadataBuilder"]: ...

    @classmethod
    def make_metadata_builder(cls, *args, **kwargs) -> "AttentionMetadataBuilder":
        return cls.get_builder_cls()(*args, **kwargs)

    @staticmethod
    @abstractmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]: ...

    @staticmethod
    @abstractmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None: ...

    @staticmethod
    @abstractmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None: ...


@dataclass
class AttentionMetadata:       # abstract.py:103-155
    num_prefills: int
    num_prefill_tokens: int
    num_decode_tokens: int
    slot_mapping: torch.Tensor
    multi_modal_placeholder_index_maps: Optional[Dict[str, Any]]

    @property
    def prefill_metadata(self) -> Optional["AttentionMetadata"]:
        raise NotImplementedError

    @property
    def decode_metadata(self) -> Optional["AttentionMetadata"]:
        raise NotImplementedError


T = TypeVar("T", bound=AttentionMetadata)


class AttentionMetadataBuilder(ABC, Generic[T]):   # abstract.py:208-219
    @abstractmethod
    def __init__(self, input_builder) -> None: ...

    @abstractmethod
    def build(self, seq_lens: List[int], query_lens: List[int], cuda_graph_pad_size: int,
              batch_size: int) -> T: ...


class AttentionImpl(ABC, Generic[T]):              # abstract.py:222-258
    @abstractmethod
    def __init__(self, num_heads: int, head_size: int, scale: float,
                 num_kv_heads: Optional[int] = None, alibi_slopes: Optional[List[float]] = None,
                 sliding_window: Optional[int] = None, kv_cache_dtype: str = "auto",
                 blocksparse_params: Optional[Dict[str, Any]] = None,
                 logits_soft_cap: Optional[float] = None, **kwargs) -> None: ...

    @abstractmethod
    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: torch.Tensor, attn_metadata: T, k_scale: float = 1.0, v_scale: float = 1.0,
                attn_type: str = AttentionType.DECODER,
                output: Optional[torch.Tensor] = None) -> torch.Tensor: ...

    @abstractmethod
    def get_rmv_block_idx(self) -> Dict[int, Tuple[str, int]]:
        """{batch_row: (request_id, block_table_index)} evicted this step (abstract.py:255-258)."""


class CommonAttentionState:
    """Stand-in for vllm/attention/backends/utils.py CommonAttentionState: CUDA-graph
    capture state.  Eviction changes table contents every trigger step and every
    documented run is --enforce-eager, so only the eager surface is provided."""

    def __init__(self, runner=None):
        self.runner = runner
        self._is_graph_capturing = False

    def begin_forward(self, model_input) -> None:
        return
