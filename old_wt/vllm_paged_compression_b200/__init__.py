"""B200-native (sm_100a) PagedCompression hot path: a drop-in for the fork's TORCH_SDPA
paged-eviction backend (vllm/attention/backends/torch_cuda_sdpa.py of
krishnateja95/vllm-Paged-Compression).

Python here is host plumbing only (metadata, block bookkeeping, op wrappers); all
device work goes through the C ABI of libpagedevict_b200.so (include/pagedevict_b200.h).
There is no CPU / PyTorch fallback: importing `ops` without the built library raises.
"""
from .config import PagedEvictConfig, CacheConfig, add_cli_args, paged_evict_config_from_args  # noqa: F401

__version__ = "0.1.0"
