"""`torch.ops.pagedevict_b200.*`: the thin torch C++ extension over the C ABI
(csrc/torch_shim.cpp -> libpagedevict_b200_torch.so).

The fork binds this path through TORCH_LIBRARY ops (`torch.ops._C.paged_attention_v1/_v2`,
`torch.ops._C_cache_ops.reshape_and_cache`, csrc/torch_bindings.cpp:27-51,444-450, called from
vllm/_custom_ops.py:70-130,984-996).  `load()` registers the same schemas under the namespace
`pagedevict_b200`; `alias_vllm_ops()` additionally defines them under the fork's own names, so an
UNPATCHED vllm/_custom_ops.py lands here -- call it instead of importing the fork's `vllm._C`.
The ctypes wrappers in ops.py remain the complete binding (fused / batched entry points)."""
from __future__ import annotations

import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PC_TORCH_LIB_PATH") or os.path.join(_HERE, "libpagedevict_b200_torch.so")
_loaded = False


def load():
    """-> torch.ops.pagedevict_b200 (raises when the extension has not been built)."""
    global _loaded
    if not _loaded:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C vllm_paged_compression_b200/csrc torch`")
        torch.ops.load_library(LIB_PATH)
        from ._lib import ABI_VERSION
        got = int(torch.ops.pagedevict_b200.abi_version())
        if got != ABI_VERSION:
            raise RuntimeError(f"libpagedevict_b200_torch.so is linked against ABI {got}, expected {ABI_VERSION}")
        _loaded = True
    return torch.ops.pagedevict_b200


def alias_vllm_ops() -> None:
    """Define _C::paged_attention_v1/_v2 and _C_cache_ops::reshape_and_cache, implemented here."""
    load().alias_vllm_ops()
