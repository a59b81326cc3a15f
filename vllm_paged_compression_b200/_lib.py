"""ctypes binding of libpagedevict_b200.so -- the only door to the device code.

Mirrors the role of vllm/_custom_ops.py's `torch.ops._C` handle.  Fails loudly when the
library is missing or stale: there is deliberately no fallback path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PC_LIB_PATH") or os.path.join(_HERE, "libpagedevict_b200.so")
ABI_VERSION = 6
MAX_PEERS = 8


class PcEvictPolicy(C.Structure):
    """struct pc_evict_policy (include/pagedevict_b200.h)."""
    _fields_ = [("prune_type", C.c_int32), ("block_size", C.c_int32),
                ("cache_budget", C.c_int32), ("evict_freq", C.c_int32),
                ("initial_blocks", C.c_int32), ("num_block_merge", C.c_int32),
                ("decode_method", C.c_int32), ("topk_blocks", C.c_int32)]


class PcScorePool(C.Structure):
    """struct pc_score_pool."""
    _fields_ = [("pools", C.c_void_p * MAX_PEERS), ("n_pools", C.c_int32), ("local", C.c_int32),
                ("heads_total", C.c_int32), ("head_offset", C.c_int32)]


class PcDecodeStep(C.Structure):
    """struct pc_decode_step."""
    _fields_ = [("slot_mapping", C.c_void_p), ("block_tables", C.c_void_p), ("table_stride", C.c_int64),
                ("kv_lens", C.c_void_p), ("seq_lens", C.c_void_p), ("n_blocks", C.c_void_p),
                ("compact_tables", C.c_void_p), ("compact_stride", C.c_int64),
                ("batch", C.c_int32), ("num_heads", C.c_int32), ("num_kv_heads", C.c_int32),
                ("head_size", C.c_int32), ("block_size", C.c_int32), ("max_kv_len", C.c_int32),
                ("scale", C.c_float), ("any_trigger", C.c_int32), ("attend_uncompacted", C.c_int32),
                ("q_stride", C.c_int64), ("key_stride", C.c_int64), ("value_stride", C.c_int64),
                ("out_stride", C.c_int64), ("kv_block_stride", C.c_int64),
                ("policy", PcEvictPolicy), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
                ("dtype", C.c_int32), ("stream", C.c_void_p),
                ("fuse_cache_write", C.c_int32), ("plan", C.c_void_p),
                ("pdl", C.c_int32)]


_P, _I32, _I64, _F, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t
_SP = C.POINTER(PcScorePool)

# name -> (restype, argtypes); must list every symbol declared in include/pagedevict_b200.h
PROTOTYPES = {
    "pc_abi_version": (C.c_int, []),
    "pc_sizeof_struct": (_SZ, [_I32]),
    "pc_last_error": (C.c_char_p, []),
    "pc_device_sm_count": (C.c_int, []),
    "pc_reshape_and_cache": (C.c_int, [_P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I64, _I64, _I64,
                                       _I32, _P]),
    "pc_prune_prompt_and_cache": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32,
                                            _I64, _I64, _I64, _SP, _I32, _P]),
    "pc_evict_select_compact": (C.c_int, [_P, _I64, _P, _P, _P, _I32, C.POINTER(PcEvictPolicy), _SP,
                                          _I32, _P, _I64, _P, _P, _I32, _P]),
    "pc_paged_decode_workspace_bytes": (_SZ, [_I32, _I32, _I32, _I32, _I32, _I32]),
    "pc_paged_decode_plan": (C.c_int, [_I32, _I32, _I32, _I32, _I32, _I32, _I32, C.POINTER(C.c_int32)]),
    "pc_paged_decode_attn_score": (C.c_int, [_P, _P, _P, _P, _P, _I64, _P, _I32, _I32, _I32, _I32, _I32,
                                             _I32, _F, _I64, _I64, _I64, _P, _SP, _P, _SZ,
                                             _I32, _P]),
    # no caller workspace: the library's file-scope device scratch (PC_DECODE_STATIC_SCRATCH_BYTES)
    "pc_paged_decode_attn_score_static": (C.c_int, [_P, _P, _P, _P, _P, _I64, _P, _I32, _I32, _I32, _I32, _I32,
                                                    _I32, _F, _I64, _I64, _I64, _P, _SP, _I32, _P]),
    "pc_decode_layer": (C.c_int, [C.POINTER(PcDecodeStep), _P, _P, _P, _P, _P, _P, _SP, _P, _P, _I64]),
    "pc_evict_select_compact_layers": (C.c_int, [_P, _I64, _P, _P, _I32, _I32, C.POINTER(PcEvictPolicy), _P, _I32,
                                                 _P, _I64, _I64, _P, _P]),
    "pc_prefill_attn_topleft_causal": (C.c_int, [_P, _P, _P, _P, _P, _I32, _I32, _I64, _I32, _I32, _I32, _F,
                                                 _I64, _I64, _I64, _I64, _P, _I32, _P]),
    "pc_prompt_token_prune_workspace_bytes": (_SZ, [_I64, _I32]),
    "pc_prompt_token_prune": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _I32, _I64, _I64, _I32, _I32, _I32, _I32,
                                        _I64, _I64, _I64, _I64, _P, _SZ, _I32, _P]),
    "pc_zero_cache_slots": (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _I32, _I64, _P]),
    "pc_zero_cache_slots_layers": (C.c_int, [_P, _P, _P, _I64, _I32, _I32, _I32, _I32, _I32, _I64, _P]),
    "pc_copy_blocks": (C.c_int, [_P, _P, _P, _I32, _I64, _P]),
    "pc_swap_blocks": (C.c_int, [_P, _P, _P, _I32, _I64, _P]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C vllm_paged_compression_b200/csrc`).  This package has no CPU / PyTorch "
            "fallback for the paged-eviction path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is missing: stale build
        fn.restype = res
        fn.argtypes = args
    if lib.pc_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libpagedevict_b200.so ABI {lib.pc_abi_version()} != expected {ABI_VERSION}")
    for which, cls in enumerate((PcEvictPolicy, PcScorePool, PcDecodeStep)):
        if lib.pc_sizeof_struct(which) != C.sizeof(cls):
            raise RuntimeError(f"struct layout mismatch for {cls.__name__}: library "
                               f"{lib.pc_sizeof_struct(which)} bytes, binding {C.sizeof(cls)}")
    _lib = lib
    return lib


def check(rc: int) -> None:
    """Native error convention -> RuntimeError, like TORCH_CHECK in the reference ops."""
    if rc != 0:
        raise RuntimeError(load().pc_last_error().decode())
