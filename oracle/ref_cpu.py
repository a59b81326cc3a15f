"""Loader for oracle/_ref/libpc_ref_cpu.so = the reference's OWN CPU kernels
(csrc/cpu/attention.cpp, csrc/cpu/cache.cpp) compiled by oracle/Makefile.
TEST INFRASTRUCTURE ONLY: used to validate the restatement and as the
`cpu_baseline.kind == "reference"` timing arm.  Limits inherited from the reference:
block_size 16 only (attention.cpp:275,449-453), k_scale == v_scale == 1 (:467).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libpc_ref_cpu.so")
_loaded: Optional[bool] = None
_why = ""


def _host_has_avx512() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            flags = f.read()
        return all(k in flags for k in ("avx512f", "avx512bw", "avx512vl", "avx512dq"))
    except OSError:
        return False


def available() -> bool:
    global _loaded, _why
    if _loaded is None:
        if not os.path.exists(_SO):
            _loaded, _why = False, "oracle/_ref/libpc_ref_cpu.so not built"
        elif not _host_has_avx512():
            _loaded, _why = False, "host lacks AVX-512 (library built with -mavx512f/bw/vl/dq)"
        else:
            try:
                torch.ops.load_library(_SO)
                _loaded = True
            except OSError as e:  # pragma: no cover
                _loaded, _why = False, f"dlopen failed: {e}"
    return bool(_loaded)


def why_unavailable() -> str:
    available()
    return _why


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len):
    assert available(), why_unavailable()
    torch.ops.pc_ref_cpu.paged_attention_v1(
        out, query, key_cache, value_cache, num_kv_heads, float(scale), block_tables,
        seq_lens, block_size, max_seq_len, None, "auto", 1.0, 1.0, 0, 0, 0, 64, 0)


def paged_attention_v2(out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len, partition=512):
    assert available(), why_unavailable()
    B, Hq, D = query.shape
    P = (max_seq_len + partition - 1) // partition
    tmp = torch.empty(B, Hq, P, D, dtype=out.dtype)
    es = torch.empty(B, Hq, P, dtype=torch.float32)
    ml = torch.empty_like(es)
    torch.ops.pc_ref_cpu.paged_attention_v2(
        out, es, ml, tmp, query, key_cache, value_cache, num_kv_heads, float(scale),
        block_tables, seq_lens, block_size, max_seq_len, None, "auto", 1.0, 1.0, 0, 0, 0, 64, 0)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping):
    assert available(), why_unavailable()
    torch.ops.pc_ref_cpu.reshape_and_cache(key, value, key_cache, value_cache,
                                           slot_mapping, "auto", 1.0, 1.0)
