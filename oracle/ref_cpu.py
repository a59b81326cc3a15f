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


class RefCpuOps:
    """reshape_and_cache / paged_attention with the call shapes of oracle/paged_ops.py,
    executed by the reference's own compiled CPU kernels (block_size 16 only).  The
    v1/v2 choice follows PagedAttention.forward_decode (paged_attn.py:117-127)."""

    @staticmethod
    def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping):
        reshape_and_cache(key.contiguous(), value.contiguous(), key_cache, value_cache,
                          slot_mapping.flatten().to(torch.long))

    @staticmethod
    def paged_attention(query, key_cache, value_cache, block_tables, kv_lens, scale,
                        out_dtype=None):
        B, Hq, D = query.shape
        bs = value_cache.shape[3]
        width = max(1, max((len(t) for t in block_tables), default=1))
        bt = torch.zeros(B, width, dtype=torch.int32)
        for r, t in enumerate(block_tables):
            if len(t):
                bt[r, :len(t)] = torch.as_tensor(list(t), dtype=torch.int32)
        lens = torch.as_tensor(list(kv_lens), dtype=torch.int32)
        max_len = int(max(kv_lens)) if len(kv_lens) else 0
        out = torch.empty_like(query)
        partitions = (max_len + 511) // 512
        q = query.contiguous()
        if max_len <= 8192 and (partitions == 1 or B * Hq > 512):
            paged_attention_v1(out, q, key_cache, value_cache, value_cache.shape[1], scale, bt,
                               lens, bs, max_len)
        else:
            paged_attention_v2(out, q, key_cache, value_cache, value_cache.shape[1], scale, bt,
                               lens, bs, max_len)
        return out
