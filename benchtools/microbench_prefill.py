"""Prefill-attention (A1) microbench of the tcgen05/TMEM kernel.

flops = 4 * D * Hq * sum_req sum_i min(i + 1, L')  (top-left causal over the kept prompt,
SURVEY.md 8d).  CUDA events around `reps` back-to-back launches after warm-up; tensor-pipe
fraction is reported against MEASURED_PEAKS.json bf16_tflops (burst, kernel timed alone).
    python -m benchtools.microbench_prefill [--reqs 16 --len 1024 --evict-freq 2]
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os

import torch

from vllm_paged_compression_b200 import _lib, ops


def _peak_tflops():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["bf16_tflops"])
    return 2250.0


def prune_layout(L: int, bs: int, evict_freq: int, initial_blocks: int = 1, num_block_merge: int = 2):
    """Percentage-mode streamingLLM prompt prune (kvcache_prunner_old.py:141-200)."""
    if evict_freq <= 0 or L <= (initial_blocks + num_block_merge) * bs:
        return L, 0, L
    rem = L % bs
    first = initial_blocks * bs
    end_mid = L - num_block_merge * bs - rem
    pruned = bs * (int((end_mid - first) / evict_freq) // bs)
    return L - pruned, pruned, first


def _debug_knobs(lib, tile, persistent):
    """kv tile width / launch mode overrides exist only in the PC_DEBUG library
    (`make -C vllm_paged_compression_b200/csrc debug`, PC_LIB_PATH=...); the product library
    always runs its measured rule (64-wide tile, persistent for prompts <= 2048)."""
    if not hasattr(lib, "pc_debug_set_prefill_tc_tile"):
        if tile != 64 or persistent != -1:
            raise RuntimeError("tile / persistent overrides need the PC_DEBUG library")
        return
    for name, v in (("pc_debug_set_prefill_tc_tile", tile), ("pc_debug_set_prefill_tc_persistent", persistent)):
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = [ctypes.c_int32], None
        fn(v)


def run(device, reqs, L, Hq, Hkv, D, evict_freq, reps, variant=0, check=None, tile=64, persistent=-1, dq=-1):
    g = torch.Generator(device=device)
    g.manual_seed(0)
    T = reqs * L
    s = D ** -0.5
    q = ((torch.rand(T, Hq, D, device=device, generator=g) * 2 - 1)).to(torch.bfloat16)
    k = ((torch.rand(T, Hkv, D, device=device, generator=g) * 2 - 1)).to(torch.bfloat16)
    v = ((torch.rand(T, Hkv, D, device=device, generator=g) * 2 - 1)).to(torch.bfloat16)
    kept, pruned, first = prune_layout(L, 16, evict_freq)
    meta = torch.tensor([[r * L, L, r * L, kept, first, pruned] for r in range(reqs)],
                        dtype=torch.int32, device=device)
    out = torch.zeros(T, Hq, D, dtype=torch.bfloat16, device=device)
    lib = _lib.load()
    _debug_knobs(lib, tile, persistent)

    def launch():
        ops.prefill_attn_topleft_causal(out, q, k, v, meta, L, s)
    for _ in range(3):
        launch()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        launch()
    e1.record()
    torch.cuda.synchronize()
    _debug_knobs(lib, 64, -1)
    sec = e0.elapsed_time(e1) / 1e3 / reps
    vis = sum(min(i + 1, kept) for i in range(L))
    flops = 4.0 * D * Hq * reqs * vis
    mode = {-1: "auto_", 0: "oneshot_", 1: "persist_", 2: "persist_localorder_"}[persistent]
    # the library's own rule (prefill_attn_tc.cu: prefill_attn): dual-query-tile kernel for head size 128, an even
    # number of q heads per kv head and prompts <= 6144 tokens; PC_DEBUG knobs (--dq, tile) override it
    rule_dq = D == 128 and (Hq // Hkv) % 2 == 0 and L <= 6144 and tile == 64 and persistent == -1
    use_dq = rule_dq if dq < 0 else (dq == 1 and (Hq // Hkv) % 2 == 0)
    row = {"kernel": "prefill_attn_dq_kernel" if use_dq else f"prefill_attn_tc_{mode}kernel<TN={tile}>",
           "reqs": reqs, "q_len": L, "kv_kept": kept, "Hq": Hq, "Hkv": Hkv, "D": D,
           "us": sec * 1e6, "tflops": flops / sec / 1e12,
           "frac_of_measured_bf16_peak": flops / sec / 1e12 / _peak_tflops()}
    if check is not None:
        row["max_abs_diff_vs_other_variant"] = float((out.float() - check.float()).abs().max())
    return row, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reqs", type=int, default=16)
    ap.add_argument("--len", type=int, default=1024)
    ap.add_argument("--Hq", type=int, default=32)
    ap.add_argument("--Hkv", type=int, default=8)
    ap.add_argument("--D", type=int, default=128)
    ap.add_argument("--evict-freq", type=int, default=2)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--pp", type=int, default=-1, help="PC_DEBUG library: 1 = ping-pong kernel, 0 = off, -1 = rule")
    ap.add_argument("--modes", action="store_true", help="PC_DEBUG library: one-shot / persistent side by side")
    ap.add_argument("--xitem", type=int, default=-1, help="PC_DEBUG library: 0 = no S MMA issue across item boundaries")
    ap.add_argument("--dq", type=int, default=-1, help="PC_DEBUG library: 1 = the dual-query-tile kernel always, 0 = never, -1 = rule")
    a = ap.parse_args()
    if a.xitem >= 0:
        import ctypes
        fn = _lib.load().pc_debug_set_prefill_tc_xitem
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(a.xitem)
    if a.dq >= 0:
        import ctypes
        fn = _lib.load().pc_debug_set_prefill_dq
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(a.dq)
    if a.pp >= 0:
        import ctypes
        fn = _lib.load().pc_debug_set_prefill_pp
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(a.pp)
    dev = torch.device("cuda", 0)
    row, ref = run(dev, a.reqs, a.len, a.Hq, a.Hkv, a.D, a.evict_freq, a.reps, dq=a.dq)
    print(json.dumps(row), flush=True)
    if a.modes:
        for tile, pers in ((64, 0), (64, 1), (64, 2)):
            row, _ = run(dev, a.reqs, a.len, a.Hq, a.Hkv, a.D, a.evict_freq, a.reps, check=ref, tile=tile,
                         persistent=pers)
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
