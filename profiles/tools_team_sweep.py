"""Calibration sweep of the decode planner (PC_DEBUG library): for each launch shape, time the
kernel with forced (chunk_pages, team) next to the planner's own choice.
usage (GPU box): PC_LIB_PATH=.../libpagedevict_b200_debug.so python profiles/tools_team_sweep.py"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_paged_compression_b200 import _lib, ops
from benchtools import microbench

lib = _lib.load()
lib.pc_debug_set_decode_plan.argtypes = [ctypes.c_int32, ctypes.c_int32]
lib.pc_debug_set_decode_plan.restype = None
dev = torch.device("cuda", 0)
SHAPES = [(64, 592, 32, 8), (64, 848, 32, 8), (64, 1056, 32, 8), (512, 592, 4, 1), (256, 592, 8, 2),
          (16, 2048, 32, 8), (16, 1024, 32, 8), (4, 4096, 32, 8), (1, 32768, 32, 8), (1, 4096, 32, 8),
          (64, 8192, 32, 8), (16, 4096, 64, 8)]
if len(sys.argv) > 1:
    SHAPES = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]]
for B, ctx, Hq, Hkv in SHAPES:
    pages = -(-ctx // 16)
    lib.pc_debug_set_decode_plan(0, 0)
    plan = ops.paged_decode_plan(B, Hq, Hkv, 128, 16, ctx)
    r = microbench.time_shape(dev, B, ctx, Hq, Hkv)
    print(f"B {B} ctx {ctx} Hq {Hq} Hkv {Hkv} pages/group {pages}: planner cp {plan[0]} T {plan[1]} -> "
          f"{r['us']:.1f} us  {r['frac_of_measured_peak']:.3f}", flush=True)
    cands = []
    for T in (1, 2, 4, 8):
        for n in (1, 2, 3, 4, 6, 9, 18, 37):
            cp = -(-pages // n)
            if cp < 2 or (n > 1 and cp < T):
                continue
            groups = B * Hkv * (-(-(Hq // Hkv) // 8))
            chunks = groups * (-(-pages // cp))
            if T > 1 and chunks > 148 * 8 // T:
                continue
            if T == 1 and chunks > 8 * 1184:
                continue
            cands.append((T, cp, chunks))
    seen = set()
    for T, cp, chunks in cands:
        if (T, cp) in seen:
            continue
        seen.add((T, cp))
        lib.pc_debug_set_decode_plan(cp, T)
        r = microbench.time_shape(dev, B, ctx, Hq, Hkv, reps=30)
        print(f"    T {T} cp {cp:4d} chunks {chunks:5d}: {r['us']:7.1f} us  {r['frac_of_measured_peak']:.3f}", flush=True)
    lib.pc_debug_set_decode_plan(0, 0)
    torch.cuda.empty_cache()
