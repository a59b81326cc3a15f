"""Clock trace of CTA 0 of the dual-query-tile prefill kernel (PC_DEBUG library only):
  PC_LIB_PATH=.../libpagedevict_b200_debug.so python profiles/tools_trace_prefill_dq.py [--reqs 64 --len 1024]
Every role of the CTA (softmax warpgroup of stream 0 / 1, MMA issuer 0 / 1, epilogue, K/V producer,
Q producer) stamps clock64 around its waits (csrc/prefill_attn_dq.cu, DQ_TR); this script prints where
the time between and inside work items goes (means over the steady part of the launch) and a merged
timeline of a few consecutive items."""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchtools import microbench_prefill  # noqa: E402
from vllm_paged_compression_b200 import _lib  # noqa: E402

ROLES, CAP = 8, 8192
NAMES = {1: "sm: before S wait (first tile)", 2: "sm: before S wait", 3: "sm: S seen", 4: "sm: P arrived",
         10: "mma: K ready (S issue)", 11: "mma: S part done", 12: "mma: V ready", 13: "mma: P ready",
         14: "mma: O free (PV issue)", 15: "mma: before Q wait", 16: "mma: Q ready",
         20: "epi: s0 before l wait", 21: "epi: s1 before l wait", 22: "epi: s0 l seen", 23: "epi: s1 l seen",
         24: "epi: s0 PV done seen", 25: "epi: s1 PV done seen", 26: "epi: s0 O free", 27: "epi: s1 O free",
         30: "kv: K stage free", 32: "kv: V stage free", 40: "q: Q set free"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reqs", type=int, default=64)
    ap.add_argument("--len", type=int, default=1024)
    ap.add_argument("--evict-freq", type=int, default=2)
    ap.add_argument("--dump-items", type=int, default=2)
    a = ap.parse_args()
    lib = _lib.load()
    fn = lib.pc_debug_prefill_dq_trace
    fn.argtypes, fn.restype = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_size_t], ctypes.c_int32
    dev = torch.device("cuda", 0)
    row, _ = microbench_prefill.run(dev, a.reqs, a.len, 32, 8, 128, a.evict_freq, 5)
    print("untraced:", round(row["us"], 1), "us", round(row["tflops"], 1), "TFLOP/s", row["kernel"])
    fn(1, None, 0)
    row, _ = microbench_prefill.run(dev, a.reqs, a.len, 32, 8, 128, a.evict_freq, 1)   # 3 warm-up launches + 1: the last one stays
    torch.cuda.synchronize()
    host = np.zeros(ROLES * CAP, dtype=np.uint64)
    fn(0, host.ctypes.data_as(ctypes.c_void_p), host.nbytes)
    print("traced launch:", round(row["us"], 1), "us")
    ev = {}
    for r in range(ROLES):
        w = host[r * CAP:(r + 1) * CAP]
        w = w[w != 0]
        code = (w >> np.uint64(56)).astype(np.int64)
        g = ((w >> np.uint64(40)) & np.uint64(0xffff)).astype(np.int64)
        clk = (w & np.uint64(0xffffffffff)).astype(np.int64)
        ev[r] = (code, g, clk)
        print(f"role {r}: {len(w)} stamps")

    def stats(name, x):
        x = np.asarray(x, dtype=np.float64)
        if len(x):
            print(f"  {name:58s} n {len(x):5d}  mean {x.mean():8.0f}  p50 {np.median(x):8.0f}  p90 {np.percentile(x, 90):8.0f}")

    # ---- softmax warps
    first_tiles = None
    for r in (0, 1):
        code, g, clk = ev[r]
        t_before = {int(gg): int(c) for cc, gg, c in zip(code, g, clk) if cc in (1, 2)}
        t_seen = {int(gg): int(c) for cc, gg, c in zip(code, g, clk) if cc == 3}
        t_p = {int(gg): int(c) for cc, gg, c in zip(code, g, clk) if cc == 4}
        firsts = sorted(int(gg) for cc, gg in zip(code, g) if cc == 1)
        first_tiles = firsts
        fs = set(firsts)
        tiles = sorted(t_p)
        tiles = [t for t in tiles if t > tiles[len(tiles) // 10]][: int(len(tiles) * 0.8)]   # steady part
        print(f"softmax warp of stream {r} (clk):")
        stats("wait for S, first tile of an item", [t_seen[t] - t_before[t] for t in tiles if t in fs])
        stats("wait for S, other tiles", [t_seen[t] - t_before[t] for t in tiles if t not in fs])
        stats("S seen -> P arrived, first tile", [t_p[t] - t_seen[t] for t in tiles if t in fs])
        stats("S seen -> P arrived, other tiles", [t_p[t] - t_seen[t] for t in tiles if t not in fs])
        stats("P arrived (last tile) -> before S wait (next item)", [t_before[t] - t_p[t - 1] for t in tiles if t in fs and t - 1 in t_p])
        stats("P arrived -> before S wait (same item)", [t_before[t] - t_p[t - 1] for t in tiles if t not in fs and t - 1 in t_p])
        stats("tile period, all", np.diff([t_p[t] for t in tiles]))
        nt = np.diff(firsts)
        per_item = [t_p[firsts[i + 1] - 1] - t_p[firsts[i] - 1] for i in range(1, len(firsts) - 1) if firsts[i] - 1 in t_p and firsts[i + 1] - 1 in t_p]
        stats("item period", per_item)
        print("  tiles per item: mean %.2f" % nt.mean())
    # ---- issuers
    for r in (2, 3):
        code, g, clk = ev[r]
        fs = set(first_tiles)
        d = {c: {} for c in (10, 11, 12, 13, 14, 15, 16)}
        for cc, gg, c in zip(code, g, clk):
            d[int(cc)][int(gg)] = int(c)
        tiles = sorted(d[14])
        tiles = tiles[len(tiles) // 10: int(len(tiles) * 0.9)]
        print(f"MMA issuer of stream {r - 2} (clk):")
        for nm, a0, b0 in (("wait for V", 11, 12), ("wait for P", 12, 13), ("wait for O free", 13, 14)):
            stats(nm + ", first tile", [d[b0][t] - d[a0][t] for t in tiles if t in fs and t in d[a0]])
            stats(nm + ", other tiles", [d[b0][t] - d[a0][t] for t in tiles if t not in fs and t in d[a0]])
        stats("wait for Q (at the last tile of an item)", [d[16][t] - d[15][t] for t in d[16] if t in d[15]])
        stats("S part (K wait + 8 MMAs + commits): PV issue -> S part done", [d[11][t] - d[14][t - 1] for t in tiles if t - 1 in d[14]])
    # ---- epilogue
    code, g, clk = ev[4]
    d = {c: {} for c in range(20, 28)}
    for cc, gg, c in zip(code, g, clk):
        d[int(cc)][int(gg)] = int(c)
    print("epilogue warp (clk):")
    for i in (0, 1):
        ks = sorted(d[26 + i])
        stats(f"stream {i}: wait for l", [d[22 + i][t] - d[20 + i][t] for t in ks])
        stats(f"stream {i}: wait for the last P.V", [d[24 + i][t] - d[22 + i][t] for t in ks])
        stats(f"stream {i}: P.V done seen -> O free", [d[26 + i][t] - d[24 + i][t] for t in ks])
    ks = sorted(d[26])
    stats("stream 0 O free -> stream 1 before l wait (rest of stream 0's stores)", [d[21][t] - d[26][t] for t in ks if t in d[21]])
    # cross-role: softmax P arrived (last tile of item) -> epilogue O free
    for i in (0, 1):
        code, g, clk = ev[i]
        t_p = {int(gg): int(c) for cc, gg, c in zip(code, g, clk) if cc == 4}
        stats(f"stream {i}: softmax P of the last tile -> O free", [d[26 + i][t] - t_p[t] for t in sorted(d[26 + i]) if t in t_p])
        stats(f"stream {i}: softmax P of the last tile -> P.V done seen by the epilogue", [d[24 + i][t] - t_p[t] for t in sorted(d[24 + i]) if t in t_p])
    # ---- merged timeline of a few items from the middle
    mid = first_tiles[len(first_tiles) // 2]
    end = first_tiles[len(first_tiles) // 2 + a.dump_items]
    rows = []
    for r in range(7):
        code, g, clk = ev[r]
        for cc, gg, c in zip(code, g, clk):
            if r == 6:
                continue
            if mid - 1 <= gg <= end:
                rows.append((int(c), r, int(cc), int(gg)))
    rows.sort()
    t0 = rows[0][0]
    print(f"timeline, tiles {mid - 1}..{end} (first tiles of items: {[t for t in first_tiles if mid - 1 <= t <= end]}):")
    for c, r, cc, gg in rows:
        print(f"  {c - t0:7d}  role {r}  tile {gg:4d}  {NAMES.get(cc, cc)}")


if __name__ == "__main__":
    main()
