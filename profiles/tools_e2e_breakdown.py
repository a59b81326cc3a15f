"""Where the e2e step time goes: PCIe bandwidth of the box, host time of a step (cProfile),
GPU time of a step.  usage (GPU box): python profiles/tools_e2e_breakdown.py"""
import argparse
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def bw():
    n = 24 << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for name, a, b in (("h2d", d, h), ("d2h", h, d)):
        for _ in range(3):
            a.copy_(b, non_blocking=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            a.copy_(b, non_blocking=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 10
        print(f"{name}: {n / dt / 1e9:.1f} GB/s ({dt * 1e6:.0f} us per {n >> 20} MiB)")
    s2 = torch.cuda.Stream()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            h2 = getattr(bw, "h2", None)
            if h2 is None:
                bw.h2 = h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
                bw.d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
            h2.copy_(bw.d2, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(f"duplex: {n / dt / 1e9:.1f} GB/s each way")


def main():
    bw()
    sys.argv = [sys.argv[0]]
    args = bench.parse_args()
    dev = torch.device("cuda", 0)
    for host_io in (False, True):
        arm = bench.Arm(args, dev, 0, 1, host_io=host_io)
        arm.prefill()
        for _ in range(8):
            arm.step()
        torch.cuda.synchronize()
        # host-only time: how long the Python side of a step takes when it never waits
        t0 = time.perf_counter()
        for _ in range(32):
            arm.step()
        t_host = (time.perf_counter() - t0) / 32
        torch.cuda.synchronize()
        t_all = (time.perf_counter() - t0) / 32
        print(f"host_io={host_io}: python issue time {t_host * 1e3:.3f} ms/step, wall incl. drain {t_all * 1e3:.3f} ms/step")
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(16):
            arm.step()
        pr.disable()
        torch.cuda.synchronize()
        st = pstats.Stats(pr)
        st.sort_stats("cumulative").print_stats(22)
        del arm
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
