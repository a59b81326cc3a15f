"""Host <-> device copy bandwidth of this box, as a rank of the e2e leg sees it.
usage (GPU box):  python profiles/tools_pcie.py [--bind]            one GPU
                  torchrun --nproc-per-node N profiles/tools_pcie.py [--bind]   N ranks at once
Sizes bracket what PinnedLayerIO moves (one layer's fused qkv rows 0.8 MB ... whole step 25 MB).
Reports H2D alone, D2H alone, both at once on two streams, per rank (min over ranks printed by rank 0).
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bind", action="store_true", help="bind to the GPU's NUMA node before allocating")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    info = None
    if a.bind:
        from vllm_paged_compression_b200 import hostmem
        info = hostmem.bind_to_gpu_numa_node(local)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    rows = []
    for mb in (0.75, 3, 12, 25, 96):
        n = int(mb * (1 << 20))
        h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
        h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
        d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
        d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
        h_in.fill_(1); h_out.fill_(2)
        reps = max(5, int(400 / mb))

        def timed(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / reps

        def h2d():
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)

        def d2h():
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)

        def both():
            h2d(); d2h()
        t_h, t_d, t_b = timed(h2d), timed(d2h), timed(both)
        rows.append((mb, n / t_h / 1e9, n / t_d / 1e9, n / t_b / 1e9))
    if world > 1:
        t = torch.tensor(rows, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        rows = t.tolist()
    if rank == 0:
        print(f"ranks {world}, numa bind {info}")
        print(" MB     H2D GB/s   D2H GB/s   duplex GB/s each way   (slowest rank)")
        for mb, a_, b_, c_ in rows:
            print(f"{mb:6.2f} {a_:9.1f} {b_:10.1f} {c_:12.1f}")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
