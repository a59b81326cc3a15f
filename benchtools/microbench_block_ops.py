"""copy_blocks / swap_blocks (backend surface, torch_cuda_sdpa.py:75-88) on B200: bytes moved per
second for the 8B block shape (block 16 x 8 kv heads x 128 x bf16 = 32 KiB per pool per block).
    python -m benchtools.microbench_block_ops
copy_blocks: one launch copies `pairs` (src, dst) blocks inside the K and the V pool (read + write
bytes counted).  swap_blocks: one cudaMemcpyAsync per run of consecutive pairs between a device pool
and a pinned host pool (the reference issues one per pair, csrc/cache_kernels.cu:24-66); the lists
used here are single runs."""
from __future__ import annotations

import json

import torch

from vllm_paged_compression_b200 import ops


def main():
    dev = torch.device("cuda", 0)
    Hkv, D, bs = 8, 128, 16
    nb = 16384                                   # 2 x 512 MiB pools
    kv = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16, device=dev)
    kc = kv[0].view(nb, Hkv, D // 8, bs, 8)
    vc = kv[1].view(nb, Hkv, D, bs)
    block_bytes = bs * Hkv * D * 2
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rows = []
    for pairs in (64, 1024, 8192):
        g = torch.Generator().manual_seed(pairs)
        perm = torch.randperm(nb, generator=g)
        maps = [torch.stack([perm[i * 2 * pairs:i * 2 * pairs + pairs],
                             perm[i * 2 * pairs + pairs:(i + 1) * 2 * pairs]], dim=1).contiguous().to(dev)
                for i in range(max(1, nb // (2 * pairs)))][:8]
        for m in maps:
            ops.copy_blocks(kc, vc, m)
        torch.cuda.synchronize()
        reps = 20
        e0.record()
        for i in range(reps):
            ops.copy_blocks(kc, vc, maps[i % len(maps)])
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        moved = 2 * 2 * pairs * block_bytes      # K and V pool, read + write
        rows.append({"op": "copy_blocks", "pairs": pairs, "us": us, "GBps": moved / us / 1e3})
    host = torch.zeros(1024, bs * Hkv * D, dtype=torch.bfloat16).pin_memory()
    for pairs in (16, 256, 1024):
        m = torch.stack([torch.arange(pairs), torch.arange(pairs)], dim=1).to(torch.int64)
        for direction, (src, dst) in (("d2h", (kv[0], host)), ("h2d", (host, kv[0]))):
            ops.swap_blocks(src, dst, m)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                ops.swap_blocks(src, dst, m)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / 5
            rows.append({"op": f"swap_blocks_{direction}", "pairs": pairs, "us": us,
                         "GBps": pairs * block_bytes / us / 1e3})
    for r in rows:
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
