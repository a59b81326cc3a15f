"""Kernel microbench of BASELINE.json configs[1]: paged decode attention + fused value_l2
scoring, batch 1-256, ctx 1K-32K, GQA 8:1 (and 4:1), block 16, bf16.

q ~ U(-s, s), KV pool ~ U(-s, s), s = D^-0.5 (tests/kernels/test_attention.py:146-148);
block tables are a seeded random permutation of the pool (non-contiguous pages).  Every
timed launch reads a DIFFERENT set of pages: the pool holds enough disjoint table sets
that at least 512 MB (4x L2) is touched before a page is read again, so the numbers are
HBM reads, not L2 hits.  CUDA events around `reps` back-to-back launches, after warm-up.
Prints one JSON object per shape (used by bench.py --sweep; summaries go to profiles/).
"""
from __future__ import annotations

import json
import os

import torch

from vllm_paged_compression_b200 import ops


def _peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"])
    return 6650.0


def time_shape(device, B, ctx, Hq, Hkv, D=128, bs=16, ragged=False, score=True, reps=None,
               max_pool_bytes=48 << 30, seed=0, sequential=False):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    if ragged:
        kv_lens = torch.randint(ctx // 2, ctx + 1, (B,), device=device, generator=g, dtype=torch.int32)
    else:
        kv_lens = torch.full((B,), ctx, device=device, dtype=torch.int32)
    if score:
        kv_lens = (kv_lens // bs) * bs                      # last block just filled -> scored
    lens_host = kv_lens.tolist()
    M = -(-max(lens_host) // bs)
    blocks_per_set = B * M
    bytes_per_launch = sum(lens_host) * Hkv * D * 2 * 2
    block_bytes = bs * Hkv * D * 2
    n_sets = max(2, min(64, -(-(512 << 20) // max(bytes_per_launch, 1))))
    n_sets = max(1, min(n_sets, max_pool_bytes // (2 * blocks_per_set * block_bytes)))
    nb = n_sets * blocks_per_set
    s = D ** -0.5
    kv = torch.empty(2, nb, bs * Hkv * D, dtype=torch.bfloat16, device=device)
    chunk = 1 << 16
    for i in range(0, nb, chunk):                           # fill without a pool-sized fp32 temp
        n = min(chunk, nb - i)
        kv[:, i:i + n] = (torch.rand(2, n, bs * Hkv * D, device=device, generator=g) * 2 * s - s).to(torch.bfloat16)
    kc = kv[0].view(nb, Hkv, D // 8, bs, 8)
    vc = kv[1].view(nb, Hkv, D, bs)
    perm = torch.randperm(nb, device=device, generator=g).to(torch.int32)
    if sequential:                                          # a sequence's blocks are consecutive ids (fresh allocator)
        perm = torch.arange(nb, device=device, dtype=torch.int32)
    tables = [perm[i * blocks_per_set:(i + 1) * blocks_per_set].view(B, M).contiguous() for i in range(n_sets)]
    q = (torch.rand(B, Hq, D, device=device, generator=g) * 2 * s - s).to(torch.bfloat16)
    out = torch.empty_like(q)
    seq_lens = kv_lens.clone() if score else None
    pool = ops.make_score_pool(torch.zeros(nb, Hkv, dtype=torch.float32, device=device), 0) if score else None

    def launch(i):
        ops.paged_decode_attn_score(out, q, kc, vc, tables[i % n_sets], kv_lens, s, max(lens_host),
                                    seq_lens, pool)
    for i in range(max(3, n_sets)):
        launch(i)
    torch.cuda.synchronize()
    if reps is None:
        reps = max(10, min(200, int(0.2 / max(bytes_per_launch / 5e12, 2e-6))))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        launch(i)
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3 / reps
    gbs = bytes_per_launch / sec / 1e9
    del kv, kc, vc
    return {"B": B, "ctx": ctx, "Hq": Hq, "Hkv": Hkv, "D": D, "block": bs, "ragged": ragged,
            "us": sec * 1e6, "kv_bytes": bytes_per_launch, "kv_read_GBps": gbs,
            "frac_of_measured_peak": gbs / _peak(), "table_sets": n_sets, "reps": reps}


def sweep(device, batches=(1, 4, 16, 64, 256), ctxs=(1024, 2048, 4096, 8192, 16384, 32768),
          shapes=((64, 8), (32, 8)), max_tokens=8 << 20):
    rows = []
    for Hq, Hkv in shapes:
        for B in batches:
            for ctx in ctxs:
                if B * ctx > max_tokens:
                    continue
                r = time_shape(device, B, ctx, Hq, Hkv)
                rows.append(r)
                print(json.dumps(r), flush=True)
                torch.cuda.empty_cache()
    return rows


def main():
    """python -m vllm_paged_compression_b200.microbench --B 64 --ctx 592 [--Hq 32 --Hkv 8] : one shape
    (the target of `ncu -k regex:paged_decode`)."""
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=592)
    ap.add_argument("--Hq", type=int, default=32)
    ap.add_argument("--Hkv", type=int, default=8)
    ap.add_argument("--D", type=int, default=128)
    ap.add_argument("--block", type=int, default=16)
    ap.add_argument("--ragged", action="store_true")
    ap.add_argument("--sequential", action="store_true", help="consecutive block ids per sequence")
    ap.add_argument("--no-score", action="store_true")
    ap.add_argument("--reps", type=int, default=None)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    print(json.dumps(time_shape(dev, a.B, a.ctx, a.Hq, a.Hkv, a.D, a.block, a.ragged, score=not a.no_score, reps=a.reps,
                                sequential=a.sequential)), flush=True)


if __name__ == "__main__":
    main()
