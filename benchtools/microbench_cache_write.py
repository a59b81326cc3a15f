"""Prompt-time cache write (W1 + R1 + S3) microbench: one layer's streamingLLM prune gather +
reshape_and_cache + prompt block scoring for a batch of equal prompts.

Algorithmic bytes per launch = kept_tokens * Hkv * D * 2 (K, V) * 2 B read + the
same written.  Every timed launch works on a DIFFERENT source / pool set (enough sets that
>= 512 MB is touched before reuse: HBM traffic, not L2 hits).  CUDA events, after warm-up.
    python -m benchtools.microbench_cache_write [--reqs 16 --len 1024 --evict-freq 2]
"""
from __future__ import annotations

import argparse
import json

import numpy as np
import torch

from vllm_paged_compression_b200 import ops
from benchtools.microbench_prefill import prune_layout


def run(device, reqs, L, Hkv, D, bs, evict_freq, reps=40, score=True):
    g = torch.Generator(device=device)
    g.manual_seed(0)
    kept, pruned, first = prune_layout(L, bs, evict_freq)
    T = reqs * L
    nblk_req = -(-kept // bs)
    nb = reqs * nblk_req
    bytes_rw = reqs * kept * Hkv * D * 2 * 2
    n_sets = max(2, min(24, -(-(512 << 20) // (2 * bytes_rw))))
    s = D ** -0.5
    keys = [(torch.rand(T, Hkv * D, device=device, generator=g) * 2 * s - s).to(torch.bfloat16) for _ in range(n_sets)]
    vals = [(torch.rand(T, Hkv * D, device=device, generator=g) * 2 * s - s).to(torch.bfloat16) for _ in range(n_sets)]
    pools = [torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16, device=device) for _ in range(n_sets)]
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(1)).numpy()
    tabs = perm.reshape(reqs, nblk_req)
    slots = []
    for r in range(reqs):
        sl = (tabs[r][:, None].astype(np.int64) * bs + np.arange(bs)[None, :]).reshape(-1)[:kept]
        slots.append(sl)
    slot = torch.from_numpy(np.concatenate(slots)).to(device)
    m5 = torch.tensor([[r * L, kept, first, pruned, r * kept] for r in range(reqs)], dtype=torch.int32, device=device)
    toff = torch.tensor([r * nblk_req for r in range(reqs + 1)], dtype=torch.int32, device=device)
    pool_sc = ops.make_score_pool(torch.zeros(nb, Hkv, dtype=torch.float32, device=device), 0) if score else None
    views = []
    for p in pools:
        kc = p[0].view(nb, Hkv, D // 8, bs, 8)
        vc = p[1].view(nb, Hkv, D, bs)
        views.append((kc, vc))

    def launch(i):
        kc, vc = views[i % n_sets]
        ops.prune_prompt_and_cache(keys[i % n_sets].view(T, Hkv, D), vals[i % n_sets].view(T, Hkv, D), kc, vc, slot,
                                   m5, toff, reqs * nblk_req, pool_sc)
    for i in range(max(3, n_sets)):
        launch(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        launch(i)
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3 / reps
    # the floor at this size: a plain device copy of the same bytes (read bytes_rw, write bytes_rw),
    # rotating over the same number of buffer sets
    nel = bytes_rw // 2
    srcs = [torch.empty(nel, dtype=torch.bfloat16, device=device).normal_() for _ in range(n_sets)]
    dsts = [torch.empty(nel, dtype=torch.bfloat16, device=device) for _ in range(n_sets)]
    for i in range(n_sets):
        dsts[i].copy_(srcs[i])
    torch.cuda.synchronize()
    e0.record()
    for i in range(reps):
        dsts[i % n_sets].copy_(srcs[i % n_sets])
    e1.record()
    torch.cuda.synchronize()
    copy_us = e0.elapsed_time(e1) * 1e3 / reps
    return {"us": sec * 1e6, "copy_same_bytes_us": copy_us, "bytes": 2 * bytes_rw, "gbs": 2 * bytes_rw / sec / 1e9, "kept": kept, "sets": n_sets,
            "launch": f"{reqs} prompts x {L} tokens, {kept} kept, Hkv {Hkv}, D {D}, block {bs}, one layer, "
                      f"{n_sets} rotating source/pool sets, scoring {'on' if score else 'off'}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reqs", type=int, default=16)
    ap.add_argument("--len", type=int, default=1024)
    ap.add_argument("--Hkv", type=int, default=8)
    ap.add_argument("--D", type=int, default=128)
    ap.add_argument("--block", type=int, default=16)
    ap.add_argument("--evict-freq", type=int, default=2)
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--no-score", action="store_true")
    ap.add_argument("--hc", type=int, default=0,
                    help="heads per CTA override (PC_DEBUG library only, PC_LIB_PATH=...debug.so)")
    a = ap.parse_args()
    if a.hc:
        from vllm_paged_compression_b200._lib import load
        import ctypes
        fn = load().pc_debug_set_cache_write_block_heads
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(a.hc)
    r = run(torch.device("cuda", 0), a.reqs, a.len, a.Hkv, a.D, a.block, a.evict_freq, a.reps, score=not a.no_score)
    if a.hc:
        r["hc"] = a.hc
    print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
