#!/usr/bin/env python
"""Generate tests/golden/ref_token_prune.npz by running the REFERENCE's own
KVCachePruner.prune_prompt_inplace (vllm/attention/kvcache_prunner.py:220-279) on CPU, in fp32,
for the token-level methods (global / local -> value_l2, inverse_key_l2).

    python tests/golden/make_golden_token_prune.py        (needs /root/reference)

Inputs are bf16-representable fp32 values from a seeded numpy stream.  A case is kept only when,
for every head, the k-th and (k+1)-th smallest reference scores differ by more than 1e-4
relative, so that the fp32 reduction order (the one declared deviation) cannot change which
tokens go; otherwise the seed is bumped.  Stored: the case parameters and the reference's pruned
K / V as int16 views of bf16 (the prune is a pure gather, so bf16 holds it exactly).
"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
OUT = os.path.join(REPO, "tests", "golden", "ref_token_prune.npz")

import importlib.util  # noqa: E402

import torch  # noqa: E402

spec = importlib.util.spec_from_file_location("ref_pruner", "/root/reference/vllm/attention/kvcache_prunner.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)


from tests.util import token_prune_data as data  # noqa: E402


def separated(pr, k, v, method, bs, budget):
    q_len = k.shape[0]
    rem = q_len % bs
    end_mid = q_len - bs - rem
    mid = end_mid - bs
    kd = mid - (budget - 2 * bs)
    if kd <= 0:
        return True
    sc = pr.get_token_score(k[bs:end_mid], v[bs:end_mid], method)
    s, _ = torch.sort(sc, dim=0)
    a, b = s[kd - 1], s[kd]
    return bool(((b - a) / b.abs().clamp_min(1e-30) > 1e-4).all())


def main():
    cases = [  # (method, bs, budget, q_len, H, D)
        ("global", 16, 128, 200, 2, 64), ("global", 16, 128, 131, 4, 64), ("local", 16, 256, 777, 2, 128),
        ("global", 32, 256, 1000, 2, 64), ("inverse_key_l2", 16, 128, 300, 2, 64),
        ("inverse_key_l2", 16, 512, 1030, 4, 128), ("global", 16, 128, 128, 2, 64), ("global", 16, 128, 161, 8, 64),
    ]
    out = {}
    meta = []
    for ci, (method, bs, budget, q_len, H, D) in enumerate(cases):
        pr = ref.KVCachePruner("budget", method, bs, budget, 1)
        seed = 100 + ci
        while True:
            k, v = data(seed, q_len, H, D)
            if separated(pr, k, v, method, bs, budget):
                break
            seed += 1000
        L = pr.get_pruned_length(q_len)
        ok, ov = torch.empty(L, H, D), torch.empty(L, H, D)
        pr.prune_prompt_inplace(k, v, ok, ov)
        meta.append([{"global": 0, "local": 1, "inverse_key_l2": 2}[method], bs, budget, q_len, H, D, seed, L])
        out[f"k_{ci}"] = ok.to(torch.bfloat16).view(torch.int16).numpy()
        out[f"v_{ci}"] = ov.to(torch.bfloat16).view(torch.int16).numpy()
    out["meta"] = np.asarray(meta, dtype=np.int64)
    np.savez_compressed(OUT, **out)
    print(OUT, os.path.getsize(OUT), "bytes;", len(cases), "cases")


if __name__ == "__main__":
    main()
