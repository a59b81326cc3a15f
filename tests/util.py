"""Shared test helpers (seeded synthetic data identical to tests/golden/make_golden.py)."""
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rnd(seed, *shape):
    a = np.random.RandomState(seed).standard_normal(shape).astype(np.float32)
    return torch.from_numpy(a).to(torch.bfloat16)


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def load_npz(name):
    return np.load(os.path.join(GOLDEN, name))


def step_inputs(seed, step, n_tok, Hq, Hkv, D, layer=0):
    """q/k/v of one engine step (of attention layer `layer`), as make_golden.drive_forward
    generated them."""
    sd = seed * 1000 + step * 3 + layer * 500000
    q = (rnd(sd + 0, n_tok, Hq * D).float() * 0.5).to(torch.bfloat16)
    k = (rnd(sd + 1, n_tok, Hkv * D).float() * 0.5).to(torch.bfloat16)
    v = rnd(sd + 2, n_tok, Hkv * D)
    return q, k, v


HEAD_POLICY_TRACES = ["head_global_random_b1_l2", "v1_global_random_b3_l2", "v1_local_random_b3",
                      "head_stream1_b1", "v1_stream1_b3", "v1_invkey_b2_l2"]


def trace_decode_method(cfg):
    """--evict-method of a golden trace -> (decode method, topk) as the reference ran it."""
    m = cfg["method"]
    if m in ("local", "global"):
        return ("random" if cfg.get("random_evict") else "value_l2"), (cfg.get("topk", 3) if m == "local" else -1)
    return m, -1


def token_prune_data(seed, q_len, H, D):
    """K / V [q_len, H, D] fp32 holding bf16-representable values with spread-out per-token norms
    (the stream tests/golden/make_golden_token_prune.py fed to the reference)."""
    rs = np.random.RandomState(seed)
    k = torch.from_numpy(rs.standard_normal((q_len, H, D)).astype(np.float32)).to(torch.bfloat16).float()
    sc = torch.from_numpy(rs.uniform(0.3, 3.0, size=(q_len, H, 1)).astype(np.float32))
    v = (torch.from_numpy(rs.standard_normal((q_len, H, D)).astype(np.float32)) * sc).to(torch.bfloat16).float()
    k = (k * torch.from_numpy(rs.uniform(0.3, 3.0, size=(q_len, H, 1)).astype(np.float32))).to(torch.bfloat16).float()
    return k, v


TOKEN_PRUNE_METHODS = ["global", "local", "inverse_key_l2"]
