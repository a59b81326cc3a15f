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


def step_inputs(seed, step, n_tok, Hq, Hkv, D):
    """q/k/v of one engine step, as make_golden.drive_forward generated them."""
    q = (rnd(seed * 1000 + step * 3 + 0, n_tok, Hq * D).float() * 0.5).to(torch.bfloat16)
    k = (rnd(seed * 1000 + step * 3 + 1, n_tok, Hkv * D).float() * 0.5).to(torch.bfloat16)
    v = rnd(seed * 1000 + step * 3 + 2, n_tok, Hkv * D)
    return q, k, v
