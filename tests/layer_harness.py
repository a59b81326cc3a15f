"""Layer-level parity harness: drive the product engine (CUDA path through the C ABI) and
the CPU oracle (oracle/backend.py) with the same seeded activations, step by step, and
compare everything the path produces.  Used by tests/test_gpu_layer.py and
__graft_entry__.smoke().

Bit-exact: slot mappings, per-layer victims (`get_rmv_block_idx`), tables handed to
attention, per-head block scores, final KV-cache bytes.  atol 2e-2: attention outputs
(BASELINE.json north_star tolerance).
"""
from typing import Dict

import numpy as np
import torch

from oracle import scoring
from oracle.backend import OracleLayer
from oracle.policy import Policy
from tests.util import rnd


def make_configs(c: Dict):
    from vllm_paged_compression_b200.config import CacheConfig, PagedEvictConfig
    bs = c["bs"]
    if c.get("prune") is None:
        return CacheConfig(block_size=bs), None
    topk = c.get("topk", -1)
    decode = c.get("decode", "value_l2")
    pe = PagedEvictConfig(
        disable_evict_prefill=c.get("disable_evict_prefill", False), cache_prune_type=c["prune"],
        evict_method=None if topk < 0 else "local", cache_budget=c.get("budget", 128),
        topk_blocks=topk if topk > 0 else 3, prompt_evict_method="streamingLLM",
        decode_evict_method=decode, evict_freq=c.get("evict_freq", 2))
    if topk < 0:
        pe.topk_blocks = -1
    pol = Policy(block_size=bs, cache_prune_type=c["prune"], cache_budget=c.get("budget", 128),
                 evict_freq=c.get("evict_freq", 2), decode_evict_method=decode, topk_blocks=topk,
                 disable_evict_prefill=c.get("disable_evict_prefill", False))
    return CacheConfig(block_size=bs, paged_evict_config=pe, max_num_seqs=c.get("max_num_seqs", 16)), pol


def layer_inputs(seed, layer, step, n_tok, Hq, Hkv, D):
    base = seed * 100003 + layer * 7919 + step * 3
    q = (rnd(base, n_tok, Hq * D).float() * 0.5).to(torch.bfloat16)
    k = (rnd(base + 1, n_tok, Hkv * D).float() * 0.5).to(torch.bfloat16)
    # layer-dependent V scale so that layers disagree on the victim
    v = (rnd(base + 2, n_tok, Hkv * D).float() * (1.0 + 0.37 * layer)).to(torch.bfloat16)
    return q, k, v


def run_case(c: Dict, device="cuda:0", num_layers=2, num_blocks=256, check_outputs=True) -> Dict:
    from vllm_paged_compression_b200.engine import PagedEvictEngine
    Hq, Hkv, D, bs = c["Hq"], c["Hkv"], c["D"], c["bs"]
    cache_config, pol = make_configs(c)
    eng = PagedEvictEngine(num_layers, Hq, Hkv, D, cache_config, num_blocks, device=device,
                           max_num_batched_tokens=c.get("max_num_batched_tokens", 1 << 30))
    assert pol is not None, "the harness drives eviction-enabled cases"
    layers = [OracleLayer(Hq, D, D ** -0.5, Hkv, pol) for _ in range(num_layers)]
    kv_ref = [torch.zeros(2, num_blocks, bs * Hkv * D, dtype=torch.bfloat16) for _ in range(num_layers)]
    for i, L in enumerate(c["prompt_lens"]):
        eng.add_request(f"r{i}", L)
    stats = {"evictions": 0, "layer_disagreements": 0, "steps": 0, "max_err": 0.0}
    step = 0
    arrivals = {int(k): v for k, v in c.get("arrivals", {}).items()}
    n_req = len(c["prompt_lens"])
    while step < c["steps"] + 8 and (step < c["steps"] or eng.waiting):
        cpu_in = {}
        for L in arrivals.get(step, []):          # requests that arrive while others decode
            eng.add_request(f"r{n_req}", L)
            n_req += 1

        def qkv(layer, n_tok):
            q, k, v = layer_inputs(c["seed"], layer, step, n_tok, Hq, Hkv, D)
            cpu_in[layer] = (q, k, v)
            return q.to(device), k.to(device), v.to(device)

        so = eng.step(qkv)
        md = so.metadata
        dev_slots = md.slot_mapping.cpu().tolist()
        for layer in range(num_layers):
            q, k, v = cpu_in[layer]
            orc = layers[layer]
            if so.is_prompt:
                ref = orc.prefill(q, k, v, kv_ref[layer], so.seq_lens, so.block_tables, so.request_ids)
                assert orc.last["kept"] == so.seq_kv_lens, (orc.last["kept"], so.seq_kv_lens)
                rmv_ref = {}
            else:
                ref, rmv_ref, _ = orc.decode(q, k, v, kv_ref[layer], so.seq_lens, so.seq_kv_lens,
                                             so.block_tables, so.request_ids)
            assert dev_slots == orc.last["slots"], f"slot mapping differs at step {step}"
            got_rmv = eng.impls[layer].get_rmv_block_idx()
            assert got_rmv == rmv_ref, f"step {step} layer {layer}: victims {got_rmv} != oracle {rmv_ref}"
            if layer > 0 and got_rmv != eng.impls[0].get_rmv_block_idx():
                stats["layer_disagreements"] += 1
            if check_outputs:
                got = so.outputs[layer].float().cpu()
                assert torch.isfinite(got).all()
                err = (got - ref.float()).abs().max().item()
                stats["max_err"] = max(stats["max_err"], err)
                assert err <= 2e-2, f"step {step} layer {layer}: attention output err {err}"
        stats["evictions"] += len(so.rmv)
        stats["steps"] += 1
        step += 1
    torch.cuda.synchronize()
    # end state: cache bytes and every live score, per layer
    for layer in range(num_layers):
        assert np.array_equal(eng.kv_caches[layer].cpu().view(torch.int16).numpy(),
                              kv_ref[layer].view(torch.int16).numpy()), f"layer {layer}: KV cache bytes differ"
        if pol.decode_evict_method == "value_l2":
            pool = eng.impls[layer]._score_pool.cpu().numpy()          # [num_blocks, Hkv]
            for s in eng.running:
                ids = list(eng.block_manager.get_block_table(s.seq_id))
                n_full = eng.block_manager.get_seq_kv_len(s.seq_id) // bs     # every full block is scored
                want = layers[layer].scores_of(ids, n_full)
                got = pool[ids[:n_full]]
                assert np.array_equal(got.view(np.int32), want.view(np.int32)), \
                    f"layer {layer} req {s.request_id}: block scores differ"
    return stats
