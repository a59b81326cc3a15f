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
        topk_blocks=topk if topk > 0 else 3, prompt_evict_method=c.get("prompt", "streamingLLM"),
        decode_evict_method=decode, evict_freq=c.get("evict_freq", 2))
    if topk < 0:
        pe.topk_blocks = -1
    pol = Policy(block_size=bs, cache_prune_type=c["prune"], cache_budget=c.get("budget", 128),
                 evict_freq=c.get("evict_freq", 2), decode_evict_method=decode, topk_blocks=topk,
                 prompt_evict_method=c.get("prompt", "streamingLLM"),
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
    import random
    rng = random.Random(1234)        # the oracle's twin of the engine's host RNG (HEAD: shared by all layers)
    layers = [OracleLayer(Hq, D, D ** -0.5, Hkv, pol, rng=rng) for _ in range(num_layers)]
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


def _sampled_prefill_ref(q, k, v, seq_lens, kept_idx, rows_per_req, Hq, Hkv, D, scale, seed):
    """Oracle prefill attention (top-left causal over the kept rows, fp32) for a SAMPLE of
    query rows per request -- the full [Hq, Lq, Lkv] score tensor of an 8192-token prompt
    does not fit a test.  -> (global row indices, ref [n, Hq*D])."""
    G = Hq // Hkv
    rs = np.random.RandomState(seed)
    rows, refs = [], []
    s = 0
    for L, gi in zip(seq_lens, kept_idx):
        idx = np.unique(np.concatenate([[0, L - 1, min(L - 1, 127), min(L - 1, 128)],
                                        rs.randint(0, L, size=rows_per_req)]))
        kk = k[s:s + L].view(L, Hkv, D)[gi].float().repeat_interleave(G, 1).transpose(0, 1)   # [Hq, L', D]
        vv = v[s:s + L].view(L, Hkv, D)[gi].float().repeat_interleave(G, 1).transpose(0, 1)
        qq = q[s:s + L].view(L, Hq, D)[idx].float().transpose(0, 1)                              # [Hq, n, D]
        sc = scale * (qq @ kk.transpose(1, 2))
        j = torch.arange(len(gi))[None, :]
        sc = sc.masked_fill(j > torch.from_numpy(idx)[:, None], float("-inf"))
        refs.append((torch.softmax(sc, -1) @ vv).transpose(0, 1).reshape(len(idx), Hq * D))
        rows.append(torch.from_numpy(idx + s))
        s += L
    return torch.cat(rows), torch.cat(refs)


def run_case_graph(c: Dict, device="cuda:0", num_layers=2, use_graph=True, pdl=None,
                   prefill_rows=None, native_decode=True) -> Dict:
    """The configuration bench.py times: prefill through PagedEvictEngine.step, decode through
    DecodeGraphRunner (packed device state, CUDA-graph replay of all layers, fused cache write,
    PDL per ops.pdl_default) -- against the oracle at every step.  Shapes are the benched ones
    (B 64 x 1024 in for configs[2], ...), so the oracle side uses the reference's own compiled
    CPU kernels (oracle/_ref) for decode when they are available (block 16)."""
    from oracle import policy as opol
    from oracle import ref_cpu
    from vllm_paged_compression_b200.engine import PagedEvictEngine
    from vllm_paged_compression_b200.graph_runner import DecodeGraphRunner
    Hq, Hkv, D, bs = c["Hq"], c["Hkv"], c["D"], c["bs"]
    cache_config, pol = make_configs(c)
    lens = c["prompt_lens"]
    B = len(lens)
    per_seq = -(-max(lens) // bs) + c["steps"] // bs + 3
    num_blocks = B * per_seq + 8
    eng = PagedEvictEngine(num_layers, Hq, Hkv, D, cache_config, num_blocks, device=device,
                           max_num_batched_tokens=c.get("max_num_batched_tokens", 16 * 1024))
    native = ref_cpu.RefCpuOps if (native_decode and bs == 16 and ref_cpu.available()) else None
    layers = [OracleLayer(Hq, D, D ** -0.5, Hkv, pol, native_ops=native) for _ in range(num_layers)]
    kv_ref = [torch.zeros(2, num_blocks, bs * Hkv * D, dtype=torch.bfloat16) for _ in range(num_layers)]
    for i, L in enumerate(lens):
        eng.add_request(f"r{i}", L)
    stats = {"evictions": 0, "layer_disagreements": 0, "steps": 0, "max_err": 0.0, "prefill_max_err": 0.0,
             "oracle_decode": "reference csrc/cpu kernels" if native else "torch restatement"}
    step = 0
    # ---------------------------------------------------------------- prefill (engine.step)
    while eng.waiting:
        cpu_in = {}

        def qkv(layer, n_tok):
            q, k, v = layer_inputs(c["seed"], layer, step, n_tok, Hq, Hkv, D)
            cpu_in[layer] = (q, k, v)
            return q.to(device), k.to(device), v.to(device)
        so = eng.step(qkv)
        assert so.is_prompt
        dev_slots = so.metadata.slot_mapping.cpu().tolist()
        for layer in range(num_layers):
            q, k, v = cpu_in[layer]
            orc = layers[layer]
            full = prefill_rows is None
            ref = orc.prefill(q, k, v, kv_ref[layer], so.seq_lens, so.block_tables, so.request_ids,
                              compute_attention=full)
            assert orc.last["kept"] == so.seq_kv_lens
            assert dev_slots == orc.last["slots"], "prefill slot mapping differs"
            got = so.outputs[layer].float().cpu()
            assert torch.isfinite(got).all()
            if full:
                err = (got - ref.float()).abs().max().item()
            else:
                kept_idx = [torch.from_numpy(opol.prompt_gather_index(L, pol)) for L in so.seq_lens]
                rows, ref = _sampled_prefill_ref(q, k, v, so.seq_lens, kept_idx, prefill_rows, Hq, Hkv, D,
                                                 D ** -0.5, c["seed"] + layer)
                err = (got[rows] - ref).abs().max().item()
            stats["prefill_max_err"] = max(stats["prefill_max_err"], err)
            assert err <= 2e-2, f"prefill layer {layer}: attention output err {err}"
        step += 1
    # ---------------------------------------------------------------- decode (graph runner)
    qs = [tuple(torch.empty(B, w * D, dtype=torch.bfloat16, device=device) for w in (Hq, Hkv, Hkv))
          for _ in range(num_layers)]
    runner = DecodeGraphRunner(eng, qs, use_graph=use_graph, pdl=pdl, max_kv_len=c.get("runner_max_kv_len"))
    runner.record = True
    req_ids = [s.request_id for s in runner.seqs]
    for _ in range(c["steps"]):
        cpu_in = []
        for layer in range(num_layers):
            q, k, v = layer_inputs(c["seed"], layer, step, B, Hq, Hkv, D)
            cpu_in.append((q, k, v))
            for dst, src in zip(qs[layer], (q, k, v)):
                dst.copy_(src)
        rmv0 = runner.step()
        rec = runner.last
        torch.cuda.synchronize()
        trig_rows = rec["trigger_rows"]
        vict = runner.d_victims.cpu().numpy() if trig_rows else None
        for layer in range(num_layers):
            q, k, v = cpu_in[layer]
            orc = layers[layer]
            ref, rmv_ref, _ = orc.decode(q, k, v, kv_ref[layer], rec["seq_lens"], rec["kv_lens"],
                                         rec["tables"], req_ids)
            assert rec["slots"] == orc.last["slots"], f"slot mapping differs at step {step}"
            if pol.decode_evict_method == "value_l2":
                got_rmv = {r: (req_ids[r], int(vict[layer][r])) for r in trig_rows if vict[layer][r] >= 0}
            else:
                got_rmv = {r: (req_ids[r], 1) for r in trig_rows}
            assert got_rmv == rmv_ref, f"step {step} layer {layer}: victims {got_rmv} != oracle {rmv_ref}"
            if layer == 0:
                assert rmv0 == rmv_ref
            elif got_rmv != rmv0:
                stats["layer_disagreements"] += 1
            got = runner.out_flat[layer].float().cpu()
            assert torch.isfinite(got).all()
            err = (got - ref.float()).abs().max().item()
            stats["max_err"] = max(stats["max_err"], err)
            assert err <= 2e-2, f"step {step} layer {layer}: attention output err {err}"
        stats["evictions"] += len(rmv0)
        stats["steps"] += 1
        step += 1
    stats["recaptures"] = runner.recaptures
    runner.sync_host()            # per-token counters of _Seq / BlockTable are booked lazily
    torch.cuda.synchronize()
    for layer in range(num_layers):
        assert np.array_equal(eng.kv_caches[layer].cpu().view(torch.int16).numpy(),
                              kv_ref[layer].view(torch.int16).numpy()), f"layer {layer}: KV cache bytes differ"
        if pol.decode_evict_method == "value_l2":
            pool = eng.impls[layer]._score_pool.cpu().numpy()
            for s in eng.running:
                ids = list(eng.block_manager.get_block_table(s.seq_id))
                n_full = eng.block_manager.get_seq_kv_len(s.seq_id) // bs
                want = layers[layer].scores_of(ids, n_full)
                assert np.array_equal(pool[ids[:n_full]].view(np.int32), want.view(np.int32)), \
                    f"layer {layer} req {s.request_id}: block scores differ"
    return stats
