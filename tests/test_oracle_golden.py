"""Pin the oracle against vectors produced by the REFERENCE's own code
(tests/golden/make_golden.py ran /root/reference's Python classes and its CPU kernels).
CPU only."""
import numpy as np
import pytest
import torch

from oracle import paged_ops, policy, scoring
from oracle.backend import OracleLayer
from oracle.policy import Policy
from tests.util import HEAD_POLICY_TRACES, trace_decode_method, load_json, load_npz, rnd, step_inputs

POL = load_json("ref_policy.json")


def test_budget_lengths():
    # page_evict_kv_util.py:6-36 and kvcache_prunner.py:284-326
    for bs, budget, q, nb, kept_util, kept_pruner in POL["budget_lengths"]:
        p = Policy(block_size=bs, cache_prune_type="budget", cache_budget=budget)
        assert policy.num_required_blocks_after_prune(q, p) == (nb, kept_util), (bs, budget, q)
        assert policy.prompt_prune_layout(q, p)[0] == kept_pruner, (bs, budget, q)


def test_budget_gather():
    # kvcache_prunner.py:183-218
    for row in POL["budget_gather"]:
        p = Policy(block_size=row["bs"], cache_budget=row["budget"])
        assert policy.prompt_gather_index(row["q_len"], p).tolist() == row["kept_src"]


def test_percentage_lengths():
    # page_evict_kv_util_bak.py:6-31, kvcache_prunner_old.py:141-200
    for bs, freq, q, nb, kept, first, pruned in POL["percentage_lengths"]:
        p = Policy(block_size=bs, cache_prune_type="percentage", evict_freq=freq)
        k, pr, f = policy.prompt_prune_layout(q, p)
        assert k == kept and policy.cdiv(k, bs) == nb, (bs, freq, q)
        if pruned:
            assert (f, pr) == (first, pruned), (bs, freq, q)
            gi = policy.prompt_gather_index(q, p)
            assert gi[first - 1] == first - 1 and gi[first] == first + pruned
        else:
            assert pr == 0


def test_readme_eviction_fractions():
    # q_len 1024, bs 16 -> freq 2 keeps 544, freq 4 keeps 784
    assert policy.prompt_prune_layout(1024, Policy(cache_prune_type="percentage", evict_freq=2))[0] == 544
    assert policy.prompt_prune_layout(1024, Policy(cache_prune_type="percentage", evict_freq=4))[0] == 784


def test_slot_mapping():
    # torch_cuda_sdpa.py:565-598
    for row in POL["slot_mapping"]:
        if row["is_prompt"]:
            got = policy.prefill_slot_mapping(row["table"], row["query_len"], row["bs"])
        else:
            got = [policy.decode_slot_mapping(row["table"], row["context_len"], row["bs"])]
        assert got == row["slots"]


def test_compaction():
    # torch_cuda_sdpa.py:1562-1586 (row 0) and torch_cuda_sdpa_v1.py list semantics
    for row in POL["compact_row0"]:
        t = np.array([row["table"]], dtype=np.int32)
        out = policy.compact_block_tables(t, [len(row["table"])], {0: row["idx"]})
        assert out[0, :len(row["out"])].tolist() == row["out"]
        assert out[0, len(row["out"]):].tolist() == [0] * (t.shape[1] - len(row["out"]))
    c = POL["compact_v1"]
    M = max(len(t) for t in c["tables"])
    t = np.zeros((len(c["tables"]), M), dtype=np.int32)
    for i, row in enumerate(c["tables"]):
        t[i, :len(row)] = row
    n = [len(r) for r in c["tables"]]
    vict = {int(k): v for k, v in c["victims"].items()}
    out = policy.compact_block_tables(t, n, vict)
    n_after = [n[i] - (1 if i in vict else 0) for i in range(len(n))]
    assert policy.narrow_compacted(out, n_after).tolist() == c["out"]


def test_argmin_and_windows():
    # torch_cuda_sdpa.py:1528-1544
    for row in POL["argmin"]:
        p = Policy(decode_evict_method="value_l2",
                   topk_blocks=-1 if row["method"] == "global" else row["topk"])
        lo, hi = policy.victim_candidates(row["n_blocks"], p)
        sc = np.array(row["scores"], dtype=np.float32)
        idx = policy.argmin_lowest_index(sc, lo, hi)
        assert idx == row["rmv_idx"], row
        t = np.arange(50, 50 + row["n_blocks"], dtype=np.int32)[None]
        out = policy.compact_block_tables(t, [row["n_blocks"]], {0: idx})
        assert out[0, :row["n_blocks"] - 1].tolist() == row["out_table"]


def test_decode_trigger_matches_block_table_trace():
    # model_runner.py:540-543 + block_table.py:136-196,462-466
    for row in POL["block_table_trace"]:
        p = Policy(block_size=row["bs"], cache_budget=row["budget"], decode_evict_method="streamingLLM")
        first = row["trace"][0]
        nb, kept = policy.num_required_blocks_after_prune(row["q_len"], p)
        assert kept == first["kv_len"] and nb == len(first["ids"])
        for rec in row["trace"][1:]:
            assert policy.decode_trigger(rec["seq_len"], rec["kv_len"], p) == rec["trigger"]


def test_native_ops_against_reference_cpu_kernels():
    g = load_npz("ref_ops.npz")
    nb, bs, Hkv, Hq, D = 48, 16, 2, 8, 64
    kv = torch.zeros(2, nb, bs * Hkv * D, dtype=torch.bfloat16)
    kc, vc = paged_ops.split_kv_cache(kv, Hkv, D)
    paged_ops.reshape_and_cache(rnd(2, 200, Hkv, D), rnd(3, 200, Hkv, D), kc, vc,
                                torch.from_numpy(g["rc_slots"]))
    assert np.array_equal(kv.view(torch.int16).numpy(), g["rc_kv_cache"])     # exact
    q = (rnd(4, 4, Hq, D).float() * 0.3).to(torch.bfloat16)
    kv2 = (rnd(5, 2, nb, bs * Hkv * D).float() * 0.5).to(torch.bfloat16)
    kc2, vc2 = paged_ops.split_kv_cache(kv2, Hkv, D)
    o = paged_ops.paged_attention(q, kc2, vc2, g["pa_block_tables"].tolist(),
                                  g["pa_kv_lens"].tolist(), D ** -0.5)
    # reference test bar: atol 1e-3 (tests/kernels/test_attention.py:323-328) on fp32 math;
    # both sides round the output to bf16 here, so allow one bf16 ulp at |x|<0.5 (2^-9)
    np.testing.assert_allclose(o.float().numpy(), g["pa_out"], atol=2e-3, rtol=0)
    # value_l2 definition vs get_block_score (kvcache_prunner.py:55) evaluated in fp32
    for blk, ref in enumerate(g["score_ref_fp32"]):
        vb = paged_ops.depage_v_block(vc2, blk).float().numpy()
        got = float(scoring.block_score(vb))
        assert abs(got - ref) <= 2e-5 * abs(ref), (blk, got, ref)
        assert abs(scoring.reference_formula_fp32(vb) - ref) <= 2e-5 * abs(ref)


@pytest.mark.parametrize("name", ["head_stream_b1", "v1_stream_b3", "v1_stream_b2_noprefillevict"])
def test_forward_trace_streaming(name):
    """Replay the reference forward() traces through the oracle layer: slot mappings,
    victims, tables handed to attention, kv lens are exact; outputs within bf16 tol."""
    tr = load_json("ref_forward.json")[name]
    arr = load_npz("ref_forward.npz")
    c = tr["cfg"]
    p = Policy(block_size=c["bs"], cache_budget=c["budget"], decode_evict_method="streamingLLM",
               disable_evict_prefill=c["disable_evict_prefill"])
    layer = OracleLayer(c["Hq"], c["D"], c["D"] ** -0.5, c["Hkv"], p, head_compat_uncompacted=True)
    kv = torch.zeros(2, 256, c["bs"] * c["Hkv"] * c["D"], dtype=torch.bfloat16)
    B = len(c["prompt_lens"])
    rids = [f"r{i}" for i in range(B)]
    for step, rec in enumerate(tr["steps"]):
        n_tok = sum(rec["seq_lens"]) if rec["is_prompt"] else B
        q, k, v = step_inputs(c["seed"], step, n_tok, c["Hq"], c["Hkv"], c["D"])
        if rec["is_prompt"]:
            out = layer.prefill(q, k, v, kv, rec["seq_lens"], rec["tables"], rids)
            assert layer.last["kept"] == rec["kv_lens"]
        else:
            out, rmv, attn_tables = layer.decode(q, k, v, kv, rec["seq_lens"], rec["kv_lens"],
                                                 rec["tables"], rids)
            assert {str(r): [rid, idx] for r, (rid, idx) in rmv.items()} == rec["rmv"]
            assert rec["attn_kv_lens"] == rec["kv_lens"]
            for r in range(B):      # tables the reference handed to paged_attention (0-padded)
                n = len(attn_tables[r])
                assert rec["attn_tables"][r][:n] == attn_tables[r]
        assert layer.last["slots"] == rec["slots"]
        ref = arr[f"{name}__out_{step}"].astype(np.float32)
        np.testing.assert_allclose(out.float().numpy(), ref, atol=2e-2, rtol=0)
    assert np.array_equal(kv.view(torch.int16).numpy(), arr[f"{name}__kv_cache_final"])


@pytest.mark.parametrize("name", HEAD_POLICY_TRACES)
def test_forward_trace_head_policies(name):
    """HEAD's other --evict-method choices as shipped: random block eviction for local / global
    (host RNG replay: ONE random.Random(1234) shared by the layers, drawn in layer order, rows in
    row order), streamingLLM-1 and inverse_key_l2 slot zeroing.  Victims of every layer, every
    slot written (step tokens + zeroed slots), tables handed to attention: exact."""
    import random
    tr = load_json("ref_forward.json")[name]
    arr = load_npz("ref_forward.npz")
    c = tr["cfg"]
    method, topk = trace_decode_method(c)
    p = Policy(block_size=c["bs"], cache_budget=c["budget"], decode_evict_method=method, topk_blocks=topk)
    L = c.get("layers", 1)
    rng = random.Random(1234)
    layers = [OracleLayer(c["Hq"], c["D"], c["D"] ** -0.5, c["Hkv"], p, head_compat_uncompacted=True, rng=rng)
              for _ in range(L)]
    kvs = [torch.zeros(2, 256, c["bs"] * c["Hkv"] * c["D"], dtype=torch.bfloat16) for _ in range(L)]
    B = len(c["prompt_lens"])
    rids = [f"r{i}" for i in range(B)]
    n_evict = n_zero = 0
    for step, rec in enumerate(tr["steps"]):
        n_tok = sum(rec["seq_lens"]) if rec["is_prompt"] else B
        for li, layer in enumerate(layers):
            sfx = "" if li == 0 else f"_l{li}"
            q, k, v = step_inputs(c["seed"], step, n_tok, c["Hq"], c["Hkv"], c["D"], layer=li)
            if rec["is_prompt"]:
                out = layer.prefill(q, k, v, kvs[li], rec["seq_lens"], rec["tables"], rids)
                assert layer.last["kept"] == rec["kv_lens"]
            else:
                out, rmv, attn_tables = layer.decode(q, k, v, kvs[li], rec["seq_lens"], rec["kv_lens"],
                                                     rec["tables"], rids)
                assert {str(r): [rid, idx] for r, (rid, idx) in rmv.items()} == rec["rmv" + sfx], (step, li)
                assert layer.last["slots"] + layer.last["zero_slots"] == rec["write_slots" + sfx], (step, li)
                n_evict += len(rmv)
                n_zero += len(layer.last["zero_slots"])
                for r in range(B):
                    n = len(attn_tables[r])
                    assert rec["attn_tables" + sfx][r][:n] == attn_tables[r]
            ref = arr[f"{name}__out_{step}{sfx}"].astype(np.float32)
            np.testing.assert_allclose(out.float().numpy(), ref, atol=2e-2, rtol=0)
    for li in range(L):
        sfx = "" if li == 0 else f"_l{li}"
        assert np.array_equal(kvs[li].view(torch.int16).numpy(), arr[f"{name}__kv_cache_final{sfx}"])
    if method == "inverse_key_l2":
        assert n_evict == 0 and n_zero > 0
    else:
        assert n_evict > 0 and (n_zero > 0) == (method == "streamingLLM-1")


def test_head_l2_branch_argmin_bookkeeping():
    """`head_global_b1_quirks`: HEAD with enable_random_evict = False.  Its block scores carry
    the reference's layout / dtype quirks and are not reproduced; what IS pinned is
    the selection over them -- strict `<` scan of [1, n-1), lowest index wins
    (torch_cuda_sdpa.py:1528-1541) -- on the score lists the reference recorded."""
    tr = load_json("ref_forward.json")["head_global_b1_quirks"]
    p = Policy(block_size=16, cache_budget=128, decode_evict_method="value_l2", topk_blocks=-1)
    n = 0
    for rec in tr["steps"]:
        if rec["is_prompt"] or not rec["rmv"]:
            continue
        sc = np.asarray(rec["l2norms"]["r0"], dtype=np.float32)
        lo, hi = policy.victim_candidates(len(sc), p)
        assert policy.argmin_lowest_index(sc, lo, hi) == rec["rmv"]["0"][1]
        n += 1
    assert n >= 3


def test_value_l2_definition_properties():
    rs = np.random.RandomState(0)
    v = torch.from_numpy(rs.standard_normal((16, 8, 128)).astype(np.float32)).to(torch.bfloat16).float().numpy()
    ph = scoring.per_head_block_score(v)
    assert ph.dtype == np.float32 and ph.shape == (8,)
    # per-head partials are independent (TP sharding by kv head relies on it)
    for h in range(8):
        assert scoring.per_head_block_score(v[:, h:h + 1])[0] == ph[h]
    ref = scoring.reference_formula_fp32(v)
    assert abs(float(scoring.combine_heads(ph)) - ref) < 1e-5 * ref


def test_token_prune_oracle_matches_reference():
    """f3: oracle/token_prune.py against the reference's own prune_prompt_inplace output
    (kvcache_prunner.py:220-279, run in fp32 on CPU by make_golden_token_prune.py): kept vectors
    and the head-mixing re-pack order, bit for bit."""
    from oracle import token_prune
    from tests.util import TOKEN_PRUNE_METHODS, token_prune_data
    g = load_npz("ref_token_prune.npz")
    n_pruned = 0
    for ci, (mi, bs, budget, q_len, H, D, seed, L) in enumerate(g["meta"].tolist()):
        method = TOKEN_PRUNE_METHODS[mi]
        k, v = token_prune_data(seed, q_len, H, D)
        kind = "inverse_key_l2" if method == "inverse_key_l2" else "value_l2"
        ok, ov = token_prune.prune_prompt(k.numpy(), v.numpy(), budget, bs, kind, repack=True)
        assert ok.shape == (L, H, D)
        want_k = torch.from_numpy(g[f"k_{ci}"]).view(torch.bfloat16).float().numpy()
        want_v = torch.from_numpy(g[f"v_{ci}"]).view(torch.bfloat16).float().numpy()
        assert np.array_equal(ok, want_k) and np.array_equal(ov, want_v), ci
        n_pruned += int(L < q_len)
        # the per-head variant keeps exactly the same (token, head) vectors, in their own columns
        pk, _ = token_prune.prune_prompt(k.numpy(), v.numpy(), budget, bs, kind, repack=False)
        assert pk.shape == ok.shape
        first, mid, kd = token_prune.prune_layout(q_len, budget, bs)
        if kd:
            assert sorted(map(bytes, pk[first:first + mid - kd].reshape(-1, D))) == \
                sorted(map(bytes, ok[first:first + mid - kd].reshape(-1, D)))
    assert n_pruned >= 6
