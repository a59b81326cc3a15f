"""CPU tests of the product's host side: policy arithmetic and block bookkeeping against
the golden vectors recorded from the reference's own classes, the metadata builder, the
CLI flag mapping, the score-row map, and the C-ABI library (loads, exports every symbol
include/pagedevict_b200.h declares -- no compute calls without a GPU)."""
import argparse
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import policy as opolicy
from oracle.policy import Policy
from tests.util import load_json
from vllm_paged_compression_b200 import policy
from vllm_paged_compression_b200.block_manager import BlockTable, NaiveBlockAllocator, SelfAttnBlockSpaceManager
from vllm_paged_compression_b200.config import (CacheConfig, PagedEvictConfig, add_cli_args,
                                                paged_evict_config_from_args)

POL = load_json("ref_policy.json")
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _budget(budget):
    return PagedEvictConfig(cache_prune_type="budget", evict_method="streamingLLM", cache_budget=budget)


def _pct(freq):
    return PagedEvictConfig(cache_prune_type="percentage", prompt_evict_method="streamingLLM",
                            decode_evict_method="value_l2", evict_freq=freq)


def test_budget_lengths_golden():
    for bs, budget, q, nb, kept_util, kept_pruner in POL["budget_lengths"]:
        cfg = _budget(budget)
        assert policy.num_required_blocks_after_prune(q, cfg, bs) == (nb, kept_util)
        assert policy.prompt_prune_layout(q, cfg, bs)[0] == kept_pruner


def test_percentage_lengths_golden():
    for bs, freq, q, nb, kept, first, pruned in POL["percentage_lengths"]:
        k, pr, f = policy.prompt_prune_layout(q, _pct(freq), bs)
        assert (k, policy.cdiv(k, bs)) == (kept, nb)
        if pruned:
            assert (f, pr) == (first, pruned)


def test_policy_matches_oracle_exhaustively():
    for bs in (16, 32):
        for q in list(range(1, 700)) + [1024, 4096, 8191, 8192]:
            for budget in (128, 256, 512):
                if budget < 3 * bs:
                    continue
                o = Policy(block_size=bs, cache_prune_type="budget", cache_budget=budget)
                assert policy.prompt_prune_layout(q, _budget(budget), bs) == opolicy.prompt_prune_layout(q, o)
                assert policy.num_required_blocks_after_prune(q, _budget(budget), bs) == \
                    opolicy.num_required_blocks_after_prune(q, o)
            for freq in (1, 2, 3, 4):
                o = Policy(block_size=bs, cache_prune_type="percentage", evict_freq=freq)
                assert policy.prompt_prune_layout(q, _pct(freq), bs) == opolicy.prompt_prune_layout(q, o)


def test_block_table_trace_golden():
    """BlockTable allocate / append / free_prunned_blocks against the reference's BlockTable
    (ids, kv lens, trigger steps)."""
    for row in POL["block_table_trace"]:
        bs, cfg = row["bs"], _budget(row["budget"])
        bt = BlockTable(bs, NaiveBlockAllocator(4096), cfg)
        bt.allocate(row["q_len"])
        first = row["trace"][0]
        assert bt.physical_block_ids == first["ids"] and bt.get_seq_kv_len() == first["kv_len"]
        for rec in row["trace"][1:]:
            bt.append_slots(1)
            assert bt.num_full_slots == rec["seq_len"] and bt.get_seq_kv_len() == rec["kv_len"]
            assert bt.physical_block_ids[:len(rec["ids"])] == rec["ids"]
            trig = policy.decode_trigger(rec["seq_len"], bt.get_seq_kv_len(), cfg, bs)
            assert trig == rec["trigger"]
            if trig:
                bt.free_prunned_blocks(1)
                assert bt.get_seq_kv_len() == rec["kv_len_after_free"]
                assert bt.physical_block_ids == rec["ids_after_free"]


def test_block_manager_accounting():
    bm = SelfAttnBlockSpaceManager(16, 40, _budget(128))
    assert bm.can_allocate(1024)            # 8 blocks after pruning, not 64
    bm.allocate(0, 1024)
    assert len(bm.get_block_table(0)) == 8 and bm.get_seq_kv_len(0) == 128
    assert not SelfAttnBlockSpaceManager(16, 7, _budget(128)).can_allocate(1024)
    free0 = bm.allocator.get_num_free_blocks()
    bm.append_slots(0)
    bm.free_prunned_blocks(0, 1)
    assert bm.allocator.get_num_free_blocks() == free0 and bm.get_seq_kv_len(0) == 113
    bm.free(0)
    assert bm.allocator.get_num_free_blocks() == 40


def test_slot_mapping_builder_golden():
    from vllm_paged_compression_b200.backend import prefill_slots_np
    for row in POL["slot_mapping"]:
        if row["is_prompt"]:
            assert prefill_slots_np(row["table"], row["query_len"], row["bs"]).tolist() == row["slots"]
        else:
            assert [row["table"][-1] * row["bs"] + row["context_len"] % row["bs"]] == row["slots"]


def test_cli_flag_sets():
    from vllm_paged_compression_b200.config import PagedEvictConfig as _PEC
    p = add_cli_args(argparse.ArgumentParser())
    # README flags (README.md:38-45)
    a = p.parse_args("--enable-paged-eviction --cache-prune-type percentage --prompt-evict-method streamingLLM "
                     "--decode-evict-method value_l2 --evict-freq 4 --block-size 16".split())
    c = paged_evict_config_from_args(a)
    assert (c.cache_prune_type, c.prompt_evict_method, c.decode_evict_method, c.evict_freq, c.topk_blocks) == \
        ("percentage", "streamingLLM", "value_l2", 4, -1)
    # HEAD flags (scripts/latency_test.sh:57-72)
    a = p.parse_args("--enable-paged-eviction --cache-prune-type budget --evict-method global --cache-budget 512 "
                     "--disable-evict-prefill".split())
    c = paged_evict_config_from_args(a)
    assert (c.decode_evict_method, c.cache_budget, c.topk_blocks, c.disable_evict_prefill) == ("value_l2", 512, -1, True)
    a = p.parse_args("--enable-paged-eviction --evict-method local --topk-blocks 5".split())
    c = paged_evict_config_from_args(a)
    assert (c.decode_evict_method, c.topk_blocks) == ("value_l2", 5)
    a = p.parse_args("--enable-paged-eviction --evict-method streamingLLM --cache-budget 64".split())
    assert paged_evict_config_from_args(a).cache_budget == 128            # floored, vllm/config.py:977-980
    assert paged_evict_config_from_args(p.parse_args([])) is None
    # every HEAD --evict-method (scripts/diff_page_sizes.sh:12) resolves to a (prompt, decode) pair
    want = {"streamingLLM": ("streamingLLM", "streamingLLM"), "streamingLLM-1": ("streamingLLM", "streamingLLM-1"),
            "inverse_key_l2": ("token_inverse_key_l2", "inverse_key_l2"),
            "local": ("streamingLLM", "value_l2"), "global": ("streamingLLM", "value_l2")}
    for m, pair in want.items():
        c = paged_evict_config_from_args(p.parse_args(f"--enable-paged-eviction --evict-method {m}".split()))
        assert (c.prompt_evict_method, c.decode_evict_method) == pair
    # HEAD as shipped: enable_random_evict = True (torch_cuda_sdpa.py:802)
    c = paged_evict_config_from_args(p.parse_args(
        "--enable-paged-eviction --evict-method global --head-random-evict".split()))
    assert (c.prompt_evict_method, c.decode_evict_method, c.topk_blocks) == ("token_value_l2", "random", -1)
    with pytest.raises(ValueError):       # the HEAD-only methods exist in budget mode only
        _PEC(cache_prune_type="percentage", evict_method="streamingLLM-1")
    with pytest.raises(ValueError):
        CacheConfig(block_size=16, paged_evict_config=_budget(200))       # budget % block_size != 0


def test_backend_surface_names():
    from vllm_paged_compression_b200.backend import (PagedAttention, TorchCUDASDPABackend,
                                                     TorchCUDASDPABackendImpl, TorchCUDASDPAMetadata,
                                                     TorchCUDASDPAMetadataBuilder)
    B = TorchCUDASDPABackend
    assert B.get_name() == "TORCH_SDPA"
    assert B.get_impl_cls() is TorchCUDASDPABackendImpl and B.get_metadata_cls() is TorchCUDASDPAMetadata
    assert B.get_builder_cls() is TorchCUDASDPAMetadataBuilder
    assert B.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16 * 8 * 128)
    impl = TorchCUDASDPABackendImpl(32, 128, 0.088, 8, None, None, "auto",
                                    cache_config=CacheConfig(16, paged_evict_config=_pct(2)),
                                    prefix="model.layers.7.self_attn.attn")
    assert impl.layer_idx == 7 and impl.enabled_paged_eviction and impl.get_rmv_block_idx() == {}
    with pytest.raises(NotImplementedError):
        TorchCUDASDPABackendImpl(32, 128, 0.088, 8, None, None, "fp8")
    with pytest.raises(ValueError):
        TorchCUDASDPABackendImpl(32, 80, 0.1, 8, None, None, "auto")
    import torch
    kv = torch.zeros(2, 3, 16 * 8 * 128, dtype=torch.bfloat16)
    kc, vc = PagedAttention.split_kv_cache(kv, 8, 128)
    assert kc.shape == (3, 8, 16, 16, 8) and vc.shape == (3, 8, 128, 16)


def test_c_abi_library_exports_every_declared_symbol():
    from vllm_paged_compression_b200 import _lib
    hdr = open(os.path.join(REPO, "include", "pagedevict_b200.h")).read()
    declared = set(re.findall(r"\b(pc_[a-z0-9_]+)\s*\(", hdr)) - {"pc_status", "pc_dtype"}
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().pc_abi_version() == _lib.ABI_VERSION
    assert ctypes.sizeof(_lib.PcEvictPolicy) == 32


def test_ops_refuse_cpu_tensors():
    import torch
    from vllm_paged_compression_b200 import ops
    kv = torch.zeros(2, 2, 16 * 2 * 64, dtype=torch.bfloat16)
    kc, vc = kv[0].view(2, 2, 8, 16, 8), kv[1].view(2, 2, 64, 16)
    k = torch.zeros(1, 2, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(1, dtype=torch.int64))


# ------------------------------------------------------------- f1: the fork's real host objects
def _product_builder_md(inter_records, seq_lens, query_lens, block_size, cfg, as_objects=None):
    """Feed the product's TorchCUDASDPAMetadataBuilder an input builder holding `inter_data_list`
    (recorded attribute values, or the reference's live objects) and build on CPU."""
    import types
    from vllm_paged_compression_b200.backend import TorchCUDASDPAMetadataBuilder
    from vllm_paged_compression_b200.config import CacheConfig as CC
    if as_objects is None:
        inter = []
        for r in inter_records:
            inter.append(types.SimpleNamespace(
                request_id=r["request_id"], is_prompt=r["is_prompt"], seq_ids=r["seq_ids"],
                block_tables=None if r["block_tables"] is None else {int(k): v for k, v in r["block_tables"].items()},
                input_tokens=[[0] * n for n in r["input_token_lens"]], orig_seq_lens=r["orig_seq_lens"],
                seq_lens=r["seq_lens"], query_lens=r["query_lens"], context_lens=r["context_lens"],
                seq_kv_lens=None if r["seq_kv_lens"] is None else {int(k): v for k, v in r["seq_kv_lens"].items()},
                prefix_cache_hit=r["prefix_cache_hit"]))
    else:
        inter = as_objects
    runner = types.SimpleNamespace(device="cpu", pin_memory=False,
                                   cache_config=CC(block_size=block_size, paged_evict_config=cfg))
    ib = types.SimpleNamespace(runner=runner, sliding_window=None, block_size=block_size,
                               chunked_prefill_enabled=False, inter_data_list=inter)
    return TorchCUDASDPAMetadataBuilder(ib).build(seq_lens, query_lens, -1, len(seq_lens))


def _assert_md_equal(md, want):
    for k in ("seq_lens", "seq_kv_lens"):
        assert list(getattr(md, k)) == want[k], k
    for k in ("num_prefills", "num_prefill_tokens", "num_decode_tokens", "max_query_len", "max_prefill_seq_len",
              "max_decode_seq_len"):
        assert int(getattr(md, k)) == want[k], k
    for k in ("slot_mapping", "seq_lens_tensor", "seq_kv_lens_tensor", "block_tables", "query_start_loc",
              "seq_start_loc", "context_lens_tensor"):
        assert getattr(md, k).cpu().tolist() == want[k], k
    got = {r: {str(s): list(t) for s, t in d.items()} for r, d in md.reqid_mapping_blocktables.items()}
    assert got == want["reqid_mapping_blocktables"]


@pytest.mark.parametrize("name", ["stream_b3", "stream1_b2_bs32", "global_b2", "invkey_b2"])
def test_metadata_builder_on_reference_host_trace(name):
    """f1: tests/golden/ref_host.json holds what the fork's REAL Scheduler + SelfAttnBlockSpaceManager
    + ModelInputForGPUBuilder produced step by step (prefill, decode, eviction feedback) and the
    metadata its own TorchCUDASDPAMetadataBuilder built from it.  (a) the product builder, fed the
    same inter_data attribute values, builds identical metadata; (b) the product's block manager +
    policy arithmetic, stepped on its own, reproduces the reference scheduler's block tables and
    kv lengths (incl. the allocator's recycling order); (c) the impl's per-step host preparation
    finds the same trigger rows the reference fed back."""
    from vllm_paged_compression_b200.backend import TorchCUDASDPABackendImpl
    from vllm_paged_compression_b200.block_manager import SelfAttnBlockSpaceManager
    from vllm_paged_compression_b200.config import CacheConfig as CC
    tr = load_json("ref_host.json")[name]
    c = tr["cfg"]
    bs = c["block_size"]
    cfg = PagedEvictConfig(cache_prune_type="budget", evict_method=c["evict_method"], cache_budget=c["budget"])
    bm = SelfAttnBlockSpaceManager(bs, 512, cfg)
    impl = TorchCUDASDPABackendImpl(8, 64, 0.125, 2, None, None, "auto",
                                    cache_config=CC(block_size=bs, paged_evict_config=cfg),
                                    prefix="model.layers.0.self_attn.attn")
    n_evict = 0
    for si, rec in enumerate(tr["steps"]):
        md = _product_builder_md(rec["inter"], rec["seq_lens"], rec["query_lens"], bs, cfg)
        _assert_md_equal(md, rec["md"])
        # (b) the product's own host bookkeeping
        if si == 0:
            for i, L in enumerate(c["prompt_lens"]):
                bm.allocate(i, L)
        else:
            for i in range(len(c["prompt_lens"])):
                bm.append_slots(i)
        for i, r in enumerate(rec["inter"]):
            assert bm.get_block_table(i) == r["block_tables"][str(i)] or si == 0
            kv = bm.get_seq_kv_len(i)
            if si > 0 and policy.decode_trigger(r["seq_lens"][0], kv, cfg, bs):
                kv -= bs                                             # model_runner.py:540-543
            assert kv == r["seq_kv_lens"][str(i)], (si, i)
        # (c) the impl's host preparation on this metadata
        if si > 0:
            st = impl._prepare_step(md, "cpu", bs)
            assert st["trigger_rows"] == sorted(int(k) for k in rec["rmv"]), si
        for row, (rid, idx) in rec["rmv"].items():
            bm.free_prunned_blocks(int(rid[1:]), idx)
            n_evict += 1
    assert (n_evict > 0) == (c["evict_method"] != "inverse_key_l2")


def test_torch_library_shim_registers_the_fork_schemas():
    """The thin torch C++ extension (csrc/torch_shim.cpp): ops exist under pagedevict_b200 with the
    fork's schemas (csrc/torch_bindings.cpp:27-51,444-450) and, after alias_vllm_ops(), under the
    fork's own names torch.ops._C / torch.ops._C_cache_ops.  No compute without a GPU; CPU tensors
    are refused loudly (no CPU fallback)."""
    import torch
    from vllm_paged_compression_b200 import torch_ops
    ns = torch_ops.load()
    for name in ("paged_attention_v1", "paged_attention_v2", "reshape_and_cache", "zero_cache_slots",
                 "prefill_attn_topleft_causal"):
        assert hasattr(ns, name)
    s = str(ns.paged_attention_v1.default._schema)
    assert "! -> ) out" in s and "int blocksparse_head_sliding_step" in s and "Tensor? alibi_slopes" in s
    s = str(ns.reshape_and_cache.default._schema)
    assert "! -> ) key_cache" in s and "str kv_cache_dtype" in s
    torch_ops.alias_vllm_ops()
    torch_ops.alias_vllm_ops()           # idempotent
    assert str(torch.ops._C.paged_attention_v2.default._schema).startswith("_C::paged_attention_v2(")
    assert str(torch.ops._C_cache_ops.reshape_and_cache.default._schema).startswith("_C_cache_ops::reshape_and_cache(")
    k = torch.zeros(2, 2, 64, dtype=torch.bfloat16)
    with pytest.raises((RuntimeError, NotImplementedError)):
        ns.reshape_and_cache(k, k, torch.zeros(1, 2, 8, 16, 8, dtype=torch.bfloat16),
                             torch.zeros(1, 2, 64, 16, dtype=torch.bfloat16), torch.zeros(2, dtype=torch.int64),
                             "auto", 1.0, 1.0)
