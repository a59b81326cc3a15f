"""world_size-2 gloo tests (CPU) of the tensor-parallel host logic: head sharding, the
"nccl"-mode score exchange (all-gather of freshly scored pool rows) and that victims chosen
from exchanged per-head partials are bit-identical to the single-rank oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import policy, scoring
from oracle.policy import Policy


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from vllm_paged_compression_b200.parallel import ScoreExchange, shard_heads
        L, nb, Hkv, D, bs = 3, 40, 4, 64, 16
        qh, kvh = shard_heads(16, Hkv, rank, world)
        hl = len(kvh)
        assert list(kvh) == list(range(rank * hl, (rank + 1) * hl)) and len(qh) == 16 // world
        ex = ScoreExchange(L, hl, mode="nccl")
        pools = [ex.alloc_pool(nb, Hkv, torch.device("cpu"))[0] for _ in range(L)]
        # every rank sees the same synthetic V blocks (as TP ranks see the same tokens) but scores its heads only
        rs = np.random.RandomState(7)
        v = rs.standard_normal((L, nb, bs, Hkv, D)).astype(np.float32)
        v = torch.from_numpy(v).to(torch.bfloat16).float().numpy()
        full = np.stack([[scoring.per_head_block_score(v[l, b]) for b in range(nb)] for l in range(L)])
        fresh_steps = [[1, 5, 9, 20], [2, 3], [39], list(range(10, 19))]
        done = set()
        for ids in fresh_steps:
            for l in range(L):
                for b in ids:       # what the scoring kernels do: local columns only
                    pools[l][b, rank * hl:(rank + 1) * hl] = torch.from_numpy(
                        scoring.per_head_block_score(v[l, b][:, rank * hl:(rank + 1) * hl]))
            ex.publish(ids)
            done.update(ids)
            for l in range(L):
                got = pools[l].numpy()
                for b in range(nb):
                    want = full[l, b] if b in done else np.zeros(Hkv, np.float32)
                    assert np.array_equal(got[b].view(np.int32), want.view(np.int32)), (l, b)
        # victims from the exchanged pool == single-rank oracle, bit for bit
        pol = Policy(block_size=bs, cache_prune_type="budget", cache_budget=128, decode_evict_method="value_l2")
        table = [1, 5, 9, 20, 2, 3, 39, 10, 11, 12]
        lo, hi = policy.victim_candidates(len(table), pol)
        for l in range(L):
            sc_tp = scoring.combine_heads(pools[l].numpy()[table[:-1]])
            sc_1 = scoring.combine_heads(full[l][table[:-1]])
            assert np.array_equal(sc_tp.view(np.int32), sc_1.view(np.int32))
            assert policy.argmin_lowest_index(sc_tp, lo, hi) == policy.argmin_lowest_index(sc_1, lo, hi)
        ex.publish([])              # nothing scored: must not communicate
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))


def test_score_exchange_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_shard_heads_errors():
    from vllm_paged_compression_b200.parallel import shard_heads
    assert shard_heads(32, 8, 3, 8) == (range(12, 16), range(3, 4))
    with pytest.raises(ValueError):
        shard_heads(32, 8, 0, 3)
