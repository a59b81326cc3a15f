"""Launched by torch.distributed.run on >= 2 GPUs (tests/test_gpu_tp.py, or by hand):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tests/run_tp_parity.py [p2p|nccl]
Every rank drives the product engine with its shard of the heads (tensor parallel over KV
heads, score exchange as given) and ALSO a single-rank engine over all heads on the same
GPU; per step and layer the victims must be identical (bit-exact decisions on every rank)
and the attention outputs of the shard must match the corresponding head columns."""
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

from tests.layer_harness import layer_inputs, make_configs  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "p2p"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from vllm_paged_compression_b200.engine import PagedEvictEngine
    from vllm_paged_compression_b200.parallel import ScoreExchange, shard_heads
    Hq, Hkv, D, bs, L, nb = 16, 4, 128, 16, 3, 512
    c = dict(Hq=Hq, Hkv=Hkv, D=D, bs=bs, prune="percentage", evict_freq=2, decode="value_l2",
             prompt_lens=[96, 130, 257, 64], steps=70, seed=11)
    qh, kvh = shard_heads(Hq, Hkv, rank, world)
    cc_tp, _ = make_configs(c)
    cc_1, _ = make_configs(c)
    ex = ScoreExchange(L, len(kvh), mode=mode)
    eng_tp = PagedEvictEngine(L, len(qh), len(kvh), D, cc_tp, nb, device=dev, score_exchange=ex,
                              tp_rank=rank, tp_size=world)
    eng_1 = PagedEvictEngine(L, Hq, Hkv, D, cc_1, nb, device=dev)
    for i, P in enumerate(c["prompt_lens"]):
        eng_tp.add_request(f"r{i}", P)
        eng_1.add_request(f"r{i}", P)
    qcols = torch.cat([torch.arange(h * D, (h + 1) * D) for h in qh])
    kcols = torch.cat([torch.arange(h * D, (h + 1) * D) for h in kvh])
    evictions, max_err = 0, 0.0
    for step in range(c["steps"]):
        full = {}

        def qkv_full(layer, n):
            q, k, v = layer_inputs(c["seed"], layer, step, n, Hq, Hkv, D)
            full[layer] = (q, k, v)
            return q.to(dev), k.to(dev), v.to(dev)

        def qkv_tp(layer, n):
            q, k, v = full[layer]
            return (q[:, qcols].contiguous().to(dev), k[:, kcols].contiguous().to(dev),
                    v[:, kcols].contiguous().to(dev))
        so1 = eng_1.step(qkv_full)
        v1 = [dict(im.get_rmv_block_idx()) for im in eng_1.impls]
        sot = eng_tp.step(qkv_tp)
        vt = [dict(im.get_rmv_block_idx()) for im in eng_tp.impls]
        assert v1 == vt, f"rank {rank} step {step}: TP victims {vt} != single-rank {v1}"
        assert so1.block_tables == sot.block_tables and so1.seq_kv_lens == sot.seq_kv_lens
        for layer in range(L):
            a = so1.outputs[layer][:, qcols.to(dev)].float()
            b = sot.outputs[layer].float()
            max_err = max(max_err, (a - b).abs().max().item())
        evictions += len(sot.rmv)
    assert max_err <= 2e-2, max_err
    assert evictions > 0
    # every rank's pools hold every head's partials, identical to the single-rank pools
    torch.cuda.synchronize()
    dist.barrier()
    for layer in range(L):
        p1 = eng_1.impls[layer]._score_pool
        pt = eng_tp.impls[layer]._score_pool
        live = sorted({b for s in eng_1.running for b in eng_1.block_manager.get_block_table(s.seq_id)[:-1]})
        idx = torch.tensor(live, device=dev)
        assert torch.equal(p1[idx], pt[idx]), f"rank {rank} layer {layer}: exchanged pool differs"
    dist.barrier()
    if rank == 0:
        print(f"TP{world} parity OK mode={mode} evictions={evictions} max_err={max_err:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
