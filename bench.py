#!/usr/bin/env python
"""bench.py -- decode throughput of the paged-eviction hot path on B200.

Contract (driver):  python bench.py --gpus N --steps K --warmup W [--impl reference]
  N > 1 is launched by torch.distributed.run (one rank per GPU, NCCL).

Workload (BASELINE.json configs[2], the configuration "decode tok/s at 25%/50% eviction"
is quoted on): Llama-3.1-8B attention shapes (32 layers, 32 q heads, 8 kv heads, head 128,
bf16), 64 prompts x 1024 tokens in, decode steps out, block 16, percentage pruning with
streamingLLM at prompt time and value_l2 block eviction every `evict_freq` blocks at
decode time.  One "step" = ONE decode step of the whole batch through the hot path of all
32 layers: reshape_and_cache -> [trigger / argmin / table compaction on eviction steps]
-> paged decode attention with fused value_l2 scoring, plus the host side either side of
it (scheduler bookkeeping, metadata build, layer-0 victim readback, block free).  The
model's linear layers / sampler are not part of the path: q/k/v per layer
are seeded synthetic activations.

The timed window ALWAYS contains eviction steps: with evict_freq f and block 16 a row
evicts one block every 16 f decode steps (all rows of the batch on the same step, the
prompts have one length), so after prefill an UNTIMED pre-roll phases the workload such
that a trigger step falls in the middle of the K timed steps.  Every step is bracketed by
CUDA events: `trigger_step_ms` / `plain_step_ms` are reported next to the whole-window mean.

  value     : tok/s with q/k/v already resident in HBM when the timed region starts.
  e2e       : same metric through the plugin API with HOST (pinned) q/k/v and host result
              buffers: H2D of every layer's q/k/v and D2H of every layer's attention
              output inside the timed region.
  roofline  : the dominant kernel (paged_decode_attn_score) timed alone with CUDA events,
              back to back over the 32 layers' pools (5+ GB >> 126 MB L2), algorithmic
              bytes = sum_seqs kv_len * Hkv * D * 2 (K,V) * 2 B per launch.
  legs      : (N = 1) the other two thirds of the metric on the same line: --evict-freq 4
              ("50 %") and eviction off, each with value, ms/step and the decode kernel's
              roofline fraction.
  cache_write_roofline / prefill_roofline : the prompt-time kernels of the same workload.
  cpu_baseline : the reference's own CPU kernels (oracle/_ref, kind "reference") or the
              oracle port, driven by the oracle's restatement of the reference's Python
              logic, on a bounded sample (a few of the 32 layers, scaled).
  --impl reference : that CPU arm alone, same JSON line, same `config`.

Multi-GPU (N > 1): tensor parallel over KV heads (rank r owns Hkv/N kv heads and their q
heads); the batch grows with N (64*N prompts) so per-GPU KV bytes stay fixed ("weak").
The only exchange is the per-head value_l2 partials of freshly filled blocks (DESIGN.md
"Multi-GPU"); `tp_victims_identical` all-gathers every layer's victims of the trigger steps
inside the timed window and compares them across ranks.  `strong` (N = 2 / 4: configs[3]
128 x 8192 in at TP = N; N = 8: configs[4] 70B shape 256 x 2048 in at TP 8) are fixed-batch
runs of the other BASELINE configs on the same ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from dataclasses import dataclass, replace

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

MODELS = {"8b": dict(name="Llama-3.1-8B", layers=32, Hq=32, Hkv=8, D=128),      # BASELINE configs[2], [3] (default)
          "70b": dict(name="Llama-3.1-70B", layers=80, Hq=64, Hkv=8, D=128),    # configs[4] (TP 8)
          "1b": dict(name="Llama-3.2-1B", layers=16, Hq=32, Hkv=8, D=64)}       # configs[0]
METRIC = "decode_tok_per_s_paged_eviction_hot_path"


PREFILL_BATCH_TOKENS = 64 * 1024


@dataclass(frozen=True)
class Spec:
    """One workload: attention shapes of a model, a batch of equal prompts, an eviction setting.
    `batch` is the GLOBAL batch; under tensor parallelism every rank sees all rows and 1/tp of
    the heads."""
    model: str
    layers: int
    Hq: int
    Hkv: int
    D: int
    batch: int
    prompt_len: int
    block_size: int
    evict_freq: int          # 0 = eviction off
    tp: int = 1              # tensor-parallel degree the job runs at (heads / tp per rank)
    tp_emulate: bool = False  # one GPU doing one rank's work of a tp-way job (profiling aid)

    @property
    def name(self) -> str:
        ev = f"evict-freq {self.evict_freq}" if self.evict_freq > 0 else "no eviction"
        return (f"{self.model} attention hot path, {self.layers} layers, {self.batch} prompts x "
                f"{self.prompt_len} in, decode steps, block {self.block_size}, percentage/streamingLLM prompt + "
                f"value_l2 decode, {ev}")

    def kept_prompt(self) -> int:
        from vllm_paged_compression_b200.policy import prompt_prune_layout
        cc = make_cache_config(self, 1)
        if cc.paged_evict_config is None:
            return self.prompt_len
        return prompt_prune_layout(self.prompt_len, cc.paged_evict_config, self.block_size)[0]

    def period(self) -> int:
        return self.block_size * self.evict_freq if self.evict_freq > 0 else 0


_RESULT_FD = None


def guard_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's
    'NCCL version ...' banner when NCCL_DEBUG=VERSION/WARN is set on the box), so fd 1 is
    pointed at stderr for the whole run and the result goes out through a saved copy."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--evict-freq", type=int, default=2,
                    help="README flag: 2 (labelled 25%%) or 4 (labelled 50%%); 0 = eviction off")
    ap.add_argument("--batch", type=int, default=64, help="prompts per GPU")
    ap.add_argument("--prompt-len", type=int, default=1024)
    ap.add_argument("--model", default="8b", choices=sorted(MODELS), help="attention shapes of the model")
    ap.add_argument("--layers", type=int, default=0, help="0 = the model's layer count")
    ap.add_argument("--block-size", type=int, default=16)
    ap.add_argument("--cpu-layers", type=int, default=0,
                    help="layers per step in the CPU arm (time scaled to the full layer count); 0 = all layers "
                         "at N = 1 (no extrapolation), fewer as the batch grows with N")
    ap.add_argument("--cpu-steps", type=int, default=0, help="cpu_baseline leg: 0 = derive from a ~15 s budget")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-prefill", action="store_true", help="skip the prompt-time kernel rooflines (A1, W1)")
    ap.add_argument("--no-legs", action="store_true", help="skip the evict-freq 4 / no-eviction legs (N = 1)")
    ap.add_argument("--no-strong", action="store_true", help="skip the fixed-batch configs[3] / [4] legs (N > 1)")
    ap.add_argument("--no-preroll", action="store_true", help="start timing right after prefill + warm-up")
    ap.add_argument("--no-graph", action="store_true",
                    help="eager PagedEvictEngine.step (one forward() per layer) instead of CUDA-graph replay")
    ap.add_argument("--no-pdl", action="store_true",
                    help="no programmatic dependent launch between consecutive layers' decode kernels")
    ap.add_argument("--force-pdl", action="store_true", help="PDL whatever the batch (default: ops.pdl_default rule)")
    ap.add_argument("--score-exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: how per-head block scores reach the other ranks (parallel.py)")
    ap.add_argument("--io-group", type=int, default=0,
                    help="e2e leg: layers per host<->device copy group (PinnedLayerIO); 0 = default schedule")
    ap.add_argument("--no-io-prefetch", action="store_true",
                    help="e2e leg: upload a step's inputs inside that step (layer groups) instead of one "
                         "copy a step ahead")
    ap.add_argument("--debug-knob", action="append", default=[], metavar="NAME=INT",
                    help="PC_DEBUG library only (PC_LIB_PATH=...debug.so): call pc_debug_set_NAME(INT) before the run, "
                         "e.g. decode_two_cta=1 (A/B runs; the product library has no such symbols)")
    ap.add_argument("--no-numa-bind", action="store_true",
                    help="do not bind the rank to the NUMA node of its GPU before allocating pinned memory")
    ap.add_argument("--tp-emulate", type=int, default=1,
                    help="debug: one GPU does ONE rank's work of a TP=N job (heads / N, batch x N, no "
                         "score exchange) -- for profiling the per-rank launch shape on a 1-GPU box")
    ap.add_argument("--sweep", action="store_true",
                    help="kernel microbench (BASELINE.json configs[1]) -> table on stdout")
    args = ap.parse_args()
    return args


def main_spec(args, world) -> Spec:
    m = MODELS[args.model]
    tp = world if world > 1 else max(1, args.tp_emulate)
    return Spec(model=m["name"], layers=args.layers if args.layers > 0 else m["layers"], Hq=m["Hq"], Hkv=m["Hkv"],
                D=m["D"], batch=args.batch * tp, prompt_len=args.prompt_len, block_size=args.block_size,
                evict_freq=max(args.evict_freq, 0), tp=tp, tp_emulate=(world == 1 and tp > 1))


def config_dict(spec: Spec, world: int, args) -> dict:
    """`config` of the JSON line.  Built from the workload only, so both arms (this
    implementation and --impl reference) print the same dict."""
    kept = spec.kept_prompt()
    gb = spec.batch * kept * spec.Hkv * spec.D * 4 * spec.layers / 1e9
    return {"workload": spec.name, "global_batch": spec.batch,
            "parallelism": (f"tp{spec.tp} over kv heads, batch x{spec.tp}" if spec.tp > 1 else "single GPU"),
            "evict_freq": spec.evict_freq, "block_size": spec.block_size, "prompt_len": spec.prompt_len,
            "prompt_tokens_kept": kept,
            "l2": f"inputs larger than L2 (per-layer KV pools, {gb:.1f} GB of KV read per step over all ranks)",
            "eviction_phase": ("off" if spec.evict_freq <= 0 else "start right after prefill" if args.no_preroll else
                               f"untimed pre-roll so that a trigger step (every {spec.period()} decode steps, "
                               "all rows) falls mid-window")}


# ------------------------------------------------------------------------ configuration
def make_cache_config(spec: Spec, max_num_seqs):
    from vllm_paged_compression_b200.config import CacheConfig, PagedEvictConfig
    if spec.evict_freq <= 0:
        return CacheConfig(block_size=spec.block_size, max_num_seqs=max_num_seqs)
    pe = PagedEvictConfig(cache_prune_type="percentage", evict_method=None,
                          prompt_evict_method="streamingLLM", decode_evict_method="value_l2",
                          evict_freq=spec.evict_freq, topk_blocks=-1)
    pe.topk_blocks = -1
    return CacheConfig(block_size=spec.block_size, paged_evict_config=pe, max_num_seqs=max_num_seqs)


def preroll_steps(spec: Spec, K: int, W: int, enabled: bool) -> int:
    """Untimed decode steps after prefill so that seq_len hits a multiple of the eviction
    period in the middle of the K timed steps."""
    period = spec.period()
    if not enabled or period <= 0:
        return 0
    return (-(spec.prompt_len + W + (K + 1) // 2)) % period


# ------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm))
        return out


# ------------------------------------------------------------------------ B200 arm
class Arm:
    """One engine + synthetic per-layer activations; a step = one decode step of the batch."""

    def __init__(self, spec: Spec, args, device, rank, world, host_io: bool, total_steps: int):
        import torch
        from vllm_paged_compression_b200.engine import PagedEvictEngine
        self.torch = torch
        self.spec, self.args, self.device, self.host_io = spec, args, device, host_io
        self.L = spec.layers
        self.Hq, self.Hkv, self.D = spec.Hq // spec.tp, spec.Hkv // spec.tp, spec.D
        self.B = spec.batch
        bs = spec.block_size
        cc = make_cache_config(spec, self.B)
        blocks_per_seq = -(-(spec.kept_prompt() + total_steps + 2 * bs) // bs) + 2
        ex = None
        if world > 1 and spec.evict_freq > 0:
            from vllm_paged_compression_b200.parallel import ScoreExchange
            ex = ScoreExchange(self.L, self.Hkv, mode=args.score_exchange)
        self.eng = PagedEvictEngine(self.L, self.Hq, self.Hkv, self.D, cc, self.B * blocks_per_seq + 8,
                                    device=device, score_exchange=ex, tp_rank=rank if world > 1 else 0,
                                    tp_size=world)
        for i in range(self.B):
            self.eng.add_request(f"r{i}", spec.prompt_len)
        g = torch.Generator(device=device)
        g.manual_seed(1234 + rank)
        self.gen = g
        s = self.D ** -0.5
        # decode activations: one [B, ...] set per layer, resident in HBM (value) or in
        # pinned host memory (e2e); U(-s, s) like tests/kernels/test_attention.py:146-148

        def u(*shape):
            return (torch.rand(*shape, device=device, generator=g) * 2 * s - s).to(torch.bfloat16)
        self.q = [u(self.B, self.Hq * self.D) for _ in range(self.L)]
        self.k = [u(self.B, self.Hkv * self.D) for _ in range(self.L)]
        # layer- and row-dependent V magnitude so victims differ between rows / layers.  The row
        # factor is a property of the token, not of the head shard: same seed on every rank.
        gr = torch.Generator(device=device)
        gr.manual_seed(4321)
        row = (1.0 + 0.5 * torch.rand(self.B, 1, device=device, generator=gr))
        self.v = [(u(self.B, self.Hkv * self.D).float() * row * (1 + 0.1 * l)).to(torch.bfloat16)
                  for l in range(self.L)]
        self.h2d = self.d2h = 0
        self.runner = None
        if host_io:
            from vllm_paged_compression_b200.engine import PinnedLayerIO
            # fused [B, (Hq + 2 Hkv) D] rows per layer in pinned host memory (what qkv_proj emits)
            # runner path: the step's inputs are uploaded as ONE copy, one step ahead (the next
            # step's activations land while this step computes)
            self.io = PinnedLayerIO(self.L, self.B, self.Hq, self.Hkv, self.D, device, group=args.io_group,
                                    prefetch=not (args.no_io_prefetch or args.no_graph))
            for l in range(self.L):
                self.io.qkv_host[l].copy_(torch.cat([self.q[l], self.k[l], self.v[l]], dim=1).cpu())
            self.h2d, self.d2h = self.io.h2d_bytes, self.io.d2h_bytes
            self.q = self.k = self.v = None

    def prefill(self):
        """Untimed prefill batches (chunked by request so activations stay small)."""
        torch = self.torch
        s = self.D ** -0.5
        g = self.gen

        def qkv(layer, n_tok):
            u = lambda w: (torch.rand(n_tok, w, device=self.device, generator=g) * 2 * s - s).to(torch.bfloat16)
            v = (u(self.Hkv * self.D).float() * (1 + 0.1 * layer)).to(torch.bfloat16)
            return u(self.Hq * self.D), u(self.Hkv * self.D), v
        # one prefill batch of up to 64 Ki tokens: the fork schedules prompts up to
        # max_num_batched_tokens (>= max_model_len without chunked prefill, vllm/config.py), so the 64 x
        # 1024-token prompts of the headline config are ONE prefill step there too
        self.eng.max_num_batched_tokens = max(PREFILL_BATCH_TOKENS, self.spec.prompt_len)
        while self.eng.waiting:
            so = self.eng.step(qkv, keep_outputs=False)
            assert so.is_prompt
        torch.cuda.synchronize()

    def make_runner(self, total_steps: int):
        """Decode steps as CUDA-graph replays (graph_runner.py) unless --no-graph."""
        if self.args.no_graph:
            self.runner = None
            return
        from vllm_paged_compression_b200.graph_runner import DecodeGraphRunner
        a = self.args
        kv_now = max(self.eng.block_manager.get_seq_kv_len(s.seq_id) for s in self.eng.running)
        bound = kv_now + total_steps + 2 * self.spec.block_size + 64
        pdl = False if a.no_pdl else (True if a.force_pdl else None)
        if self.host_io:
            self.runner = DecodeGraphRunner(self.eng, None, io=self.io, max_kv_len=bound, pdl=pdl)
        else:
            self.runner = DecodeGraphRunner(self.eng, list(zip(self.q, self.k, self.v)), max_kv_len=bound, pdl=pdl)

    def step(self) -> int:
        """One decode step of the whole batch; -> blocks evicted (layer-0 feedback)."""
        if self.runner is not None:
            # host_io: H2D of every layer's fused qkv rows and D2H of every layer's attention
            # output are nodes of the replayed graph; step() returns once out_host is valid
            return len(self.runner.step())
        if not self.host_io:
            so = self.eng.step(lambda l, n: (self.q[l], self.k[l], self.v[l]), keep_outputs=False)
            return len(so.rmv)
        self.io.begin_step()
        so = self.eng.step(self.io.qkv, keep_outputs=False, out_fn=self.io.put_out,
                           out_buf_fn=self.io.out_buf)
        self.io.end_step()                                  # the caller owns the host result now
        return len(so.rmv)


def timed_steps(torch, dist, arm, K, W, P, world):
    """P pre-roll + W warm-up steps (untimed), then K timed steps: barrier + synchronize
    either side, one CUDA event per step boundary, max over ranks.
    -> dict(sec, evictions, launches, trigger_step_ms, plain_step_ms, n_trigger_steps)"""
    from vllm_paged_compression_b200 import ops
    for _ in range(P + W):
        arm.step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if arm.runner is not None:
        arm.runner.log_victims = world > 1
        arm.runner.victim_log = []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    n0 = ops.launch_count
    per_step = []
    torch.cuda.profiler.start()          # ncu --profile-from-start off sees the timed region only
    ev[0].record()
    for i in range(K):
        per_step.append(arm.step())
        ev[i + 1].record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sec = ev[0].elapsed_time(ev[K]) / 1e3
    ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(K)]
    trig = [m for m, e in zip(ms, per_step) if e > 0]
    plain = [m for m, e in zip(ms, per_step) if e == 0]
    if world > 1:
        t = torch.tensor([sec], device=arm.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return dict(sec=sec, evictions=sum(per_step), launches=ops.launch_count - n0,
                trigger_step_ms=statistics.mean(trig) if trig else None,
                plain_step_ms=statistics.mean(plain) if plain else None, n_trigger_steps=len(trig))


def victims_identical(torch, dist, arm, world):
    """All-gather every layer's victims of the trigger steps of the timed window and compare
    them across ranks (tensor parallel: every rank must evict the same blocks)."""
    if world <= 1 or arm.runner is None or arm.spec.evict_freq <= 0:
        return None, 0
    log_ = arm.runner.victim_log
    n = torch.tensor([len(log_)], device=arm.device, dtype=torch.int64)
    dist.all_reduce(n, op=dist.ReduceOp.MIN)
    if int(n.item()) == 0 or len(log_) == 0:
        return None, 0
    mine = torch.stack(log_[:int(n.item())]).contiguous()        # [T, L, B]
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine)
    same = all(bool(torch.equal(allv[0], x)) for x in allv[1:])
    return same, int((mine[:, 0] >= 0).sum().item())


def roofline_leg(torch, dist, arm, args, world, reps=20):
    """The decode-attention + scoring kernel alone, back to back over every layer's pool
    (distinct memory per launch, total >> L2), on the state the timed region ended in."""
    from vllm_paged_compression_b200 import ops
    from vllm_paged_compression_b200.backend import PagedAttention
    eng = arm.eng
    md = eng.build_metadata(False, list(eng.running))
    kv_lens_host = list(md.seq_kv_lens or md.seq_lens)
    kv_lens = md.seq_kv_lens_tensor if md.seq_kv_lens_tensor is not None else md.seq_lens_tensor
    max_kv = max(kv_lens_host)
    caches = [PagedAttention.split_kv_cache(kv, arm.Hkv, arm.D) for kv in eng.kv_caches]
    outs = torch.empty(arm.B, arm.Hq, arm.D, dtype=torch.bfloat16, device=arm.device)
    qs = [q.view(arm.B, arm.Hq, arm.D) for q in arm.q]
    ks = [k.view(arm.B, arm.Hkv, arm.D) for k in arm.k]
    vs = [v.view(arm.B, arm.Hkv, arm.D) for v in arm.v]
    scale = arm.D ** -0.5
    pdl = (not args.no_pdl) and (args.force_pdl or ops.pdl_default(arm.B))
    # the launch the product makes (pc_decode_layer on a non-trigger step): the step's K/V row
    # written by the attention kernel itself, programmatic dependent launch behind the
    # previous layer's kernel (--no-pdl: plain stream order)
    ds = ops.make_decode_step(md.slot_mapping, md.block_tables, kv_lens, md.seq_lens_tensor, arm.B, arm.Hq,
                              arm.Hkv, arm.D, arm.spec.block_size, max_kv, scale, qs[0].stride(0), ks[0].stride(0),
                              vs[0].stride(0), outs.stride(0), caches[0][1].stride(0), fuse_cache_write=True,
                              pdl=pdl)

    def run_all():
        for l in range(arm.L):
            ops.decode_layer(ds, outs, qs[l], ks[l], vs[l], caches[l][0].data_ptr(), caches[l][1].data_ptr(),
                             None, None)
    for _ in range(3):
        run_all()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run_all()
    e1.record()
    torch.cuda.synchronize()
    per_launch_s = e0.elapsed_time(e1) / 1e3 / (reps * arm.L)
    if world > 1:
        t = torch.tensor([per_launch_s], device=arm.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per_launch_s = float(t.item())
    alg_bytes = sum(kv_lens_host) * arm.Hkv * arm.D * 2 * 2
    peak, peak_src = load_peaks()
    achieved = alg_bytes / per_launch_s / 1e9
    return {"bound": "hbm", "kernel": "paged_decode_attn_score_kernel", "achieved": achieved,
            "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
            "launch": "pc_decode_layer as in the step: fused cache write" + (", PDL" if pdl else "") +
                      ", back to back over the layers' pools, CUDA events",
            "us_per_launch": per_launch_s * 1e6, "mean_kv_len": sum(kv_lens_host) / len(kv_lens_host)}


def load_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_b200(spec: Spec, args, torch, dist, device, rank, world, K, W, e2e: bool, sampler_rank0=False):
    """One workload on this implementation: value (resident inputs) [+ e2e (host buffers)].
    -> dict with the numbers of the JSON line."""
    P = preroll_steps(spec, K, W, not args.no_preroll)
    total = P + W + K
    sampler = ClockSampler(device.index) if sampler_rank0 else None   # nvidia-smi needs ~0.3 s to start sampling
    arm = Arm(spec, args, device, rank, world, host_io=False, total_steps=total)
    arm.prefill()
    arm.make_runner(total)
    t = timed_steps(torch, dist, arm, K, W, P, world)
    if arm.runner is not None:
        arm.runner.sync_host()          # the roofline leg reads lengths from the host objects
    same, n_vict = victims_identical(torch, dist, arm, world)
    roof = roofline_leg(torch, dist, arm, args, world)
    clocks = sampler.stop() if sampler else None        # sampled over the timed steps + the roofline leg
    B = arm.B
    recaptures = arm.runner.recaptures if arm.runner is not None else None
    del arm
    torch.cuda.empty_cache()
    out = {"value": B * K / t["sec"], "ms_per_step": t["sec"] / K * 1e3, "timing": t, "roofline": roof,
           "clocks": clocks, "tp_victims_identical": same, "victims_compared": n_vict, "preroll": P,
           "graph_recaptures": recaptures, "B": B, "e2e": None}
    if e2e:
        arm = Arm(spec, args, device, rank, world, host_io=True, total_steps=total)
        arm.prefill()
        arm.make_runner(total)
        te = timed_steps(torch, dist, arm, K, W, P, world)
        out["e2e"] = {"value": B * K / te["sec"], "unit": "tok/s", "h2d_bytes_per_step": arm.h2d,
                      "d2h_bytes_per_step": arm.d2h, "ms_per_step": te["sec"] / K * 1e3,
                      "trigger_step_ms": te["trigger_step_ms"], "plain_step_ms": te["plain_step_ms"],
                      "evictions_in_timed_region": te["evictions"],
                      "io": arm.io.describe() if hasattr(arm.io, "describe") else None}
        del arm
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------ CPU arm
class CpuArm:
    """The reference path on host cores: the oracle's restatement of the reference's
    Python logic (oracle/backend.py, oracle/policy.py) over the reference's OWN compiled
    CPU kernels (oracle/_ref: csrc/cpu/attention.cpp, cache.cpp) when available
    (kind "reference"), else the torch restatement (kind "port").  Bounded sample:
    `cpu_layers` of the model's layers per step, full batch."""

    def __init__(self, spec: Spec, cpu_layers: int, total_steps: int):
        import numpy as np
        import torch
        from oracle import policy as opol
        from oracle import ref_cpu
        from oracle.backend import OracleLayer
        from oracle.policy import Policy
        self.torch, self.np, self.opol = torch, np, opol
        self.spec = spec
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.Lc = max(1, min(cpu_layers, spec.layers))
        self.Hq, self.Hkv, self.D = spec.Hq, spec.Hkv, spec.D
        self.B = spec.batch
        bs = self.bs = spec.block_size
        self.kind = "reference" if (bs == 16 and ref_cpu.available()) else "port"
        native = ref_cpu.RefCpuOps if self.kind == "reference" else None
        self.pol = Policy(block_size=bs, cache_prune_type="percentage",
                          evict_freq=max(spec.evict_freq, 1), decode_evict_method="value_l2", topk_blocks=-1)
        self.evict = spec.evict_freq > 0
        if not self.evict:
            # eviction off: nothing is ever pruned or triggered
            self.pol = Policy(block_size=bs, cache_prune_type="budget", cache_budget=1 << 30,
                              decode_evict_method="streamingLLM")
        self.layers = [OracleLayer(self.Hq, self.D, self.D ** -0.5, self.Hkv, self.pol, native_ops=native)
                       for _ in range(self.Lc)]
        P = spec.prompt_len
        nblk, kept = opol.num_required_blocks_after_prune(P, self.pol)
        per_seq = nblk + total_steps // bs + 4
        self.num_blocks = self.B * per_seq + 8
        self.kv = [torch.zeros(2, self.num_blocks, bs * self.Hkv * self.D, dtype=torch.bfloat16)
                   for _ in range(self.Lc)]
        self.free = list(range(self.num_blocks - 1, -1, -1))
        self.tables = [[self.free.pop() for _ in range(nblk)] for _ in range(self.B)]
        self.seq_len = [P] * self.B
        self.kv_len = [kept] * self.B
        self.req = [f"r{i}" for i in range(self.B)]
        g = torch.Generator().manual_seed(1234)
        s = self.D ** -0.5
        u = lambda *sh: (torch.rand(*sh, generator=g) * 2 * s - s).to(torch.bfloat16)
        # prompt KV straight into the cache + prompt block scores (no prefill attention: the
        # metric is decode throughput); chunked so host memory stays small
        for l, lay in enumerate(self.layers):
            for c0 in range(0, self.B, 8):
                rows = range(c0, min(c0 + 8, self.B))
                n = len(rows) * P
                lay.prefill(torch.empty(0), u(n, self.Hkv * self.D),
                            (u(n, self.Hkv * self.D).float() * (1 + 0.1 * l)).to(torch.bfloat16),
                            self.kv[l], [P] * len(rows), [self.tables[r] for r in rows],
                            [self.req[r] for r in rows], compute_attention=False)
        self.q = [u(self.B, self.Hq * self.D) for _ in range(self.Lc)]
        self.k = [u(self.B, self.Hkv * self.D) for _ in range(self.Lc)]
        self.v = [u(self.B, self.Hkv * self.D) for _ in range(self.Lc)]

    def step(self):
        opol, bs = self.opol, self.bs
        kv_pass = []
        for r in range(self.B):                       # scheduler + model-runner side
            if self.kv_len[r] % bs == 0:
                self.tables[r].append(self.free.pop())
            self.seq_len[r] += 1
            self.kv_len[r] += 1
            trig = self.evict and opol.decode_trigger(self.seq_len[r], len(self.tables[r]) * bs, self.pol)
            kv_pass.append(self.kv_len[r] - (bs if trig else 0))
        rmv0 = {}
        for l, lay in enumerate(self.layers):
            _, rmv, _ = lay.decode(self.q[l], self.k[l], self.v[l], self.kv[l], self.seq_len, kv_pass,
                                   self.tables, self.req)
            if l == 0:
                rmv0 = rmv
        for r, (_, idx) in rmv0.items():              # engine feedback: free the victim
            self.free.append(self.tables[r].pop(idx))
            self.kv_len[r] -= bs
        return len(rmv0)

    def run(self, K, W, P=0):
        for _ in range(P + W):
            self.step()
        ev = 0
        t0 = time.perf_counter()
        for _ in range(K):
            ev += self.step()
        dt = time.perf_counter() - t0
        scale = self.spec.layers / self.Lc
        sec_per_step = dt / K * scale
        sample = (f"{K} decode steps x {self.Lc} of {self.spec.layers} layers x full batch {self.B}"
                  + (f" (time scaled x{scale:g} to all layers)" if self.Lc != self.spec.layers else " (all layers)")
                  + f", {P} pre-roll + {W} warm-up steps, {ev} evictions in the timed steps")
        return dict(value=self.B / sec_per_step, ms_per_step=sec_per_step * 1e3, kind=self.kind,
                    cores=self.cores, sample=sample, evictions=ev)


def cpu_layers_for(args, spec: Spec) -> int:
    if args.cpu_layers > 0:
        return args.cpu_layers
    # all layers at the N = 1 batch (no extrapolation); the CPU step cost grows with the batch
    return max(2, min(spec.layers, spec.layers * 64 // max(spec.batch, 64)))


def run_reference(args, rank, world):
    if rank != 0:
        return
    spec = main_spec(args, world)
    K, W = args.steps, max(args.warmup, 1)
    P = preroll_steps(spec, K, W, not args.no_preroll)
    arm = CpuArm(spec, cpu_layers_for(args, spec), P + W + K)
    r = arm.run(K, W, P)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "tok/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "config": config_dict(spec, world, args),
        "evictions_in_timed_region": r["evictions"],
        "cpu_baseline": {"value": r["value"], "unit": "tok/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------ extra legs
def leg_summary(r):
    t = r["timing"]
    return {"value": r["value"], "unit": "tok/s", "ms_per_step": r["ms_per_step"],
            "trigger_step_ms": t["trigger_step_ms"], "plain_step_ms": t["plain_step_ms"],
            "evictions_in_timed_region": t["evictions"], "preroll_steps": r["preroll"],
            "decode_frac": r["roofline"]["frac"], "decode_us_per_launch": r["roofline"]["us_per_launch"],
            "decode_gbs": r["roofline"]["achieved"], "mean_kv_len": r["roofline"]["mean_kv_len"],
            "tp_victims_identical": r["tp_victims_identical"]}


def strong_spec(world) -> Spec:
    """Fixed-batch BASELINE configs run on the ranks of this job: configs[3] (8B long context,
    128 x 8192 in, 50 % = evict-freq 4) at TP 2 / 4, configs[4] (70B, 256 x 2048 in) at TP 8."""
    if world == 8:
        m = MODELS["70b"]
        return Spec(m["name"], m["layers"], m["Hq"], m["Hkv"], m["D"], 256, 2048, 16, 4, tp=8)
    m = MODELS["8b"]
    return Spec(m["name"], m["layers"], m["Hq"], m["Hkv"], m["D"], 128, 8192, 16, 4, tp=world)


# ------------------------------------------------------------------------ main
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if not args.sweep:
        guard_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.gpus > 1 and world == 1:
        raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")

    numa = None
    if not args.no_numa_bind:
        from vllm_paged_compression_b200 import hostmem
        numa = hostmem.bind_to_gpu_numa_node(local_rank)      # before CUDA / pinned allocations
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    spec = main_spec(args, world)
    if spec.Hkv % spec.tp != 0:
        raise SystemExit(f"tensor parallel size {spec.tp} must divide {spec.Hkv} kv heads")
    from vllm_paged_compression_b200 import _lib
    lib = _lib.load()                                  # fail loudly if the CUDA library is missing
    for kv in args.debug_knob:                         # A/B runs with the PC_DEBUG library
        import ctypes
        name, val = kv.split("=")
        fn = getattr(lib, "pc_debug_set_" + name)      # AttributeError with the product library
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(int(val))

    if args.sweep:
        from benchtools import microbench
        microbench.sweep(device)
        return

    K, W = args.steps, args.warmup
    r = run_b200(spec, args, torch, dist, device, rank, world, K, W, e2e=not args.no_e2e,
                 sampler_rank0=(rank == 0))
    t = r["timing"]
    line = {
        "metric": METRIC, "value": r["value"], "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": config_dict(spec, world, args),
        "evictions_in_timed_region": t["evictions"], "trigger_steps_in_timed_region": t["n_trigger_steps"],
        "trigger_step_ms": t["trigger_step_ms"], "plain_step_ms": t["plain_step_ms"],
        "preroll_steps": r["preroll"], "tp_victims_identical": r["tp_victims_identical"],
        "tp_victims_compared": r["victims_compared"], "graph_recaptures": r["graph_recaptures"],
        "impl_details": {"step_mode": "eager per-layer forward()" if args.no_graph else
                         "CUDA-graph replay of the whole-layer-stack step (graph_runner.py)",
                         "score_exchange": args.score_exchange if world > 1 else None,
                         "numa_bind": numa},
        "clocks": r["clocks"],
        "e2e": r["e2e"],
        "gpu_launches": t["launches"],
        "roofline": r["roofline"],
    }
    if t["evictions"] == 0 and spec.evict_freq > 0:
        line["warning"] = "eviction enabled but no eviction inside the timed window (steps < eviction period?)"
    # ncu-measured DRAM traffic of the same launch shape, when a summary is committed
    tpath = os.path.join(REPO, "profiles", "decode_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        line["roofline"]["traffic"] = tj.get("dram_bytes_per_launch")
        line["roofline"]["traffic_source"] = tj.get("source")
        line["roofline"]["traffic_measured_in_run"] = False
    if world == 1 and not args.no_legs and spec.evict_freq > 0 and not spec.tp_emulate:
        # the rest of the metric ("25 % / 50 % eviction vs no eviction") on the same line
        legs = {}
        for name, ef in (("evict_freq_4", 4), ("no_eviction", 0)):
            if ef == spec.evict_freq:
                continue
            log(f"leg {name}")
            legs[name] = leg_summary(run_b200(replace(spec, evict_freq=ef), args, torch, dist, device, rank, world,
                                              K, W, e2e=False))
        legs[f"evict_freq_{spec.evict_freq}"] = "this line"
        line["legs"] = legs
    if world > 1 and not args.no_strong and world in (2, 4, 8) and args.model == "8b":
        ss = strong_spec(world)
        log(f"strong leg: {ss.name} at TP {world}")
        sr = run_b200(ss, args, torch, dist, device, rank, world, min(K, 24), min(W, 3), e2e=False)
        line["strong"] = dict(leg_summary(sr), config=config_dict(ss, world, args), scaling="strong",
                              per_rank_shape=f"B {ss.batch}, Hq {ss.Hq // ss.tp} / Hkv {ss.Hkv // ss.tp}, "
                                             f"{ss.layers} layers")
    if rank == 0 and world == 1 and not args.no_prefill:
        from benchtools import microbench_cache_write, microbench_prefill
        peaks = {}
        ppath = os.path.join(REPO, "MEASURED_PEAKS.json")
        if os.path.exists(ppath):
            with open(ppath) as f:
                peaks = json.load(f)
        # prompt-time kernels of the same workload.  A1: one layer's launch for one prefill batch
        # (as the engine batches them: up to 64 Ki tokens), tcgen05 kernel, tensor-bound
        pre_reqs = max(1, min(spec.batch, PREFILL_BATCH_TOKENS // spec.prompt_len))
        row, _ = microbench_prefill.run(device, pre_reqs, spec.prompt_len, spec.Hq, spec.Hkv, spec.D,
                                        spec.evict_freq, 20)
        tpeak = float(peaks.get("bf16_tflops", 2250.0))
        line["prefill_roofline"] = {
            "bound": "tensor", "kernel": row["kernel"] + " (tcgen05.mma, TMEM accumulators, TMA loads)",
            "achieved": row["tflops"], "peak": tpeak, "unit": "TFLOP/s", "frac": row["tflops"] / tpeak,
            "peak_source": "measured burst (MEASURED_PEAKS.json bf16_tflops)" if peaks else "nominal dense bf16",
            "us_per_launch": row["us"], "launch": f"{pre_reqs} prompts x {spec.prompt_len} q, {row['kv_kept']} kept kv, "
            "top-left causal, one layer", "flops_per_launch": row["tflops"] * 1e12 * row["us"] * 1e-6}
        # W1 + R1 + S3: one layer's prompt prune + cache write + block scoring, HBM-bound
        cw = microbench_cache_write.run(device, pre_reqs, spec.prompt_len, spec.Hkv, spec.D, spec.block_size,
                                        spec.evict_freq)
        hpeak, hsrc = load_peaks()
        line["cache_write_roofline"] = {
            "bound": "hbm", "kernel": "cache_write_stream_kernel (prune gather + reshape_and_cache + block scores, persistent TMA pipeline)",
            "achieved": cw["gbs"], "peak": hpeak, "unit": "GB/s", "frac": cw["gbs"] / hpeak, "peak_source": hsrc,
            "traffic": None, "algorithmic_bytes_per_launch": cw["bytes"], "us_per_launch": cw["us"],
            "launch": cw["launch"],
            # the floor at this size: a plain device copy of the same bytes, timed the same way
            "copy_same_bytes_us": cw["copy_same_bytes_us"], "frac_of_same_size_copy": cw["copy_same_bytes_us"] / cw["us"]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        carm = CpuArm(spec, min(4, spec.layers), 64)
        t0 = time.perf_counter()
        carm.step()
        one = time.perf_counter() - t0
        ck = args.cpu_steps or max(3, min(32, int(15.0 / max(one, 1e-3))))
        line["cpu_baseline"] = {k: v for k, v in carm.run(ck, 1).items() if k not in ("ms_per_step", "evictions")}
        line["cpu_baseline"]["unit"] = "tok/s"
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
