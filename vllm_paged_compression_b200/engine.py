"""Step driver for the paged-eviction hot path: the callers either side of
`TorchCUDASDPABackendImpl.forward`, reduced to what the path needs.

One `step()` = the slice of `LLMEngine.step` (vllm/engine/llm_engine.py:1280-1420) that
touches this path:
  scheduler   : prefill XOR decode batch, block tables + seq_kv_lens per sequence
                (vllm/core/scheduler.py:1300-1432; chunked prefill asserted off :1306-1307)
  model runner: context/seq/query lens and the decode-time `seq_kv_lens -= block_size`
                adjust (vllm/worker/model_runner.py:500-546), metadata builder
  model       : num_layers x impl.forward(q, k, v, kv_cache[layer], metadata)
  feedback    : layer 0's get_rmv_block_idx() -> free_prunned_blocks
                (model_runner.py:1797-1802, llm_engine.py:1179-1193, scheduler.py:1879-1890)
The model's linear layers / sampler are NOT here: q/k/v come from the caller (tests and
bench.py feed seeded synthetic activations), which is what the reference's attention
backend sees too.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import random

import torch

from .backend import (TorchCUDASDPABackend, TorchCUDASDPABackendImpl, TorchCUDASDPAMetadata,
                      TorchCUDASDPAMetadataBuilder)
from .block_manager import SelfAttnBlockSpaceManager
from .config import CacheConfig
from .policy import decode_trigger


@dataclass
class _Seq:
    request_id: str
    seq_id: int
    prompt_len: int
    seq_len: int = 0          # logical tokens so far (prompt + generated)
    computed: int = 0         # tokens whose KV has been computed
    prefilled: bool = False


@dataclass
class InterDataForSeqGroup:
    """The attributes TorchCUDASDPAMetadataBuilder._add_seq_group reads from
    ModelInputForGPUBuilder.InterDataForSeqGroup (model_runner.py:255-300)."""
    request_id: str
    is_prompt: bool
    seq_ids: List[int]
    block_tables: Optional[Dict[int, List[int]]]
    input_tokens: List[List[int]]
    orig_seq_lens: List[int]
    seq_lens: List[int]
    query_lens: List[int]
    context_lens: List[int]
    seq_kv_lens: Optional[Dict[int, int]] = None
    prefix_cache_hit: bool = False
    curr_sliding_window_blocks: List[int] = field(default_factory=lambda: [0])


class _Runner:
    def __init__(self, device, cache_config, pin_memory=True):
        self.device, self.cache_config, self.pin_memory = device, cache_config, pin_memory


class _InputBuilder:
    def __init__(self, runner, block_size, inter_data_list):
        self.runner = runner
        self.sliding_window = None
        self.block_size = block_size
        self.chunked_prefill_enabled = False
        self.inter_data_list = inter_data_list


@dataclass
class StepOutput:
    is_prompt: bool
    request_ids: List[str]
    seq_lens: List[int]
    seq_kv_lens: List[int]
    block_tables: List[List[int]]
    metadata: TorchCUDASDPAMetadata
    outputs: List[torch.Tensor]                       # per layer [T, Hq*D]
    rmv: Dict[int, Tuple[str, int]]                   # layer-0 decision, applied to the host tables


QKVFn = Callable[[int, int], Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]


class PagedEvictEngine:
    def __init__(self, num_layers: int, num_heads: int, num_kv_heads: int, head_size: int,
                 cache_config: CacheConfig, num_gpu_blocks: int, device="cuda:0",
                 dtype=torch.bfloat16, max_num_batched_tokens: int = 1 << 30, score_exchange=None,
                 **impl_kwargs) -> None:
        """num_heads / num_kv_heads are the LOCAL (per tensor-parallel rank) head counts;
        tp_rank / tp_size go to the impls through impl_kwargs; score_exchange
        (parallel.ScoreExchange) is required for value_l2 eviction when tp_size > 1."""
        self.device = torch.device(device)
        self.cache_config = cache_config
        self.block_size = cache_config.block_size
        self.num_layers, self.num_heads = num_layers, num_heads
        self.num_kv_heads, self.head_size = num_kv_heads, head_size
        scale = head_size ** -0.5
        self.score_exchange = score_exchange
        # ONE host RNG for every layer, as in HEAD where all layers share Python's process-wide
        # `random` (seeded 1234, torch_cuda_sdpa.py:796); an instance keeps engines independent
        impl_kwargs.setdefault("rng", random.Random(1234))
        self.rng = impl_kwargs["rng"]
        self.impls = [TorchCUDASDPABackendImpl(num_heads, head_size, scale, num_kv_heads, None, None,
                                               "auto", cache_config=cache_config,
                                               prefix=f"model.layers.{i}.self_attn.attn",
                                               score_exchange=score_exchange, **impl_kwargs)
                      for i in range(num_layers)]
        shape = TorchCUDASDPABackend.get_kv_cache_shape(num_gpu_blocks, self.block_size, num_kv_heads,
                                                        head_size)
        # CacheEngine._allocate_kv_cache: zero-filled [2, nb, bs*Hkv*D] per layer (cache_engine.py:68-88)
        self.kv_caches = [torch.zeros(shape, dtype=dtype, device=self.device) for _ in range(num_layers)]
        self.block_manager = SelfAttnBlockSpaceManager(self.block_size, num_gpu_blocks,
                                                       cache_config.paged_evict_config)
        self.max_num_batched_tokens = max_num_batched_tokens
        self.waiting: List[_Seq] = []
        self.running: List[_Seq] = []
        self._next_seq_id = 0
        self._runner = _Runner(self.device, cache_config)

    # ------------------------------------------------------------------ scheduler side
    def add_request(self, request_id: str, prompt_len: int) -> None:
        self.waiting.append(_Seq(request_id, self._next_seq_id, prompt_len, seq_len=prompt_len))
        self._next_seq_id += 1

    def finish_request(self, request_id: str) -> None:
        for s in [s for s in self.running if s.request_id == request_id]:
            self.block_manager.free(s.seq_id)
            self.running.remove(s)

    def has_unfinished(self) -> bool:
        return bool(self.waiting or self.running)

    def _schedule(self) -> Tuple[bool, List[_Seq]]:
        """Prefill batch if anything is waiting and fits, else one decode step for all
        running sequences (scheduler.py _schedule_default)."""
        batch: List[_Seq] = []
        tokens = 0
        while self.waiting:
            s = self.waiting[0]
            if tokens + s.prompt_len > self.max_num_batched_tokens and batch:
                break
            if not self.block_manager.can_allocate(s.prompt_len):
                break
            self.block_manager.allocate(s.seq_id, s.prompt_len)
            batch.append(self.waiting.pop(0))
            tokens += s.prompt_len
        if batch:
            return True, batch
        if not self.block_manager.can_append_slots(len(self.running)):
            raise RuntimeError("out of KV blocks (preemption is outside the hot path)")
        for s in self.running:
            s.seq_len += 1
            self.block_manager.append_slots(s.seq_id)
        return False, list(self.running)

    # --------------------------------------------------------------- model-runner side
    def _inter_data(self, is_prompt: bool, batch: List[_Seq]) -> List[InterDataForSeqGroup]:
        cfg, bs = self.cache_config.paged_evict_config, self.block_size
        out = []
        for s in batch:
            table = list(self.block_manager.get_block_table(s.seq_id))
            kv_len = self.block_manager.get_seq_kv_len(s.seq_id)
            if is_prompt:
                ctx, q = 0, s.prompt_len
            else:
                ctx, q = s.seq_len - 1, 1
                # model_runner.py:540-543 (budget) / page_evict_kv_util_bak.py:69 (percentage)
                if cfg is not None and decode_trigger(s.seq_len, kv_len, cfg, bs):
                    kv_len -= bs
            out.append(InterDataForSeqGroup(
                request_id=s.request_id, is_prompt=is_prompt, seq_ids=[s.seq_id],
                block_tables={s.seq_id: table}, input_tokens=[[0] * q], orig_seq_lens=[s.seq_len],
                seq_lens=[s.seq_len], query_lens=[q], context_lens=[ctx],
                seq_kv_lens={s.seq_id: kv_len} if cfg is not None else None))
        return out

    def build_metadata(self, is_prompt: bool, batch: List[_Seq]) -> TorchCUDASDPAMetadata:
        inter = self._inter_data(is_prompt, batch)
        builder = TorchCUDASDPAMetadataBuilder(_InputBuilder(self._runner, self.block_size, inter))
        seq_lens = [d.seq_lens[0] for d in inter]
        query_lens = [d.query_lens[0] for d in inter]
        md = builder.build(seq_lens, query_lens, -1, len(inter))
        if self.cache_config.paged_evict_config is None:
            md.seq_kv_lens, md.seq_kv_lens_tensor = None, None
        return md

    def _scored_blocks(self, is_prompt: bool, batch: List[_Seq]) -> List[int]:
        """Physical ids of the blocks the layers scored during this step: every full kept
        block of a prefilled prompt, or the last block of each row that just filled."""
        cfg, bs = self.cache_config.paged_evict_config, self.block_size
        if cfg is None or cfg.decode_evict_method != "value_l2":
            return []
        ids: List[int] = []
        for s in batch:
            table = self.block_manager.get_block_table(s.seq_id)
            if is_prompt:
                ids.extend(table[:self.block_manager.get_seq_kv_len(s.seq_id) // bs])
            elif s.seq_len % bs == 0:
                ids.append(table[-1])
        return ids

    # -------------------------------------------------------------------------- step
    def step(self, qkv_fn: QKVFn, keep_outputs: bool = True,
             out_fn: Optional[Callable[[int, torch.Tensor], None]] = None,
             out_buf_fn: Optional[Callable[[int, int], torch.Tensor]] = None) -> StepOutput:
        """qkv_fn(layer, num_tokens) -> (query [T,Hq*D], key [T,Hkv*D], value [T,Hkv*D])
        on self.device.  out_fn(layer, out) receives each layer's attention output (the
        tensor the model's o_proj would consume); out_buf_fn(layer, num_tokens) supplies
        the tensor `forward(..., output=)` writes into."""
        is_prompt, batch = self._schedule()
        md = self.build_metadata(is_prompt, batch)
        n_tok = md.num_prefill_tokens + md.num_decode_tokens
        outs = []
        for layer, impl in enumerate(self.impls):
            q, k, v = qkv_fn(layer, n_tok)
            ob = out_buf_fn(layer, n_tok) if out_buf_fn is not None else None
            o = impl.forward(q, k, v, self.kv_caches[layer], md, output=ob)
            if out_fn is not None:
                out_fn(layer, o)
            if keep_outputs:
                outs.append(o)
        if self.score_exchange is not None:
            self.score_exchange.publish(self._scored_blocks(is_prompt, batch))
        # model_runner.py:1797-1802: only layer 0's decision reaches the block manager
        rmv = dict(self.impls[0].get_rmv_block_idx())
        out = StepOutput(is_prompt, [s.request_id for s in batch], list(md.seq_lens),
                         list(md.seq_kv_lens or md.seq_lens),
                         [list(self.block_manager.get_block_table(s.seq_id)) for s in batch],
                         md, outs, rmv)
        by_req = {s.request_id: s for s in batch}
        for row, (req_id, idx) in rmv.items():           # llm_engine.py:1179-1193
            self.block_manager.free_prunned_blocks(by_req[req_id].seq_id, idx)
        if is_prompt:
            for s in batch:
                s.prefilled = True
                self.running.append(s)
        return out


class PinnedLayerIO:
    """Host <-> device staging for a caller whose per-layer activations live in HOST memory.

    The reference's model runner hands `forward` device tensors; a caller that keeps q/k/v
    on the host (bench.py's e2e leg, an offloaded model) would otherwise pay one blocking
    copy per tensor per layer on the compute stream.  Here the step's inputs sit in ONE
    contiguous pinned arena as fused `[B, (Hq + 2 Hkv) D]` row blocks per layer (the layout
    `qkv_proj` produces, vllm/model_executor/models/llama.py:195-199), the outputs in a second
    one, and a step moves them with a handful of large copies:

      H2D  groups of 1, 1, 2, 4, 8, 16 ... layers on a copy stream that runs ahead of compute
           (the link moves a layer's rows about twice as fast as the layer computes, so a
           group that doubles arrives before the previous one is consumed; only the first
           layer's copy is exposed);
      D2H  the mirror image, ... 16, 8, 4, 2, 1, 1 on a SECOND stream, each group as soon as its
           last layer is done; only the last layer's copy is exposed.

    PCIe is full duplex on these boxes: 55 GB/s either way alone, 45 GB/s each way together
    (profiles/r02_pcie.txt; round 1's "12 GB/s duplex" was a measurement bug), so the two
    directions get their own streams.  A step costs max(compute, H2D, D2H) + the first H2D +
    the last D2H.

    prefetch=True (DecodeGraphRunner only): the whole step's input arena moves as ONE copy, one
    step AHEAD -- while step n computes from landing buffer n % 2, the copy stream fills the other
    one with `qkv_host`, i.e. with step n + 1's activations, so no layer ever waits on the link
    and the attention kernels keep one unbroken programmatic-dependent-launch chain (each
    cross-stream edge inside a captured step costs ~3 us, profiles/r02_e2e_breakdown.txt).
    Contract: runner.step() attends over what the PREVIOUS step() (or prime()) uploaded.

    Use:  io.begin_step(); engine.step(io.qkv, out_fn=io.put_out, out_buf_fn=io.out_buf); io.end_step()
    """

    def __init__(self, num_layers: int, batch: int, num_heads: int, num_kv_heads: int, head_size: int,
                 device, group: int = 0, dtype=torch.bfloat16, duplex: bool = True,
                 prefetch: bool = False) -> None:
        self.L, self.B = num_layers, batch
        self.prefetch = bool(prefetch)

        def doubling(n):
            sizes, left, nxt = [], n, 1
            while left > 0:
                g = min(nxt, left)
                sizes.append(g)
                left -= g
                if len(sizes) >= 2:
                    nxt *= 2
            return sizes
        if group and group > 0:       # fixed group size, both directions
            h_sizes = [min(group, num_layers - lo) for lo in range(0, num_layers, group)]
            d_sizes = list(h_sizes)
        else:
            h_sizes = doubling(num_layers)
            d_sizes = h_sizes[::-1]

        def bounds(sizes):
            out, lo = [], 0
            for n in sizes:
                out.append((lo, lo + n))
                lo += n
            return out
        self.bounds = bounds(h_sizes)          # H2D groups
        self.d_bounds = bounds(d_sizes)        # D2H groups
        self.gidx = [gi for gi, (a, b) in enumerate(self.bounds) for _ in range(a, b)]
        self.d_gidx = [gi for gi, (a, b) in enumerate(self.d_bounds) for _ in range(a, b)]
        self.is_first = [l == self.bounds[self.gidx[l]][0] and not self.prefetch for l in range(num_layers)]
        self.is_last = [l + 1 == self.d_bounds[self.d_gidx[l]][1] for l in range(num_layers)]
        self.qw, self.kw = num_heads * head_size, num_kv_heads * head_size
        w = self.qw + 2 * self.kw
        self.device = torch.device(device)
        self.qkv_host = torch.empty(num_layers, batch, w, dtype=dtype).pin_memory()
        self.out_host = torch.empty(num_layers, batch, self.qw, dtype=dtype).pin_memory()
        # landing buffers: one, or two alternating ones when the upload runs a step ahead
        self.qkv_bufs = torch.empty(2 if self.prefetch else 1, num_layers, batch, w, dtype=dtype,
                                    device=self.device)
        self.qkv_dev = self.qkv_bufs[0]
        self.out_dev = torch.empty(num_layers, batch, self.qw, dtype=dtype, device=self.device)
        self.h2d_stream = torch.cuda.Stream(self.device)
        self.d2h_stream = torch.cuda.Stream(self.device) if duplex else self.h2d_stream
        self.h2d_done = [torch.cuda.Event() for _ in self.bounds]
        self.grp_done = [torch.cuda.Event() for _ in self.d_bounds]
        self.step_done = torch.cuda.Event()
        self.h2d_bytes = self.qkv_host.numel() * self.qkv_host.element_size()
        self.d2h_bytes = self.out_host.numel() * self.out_host.element_size()

    def describe(self) -> dict:
        """Copy schedule, for the bench log."""
        h2d = [self.L] if self.prefetch else [b - a for a, b in self.bounds]
        return {"h2d_groups": h2d, "d2h_groups": [b - a for a, b in self.d_bounds],
                "duplex_streams": self.d2h_stream is not self.h2d_stream,
                "h2d_copies_per_step": len(h2d), "d2h_copies_per_step": len(self.d_bounds),
                "h2d_one_step_ahead": self.prefetch}

    def prime(self, buf: int = 0) -> None:
        """prefetch mode: upload `qkv_host` into landing buffer `buf` now (before the first step)."""
        self.qkv_bufs[buf].copy_(self.qkv_host, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()

    def begin_step(self, buf: int = 0) -> None:
        """Queue every layer group's H2D on the copy stream (after the previous step's
        compute has consumed the staging buffers).  prefetch mode: the step computes from landing
        buffer `buf`; ONE copy fills the other buffer for the next step."""
        cur = torch.cuda.current_stream(self.device)
        self.h2d_stream.wait_stream(cur)
        if self.prefetch:
            with torch.cuda.stream(self.h2d_stream):
                self.qkv_bufs[1 - buf].copy_(self.qkv_host, non_blocking=True)
            return
        with torch.cuda.stream(self.h2d_stream):
            for gi in range(len(self.h2d_done)):
                lo, hi = self.bounds[gi]
                self.qkv_dev[lo:hi].copy_(self.qkv_host[lo:hi], non_blocking=True)
                self.h2d_done[gi].record(self.h2d_stream)

    def qkv(self, layer: int, n_tok: int, wait: bool = True, buf: int = 0):
        if wait and self.is_first[layer]:
            torch.cuda.current_stream(self.device).wait_event(self.h2d_done[self.gidx[layer]])
        row = self.qkv_bufs[buf][layer]
        return row[:, :self.qw], row[:, self.qw:self.qw + self.kw], row[:, self.qw + self.kw:]

    def out_buf(self, layer: int, n_tok: int) -> torch.Tensor:
        return self.out_dev[layer]

    def put_out(self, layer: int, out: torch.Tensor) -> None:
        if out.data_ptr() != self.out_dev[layer].data_ptr():
            self.out_dev[layer].copy_(out.view(self.B, self.qw))
        if self.is_last[layer]:
            gi = self.d_gidx[layer]
            cur = torch.cuda.current_stream(self.device)
            self.grp_done[gi].record(cur)
            self.d2h_stream.wait_event(self.grp_done[gi])
            lo = self.d_bounds[gi][0]
            with torch.cuda.stream(self.d2h_stream):
                self.out_host[lo:layer + 1].copy_(self.out_dev[lo:layer + 1], non_blocking=True)

    def end_step(self) -> None:
        """Block until the step's outputs are in `out_host` (the caller owns them now)."""
        self.step_done.record(self.d2h_stream)
        self.step_done.synchronize()
