"""Decode steps of a fixed set of running sequences, replayed as CUDA graphs.

The reference runs the eviction path with `--enforce-eager` (every documented command;
`TorchCUDASDPAMetadataBuilder.build` has a graph branch, torch_cuda_sdpa.py:626-640, but the
compacted tables change width so it is unusable).  Measured on B200 the eager path of this
package is HOST-bound for the 8B shape (1.28 ms of Python + ctypes per 32-layer step against
1.23 ms of kernels, profiles/r01_e2e_breakdown.txt), so the B200-native step is:

  * all per-step integer state (slot mapping, block tables, kv / seq lengths, blocks per row,
    the decode kernel's chunk plan) lives in ONE fixed device buffer, refreshed by ONE
    pinned H2D copy per step; the host keeps a numpy mirror that it updates incrementally
    (a new block = one element, an eviction = one row shift) instead of rebuilding
    `TorchCUDASDPAMetadata` from Python lists;
  * the per-layer launches (`pc_decode_layer`: [select + compact] -> attention with the
    fused cache write and block scoring) of all layers are captured once per variant
    (trigger step / plain step) and replayed; table width, kv lengths and the chunk plan are
    read from device memory by the kernels, so one capture serves every step;
  * with host-resident activations (`PinnedLayerIO`) the H2D / D2H copies and their stream
    fork / join are part of the same graph.

Semantics are those of `PagedEvictEngine.step` (engine.py): same scheduler arithmetic, same
kernels, same layer-0 feedback; tests/test_gpu_layer.py checks both paths bit for bit.
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from .policy import cdiv, random_victim, zero_slot


_VICT_PENDING = -2      # sentinel in the pinned victims buffer (real entries are -1 or >= 1)


class DecodeGraphRunner:
    def __init__(self, engine, qkv: Optional[Sequence[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]],
                 outs: Optional[Sequence[torch.Tensor]] = None, io=None, max_kv_len: Optional[int] = None,
                 use_graph: bool = True, pdl: Optional[bool] = None) -> None:
        """engine: PagedEvictEngine whose `running` sequences are all prefilled.
        qkv[l] = (query [B,Hq*D], key [B,Hkv*D], value [B,Hkv*D]) STATIC device tensors the
        caller refills before every step (ignored when `io` is given: PinnedLayerIO supplies
        its device staging buffers and moves host data in-graph).  outs[l] [B, Hq*D] static.
        max_kv_len bounds the physical KV length of any row while this runner lives."""
        eng = self.eng = engine
        self.device = eng.device
        self.io = io
        self.use_graph = use_graph
        self.pdl = ops.pdl_default(len(eng.running)) if pdl is None else pdl   # None = measured rule
        self.seqs = list(eng.running)
        if not self.seqs:
            raise RuntimeError("DecodeGraphRunner needs running (prefilled) sequences")
        B = self.B = len(self.seqs)
        bs = self.bs = eng.block_size
        L = self.L = eng.num_layers
        self.cfg = eng.cache_config.paged_evict_config
        impl0 = eng.impls[0]
        self.value_l2 = impl0._value_l2
        self.method = impl0._decode_method       # None = eviction off
        # HEAD's remaining policies: victims drawn on the host ("random"), W2 slot zeroing
        self.host_victims = self.method == "random"
        self.zeroing = self.method in ("streamingLLM-1", "inverse_key_l2")
        self.rng = impl0.rng                     # one RNG for every layer (torch_cuda_sdpa.py:796)
        self.attend_unc = (self.method in ("streamingLLM", "streamingLLM-1")) and impl0.head_compat_uncompacted
        bm = eng.block_manager
        self.bts = [bm.block_tables[s.seq_id] for s in self.seqs]
        cur_max = max(bt.get_seq_kv_len() for bt in self.bts)
        if max_kv_len is None:
            # the bound follows the policy: evict_freq mode holds a row near where it is now;
            # budget mode lets rows below the budget grow up to it (one block per 16 steps, no
            # eviction) and then holds them there; without eviction rows just grow.  A row that
            # still reaches the bound makes step() re-capture with a larger one (_setup).
            if self.cfg is None:
                max_kv_len = cur_max + 4096
            elif self.cfg.cache_prune_type == "budget":
                max_kv_len = max(int(self.cfg.cache_budget), cur_max) + 8 * bs
            else:
                max_kv_len = cur_max + 8 * bs * max(getattr(self.cfg, "evict_freq", 1) or 1, 1)
        self._qkv_arg, self._outs_arg = qkv, outs
        self.log_victims = False  # bench (TP > 1): keep every trigger step's [L, B] victims on the device
        self.victim_log: List[torch.Tensor] = []
        self.record = False      # tests: keep the host-side inputs of the last step in self.last
        self.last: Dict[str, list] = {}
        self.recaptures = -1
        self._setup(int(max_kv_len))

    def _setup(self, max_kv_len: int) -> None:
        """(Re)build everything whose size depends on the table width: the packed device
        state, the host mirror, the workspace and the captured graphs."""
        eng, B, bs, L, io = self.eng, self.B, self.bs, self.L, self.io
        qkv, outs = self._qkv_arg, self._outs_arg
        self.recaptures += 1
        self.max_kv_len = int(max_kv_len)
        M = self.M = cdiv(self.max_kv_len, bs) + 1

        qkv_alt = None
        if io is not None:
            qkv = [io.qkv(l, B, wait=False) for l in range(L)]
            outs = [io.out_dev[l] for l in range(L)]
            if io.prefetch:      # two landing buffers, the step's parity picks one
                qkv_alt = [io.qkv(l, B, wait=False, buf=1) for l in range(L)]
        if outs is None:
            outs = [torch.empty(B, eng.num_heads * eng.head_size, dtype=torch.bfloat16, device=self.device)
                    for _ in range(L)]
        views = lambda xs: [tuple(t.view(B, -1, eng.head_size) for t in x) for x in xs]
        self.qkv_par = [views(qkv)] + ([views(qkv_alt)] if qkv_alt is not None else [])
        self.qkv = self.qkv_par[0]
        self.outs = [o.view(B, eng.num_heads, eng.head_size) for o in outs]
        self.out_flat = list(outs)

        # ---- packed int32 state: device buffer + ring of pinned host copies
        o_slot, o_kv, o_seq, o_nblk, o_plan, o_tab = 0, 2 * B, 3 * B, 4 * B, 5 * B, 5 * B + 4
        per_layer = self.value_l2 or self.host_victims
        self.Lsel = L if per_layer else 1
        # host-decided extras ride in the same packed copy: victims [Lsel, B] int32 ("random"),
        # zeroed slots [L, B] int64 (W2)
        o_vict = o_tab + B * M
        o_zero = o_vict + (self.Lsel * B if self.host_victims else 0)
        o_zero += o_zero & 1                     # int64 alignment
        self.n_meta = o_zero + (2 * L * B if self.zeroing else 0)
        self.meta = torch.zeros(self.n_meta, dtype=torch.int32, device=self.device)
        self._ring = [torch.zeros(self.n_meta, dtype=torch.int32).pin_memory() for _ in range(4)]
        self._ring_ev = [None] * len(self._ring)
        self._ring_i = 0
        m = self.meta
        self.d_slot = m[o_slot:o_kv].view(torch.int64)
        self.d_kv, self.d_seq, self.d_nblk = m[o_kv:o_seq], m[o_seq:o_nblk], m[o_nblk:o_plan]
        self.d_plan = m[o_plan:o_tab]
        self.d_tab = m[o_tab:o_vict].view(B, M)
        self._off = (o_slot, o_kv, o_seq, o_nblk, o_plan, o_tab)
        self._off_vict, self._off_zero = o_vict, o_zero
        # trigger steps: every layer's compacted tables / victims from ONE selection launch
        # (value_l2 / random: per layer; streamingLLM: the victim is index 1 for every layer, one copy)
        self.d_compact = torch.zeros(self.Lsel, B, M, dtype=torch.int32, device=self.device)
        if self.host_victims:        # IN/OUT of the selection kernel: uploaded with the packed state
            self.d_victims = m[o_vict:o_vict + self.Lsel * B].view(self.Lsel, B)
        else:
            self.d_victims = torch.full((self.Lsel, B), -1, dtype=torch.int32, device=self.device)
        self.d_zero = m[o_zero:o_zero + 2 * L * B].view(torch.int64).view(L, B) if self.zeroing else None
        self._vict_host = torch.empty(B, dtype=torch.int32).pin_memory()
        if not hasattr(self, "_vict_stream"):
            self._vict_stream = torch.cuda.Stream(self.device)

        # ---- host mirror
        self.tab = np.zeros((B, M), dtype=np.int32)
        self.nblk = np.zeros(B, dtype=np.int32)
        for r, bt in enumerate(self.bts):
            ids = bt.physical_block_ids
            if len(ids) > M:
                raise RuntimeError("max_kv_len is smaller than a sequence already is")
            self.tab[r, :len(ids)] = ids
            self.nblk[r] = len(ids)
        self.seq = np.asarray([s.seq_len for s in self.seqs], dtype=np.int64)
        self.kv = np.asarray([bt.get_seq_kv_len() for bt in self.bts], dtype=np.int64)
        self.rows = np.arange(B)
        self.cols = np.arange(M)
        # decode tokens not yet booked on the host objects (_Seq.seq_len, BlockTable slot
        # counters): the numpy mirrors are authoritative while the runner steps, the objects are
        # brought up to date when a row takes a new block and by sync_host()
        self.pend = np.zeros(B, dtype=np.int64)
        # evicted table index whose BlockTable entry is not deleted yet (the id is already free)
        self.pend_del = np.full(B, -1, dtype=np.int64)
        self._req_ids = np.asarray([s.request_id for s in self.seqs], dtype=object)

        # ---- decode workspace: pc_paged_decode_workspace_bytes covers EVERY plan the device-side
        # plan word can name for a longest row of up to max_kv_len tokens
        Hq, Hkv, D = eng.num_heads, eng.num_kv_heads, eng.head_size
        need = ops.paged_decode_workspace_bytes(B, Hq, Hkv, D, bs, M * bs)
        self.ws = torch.zeros(need, dtype=torch.uint8, device=self.device)
        self._plan_cache: Dict[int, Tuple[int, int]] = {}

        # ---- per-layer pointers
        self.layers = []
        for l, impl in enumerate(eng.impls):
            kc, vc, (kptr, vptr, blk_stride) = impl._caches(eng.kv_caches[l])
            score = impl._scores(self.device, eng.kv_caches[l].shape[1]) if self.value_l2 else None
            self.layers.append((kptr, vptr, score))
        self.blk_stride = blk_stride
        if self.zeroing:
            self.d_kptrs = torch.tensor([x[0] for x in self.layers], dtype=torch.int64, device=self.device)
            self.d_vptrs = torch.tensor([x[1] for x in self.layers], dtype=torch.int64, device=self.device)
        self.d_pool_ptrs = None
        if self.value_l2:
            ptrs = [sc.pools[sc.local] for _, _, sc in self.layers]
            self.d_pool_ptrs = torch.tensor(ptrs, dtype=torch.int64, device=self.device)
            self.heads_total = self.layers[0][2].heads_total
        self._graphs: Dict[Tuple[bool, int], object] = {}
        self._launches: Dict[bool, int] = {}
        self._ahead = None          # next step's packed host state, prepared while the device runs
        self._warm()
        if self.use_graph:          # capture every variant now, not inside somebody's timed step
            for trig in ((False, True) if self.cfg is not None else (False,)):
                for par in range(len(self.qkv_par)):
                    self._graph(trig, par)
        if io is not None and io.prefetch and self.recaptures == 0:
            self._io_par = 0
            io.prime(0)             # the first step's activations; later ones arrive a step ahead

    # ------------------------------------------------------------------ launches
    def _make_step(self, pdl: bool = False):
        eng = self.eng
        q, k, v = self.qkv[0]
        return ops.make_decode_step(
            self.d_slot, self.d_tab, self.d_kv, self.d_seq, self.B, eng.num_heads, eng.num_kv_heads,
            eng.head_size, self.bs, self.max_kv_len, eng.impls[0].scale, q.stride(0), k.stride(0),
            v.stride(0), self.outs[0].stride(0), self.blk_stride,
            policy=eng.impls[0]._policy_struct() if self.cfg is not None else None,
            n_blocks=self.d_nblk, plan=self.d_plan, workspace=self.ws, pdl=pdl)

    def _launch_layers(self, trig: bool, par: int = 0) -> None:
        """All layers of one decode step on the current stream (captured or eager).
        Trigger step: ONE selection launch for all layers (a layer's victims depend on scores
        written on earlier steps only), then the same attention launches as a plain step, each
        over its layer's compacted tables -- so the PDL chain between the layers' kernels
        survives trigger steps."""
        eng = self.eng
        ds = self._make_step()
        # PDL for a layer whose predecessor in the stream is the previous layer's attention
        # kernel (another layer's pool): its first pages load while that kernel drains
        ds_pdl = self._make_step(pdl=True) if self.pdl else ds
        self._launches[trig] = self.L + (1 if trig else 0) + (1 if self.zeroing else 0)
        io = self.io
        cur = torch.cuda.current_stream(self.device)
        if io is not None:
            io.begin_step(par)
        uncompacted = trig and self.attend_unc     # HEAD's streamingLLM behaviour (DESIGN D5)
        if self.zeroing:     # W2: every layer's zeroed slots in one launch, ahead of the first attention
            ops.zero_cache_slots_layers(self.d_kptrs, self.d_vptrs, self.d_zero, eng.num_kv_heads,
                                        eng.head_size, self.bs, self.blk_stride)
        early_victims = trig and self.value_l2
        if trig:
            ops.evict_select_compact_layers(self.d_tab, self.d_nblk, self.d_seq, ds.policy, self.d_pool_ptrs,
                                            self.heads_total if self.value_l2 else 0, self.d_compact,
                                            self.d_victims)
        if early_victims:
            # layer 0's victims leave for the host NOW, on a side stream, while the attention kernels of
            # the step run: the host frees the blocks and prepares the next step during the step instead
            # of after it (step() polls the pinned buffer for the copy; no device synchronisation)
            ev = torch.cuda.Event()
            ev.record(cur)
            self._vict_stream.wait_event(ev)
            with torch.cuda.stream(self._vict_stream):
                self._vict_host.copy_(self.d_victims[0], non_blocking=True)
        for l in range(self.L):
            if io is not None and io.is_first[l]:
                cur.wait_event(io.h2d_done[io.gidx[l]])
            q, k, v = self.qkv_par[par][l]
            kptr, vptr, score = self.layers[l]
            chained = l > 0 and (io is None or not io.is_first[l])
            tables = None
            if trig and not uncompacted:
                tables = self.d_compact[l if self.Lsel > 1 else 0]
            ops.decode_layer(ds_pdl if chained else ds, self.outs[l], q, k, v, kptr, vptr, score, None, tables)
            if io is not None:
                io.put_out(l, self.out_flat[l])
        if early_victims:
            cur.wait_stream(self._vict_stream)
        if io is not None:
            cur.wait_stream(io.d2h_stream)
            cur.wait_stream(io.h2d_stream)

    def _warm(self) -> None:
        """Load every kernel (lazy module loading must not happen under capture): one eager
        pass of both variants over rows of length zero -- no page is read or written."""
        self.meta.zero_()
        if self.zeroing:
            self.d_zero.fill_(-1)        # slot 0 is a real slot
        io, self.io = self.io, None
        try:
            for trig in ((False, True) if self.cfg is not None else (False,)):
                self._launch_layers(trig)
        finally:
            self.io = io
        torch.cuda.current_stream(self.device).synchronize()

    def _graph(self, trig: bool, par: int = 0):
        g = self._graphs.get((trig, par))
        if g is None:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_layers(trig, par)
            self._graphs[(trig, par)] = g
        return g

    # ---------------------------------------------------------------------- step
    def sync_host(self) -> None:
        """Book what the runner has stepped through on the host objects: the decode tokens
        (`_Seq.seq_len`, BlockTable slot counters) and the evicted table entries.  The allocator is
        always current (block ids are freed and handed out at the step they belong to, in the
        reference's order) and so are the runner's own mirrors (`tab`, `nblk`, `kv`); only the
        per-sequence Python objects are batched (a per-row loop per step costs more than the step at
        batch 512).  Call before reading them."""
        for r in np.nonzero(self.pend_del >= 0)[0].tolist():
            self._apply_pending_delete(r)
        for r in np.nonzero(self.pend)[0].tolist():
            n = int(self.pend[r])
            self.bts[r].append_slots(n)          # never crosses a block boundary: those rows are booked at once
            self.seqs[r].seq_len += n
        self.pend[:] = 0

    def _apply_pending_delete(self, r: int) -> None:
        """Book an eviction whose block id already went back to the allocator on the row's BlockTable."""
        bt = self.bts[r]
        del bt._blocks[int(self.pend_del[r])]
        bt._num_cached_slots -= self.bs
        self.pend_del[r] = -1

    def _trigger(self, seq: np.ndarray) -> np.ndarray:
        cfg, bs = self.cfg, self.bs
        if cfg is None or self.method == "inverse_key_l2":     # model_runner.py:545-547: never frees
            return np.zeros(self.B, dtype=bool)
        ok = self.nblk >= 3
        if cfg.cache_prune_type == "budget":
            return ok & (seq > cfg.cache_budget) & (seq % bs == 0)
        return (ok & (seq % (bs * cfg.evict_freq) == 0)
                & (self.nblk.astype(np.int64) * bs > (cfg.initial_blocks + cfg.num_block_merge) * bs))

    def _pack(self, seq: np.ndarray, kv: np.ndarray):
        """Everything the device needs for the step at logical lengths `seq` / physical lengths
        `kv` (tables as they are now), packed into the next pinned ring buffer.
        -> (trigger mask, any trigger, kv lens handed to attention, slot mapping, ring index)"""
        eng, bs, B = self.eng, self.bs, self.B
        tab, nblk = self.tab, self.nblk
        trig = self._trigger(seq)
        any_trig = bool(trig.any())
        kv_pass = kv - bs * trig if self.cfg is not None else seq               # model_runner.py:540-543
        slot = tab[self.rows, nblk - 1].astype(np.int64) * bs + (seq - 1) % bs
        mx = int(kv_pass.max())
        plan = self._plan_cache.get(mx)
        if plan is None:
            plan = self._plan_cache[mx] = ops.paged_decode_plan(B, eng.num_heads, eng.num_kv_heads,
                                                                eng.head_size, bs, mx)
        i = self._ring_i
        self._ring_i = (i + 1) % len(self._ring)
        if self._ring_ev[i] is not None:
            self._ring_ev[i].synchronize()
        host = self._ring[i].numpy()
        o_slot, o_kv, o_seq, o_nblk, o_plan, o_tab = self._off
        host[o_slot:o_kv].view(np.int64)[:] = slot
        host[o_kv:o_seq] = kv_pass
        host[o_seq:o_nblk] = seq
        host[o_nblk:o_plan] = nblk
        host[o_plan:o_plan + 2] = plan
        host[o_tab:o_tab + tab.size] = tab.reshape(-1)
        # host decisions of HEAD's remaining policies, drawn in HEAD's order: layer by layer, rows
        # in row order (every layer's forward() walks the batch, torch_cuda_sdpa.py:1389,1440,1503)
        hv = None
        if self.host_victims and any_trig:
            hv = np.full((self.Lsel, B), -1, dtype=np.int32)
            rows = np.nonzero(trig)[0].tolist()
            for l in range(self.Lsel):
                for r in rows:
                    v = random_victim(self.rng, int(nblk[r]), self.cfg)
                    if not 1 <= v < nblk[r]:
                        raise RuntimeError(f"random victim {v} outside the row's {nblk[r]} blocks")
                    hv[l, r] = v
            host[self._off_vict:self._off_vict + hv.size] = hv.reshape(-1)
        if self.zeroing:
            z = host[self._off_zero:self._off_zero + 2 * self.L * B].view(np.int64).reshape(self.L, B)
            z[:] = -1
            over = np.nonzero(seq > self.cfg.cache_budget)[0].tolist()
            if self.method == "streamingLLM-1":       # the same slot in every layer, no draw
                for r in over:
                    z[:, r] = zero_slot(None, int(seq[r]), tab[r], self.cfg, bs)
            else:
                for l in range(self.L):
                    for r in over:
                        z[l, r] = zero_slot(self.rng, int(seq[r]), tab[r, :nblk[r]], self.cfg, bs)
        return trig, any_trig, kv_pass, slot, i, hv

    def step(self) -> Dict[int, Tuple[str, int]]:
        """One decode step of every sequence.  The caller has refilled the static q/k/v
        buffers (or `io.qkv_host`).  Returns layer 0's {row: (request_id, table index)}
        after applying it to the block manager (llm_engine.py:1179-1193)."""
        eng, bs, B = self.eng, self.bs, self.B
        bm = eng.block_manager
        if not bm.can_append_slots(B):
            raise RuntimeError("out of KV blocks (preemption is outside the hot path)")
        # a row about to outgrow the captured table width: re-capture with a larger bound
        grow = self.nblk + (self.kv % bs == 0)
        if int(grow.max()) > self.M - 1:
            torch.cuda.current_stream(self.device).synchronize()
            self.sync_host()
            self._setup(max(2 * self.max_kv_len - int(self.kv.min()), (int(grow.max()) + 8) * bs))
        tab, nblk = self.tab, self.nblk
        # scheduler.py _schedule_default, vectorised: only rows whose last block is full touch
        # their BlockTable (allocator order = row order, like the per-sequence loop)
        new_rows = np.nonzero(self.kv % bs == 0)[0]
        self.pend += 1
        if len(new_rows):
            for r in new_rows.tolist():
                bt = self.bts[r]
                if self.pend_del[r] >= 0:
                    self._apply_pending_delete(r)
                bt.append_slots(int(self.pend[r]))
                self.seqs[r].seq_len += int(self.pend[r])
                tab[r, nblk[r]] = bt.physical_block_ids[-1]
            self.pend[new_rows] = 0
            nblk[new_rows] += 1
        self.seq += 1
        self.kv += 1
        ahead, self._ahead = self._ahead, None
        if ahead is None:
            ahead = self._pack(self.seq, self.kv)
        trig, any_trig, kv_pass, slot, i, host_vict = ahead
        if self.record:
            self.last = {"seq_lens": self.seq.tolist(), "kv_lens": kv_pass.tolist(), "slots": slot.tolist(),
                         "tables": [tab[r, :nblk[r]].tolist() for r in range(B)],
                         "trigger_rows": np.nonzero(trig)[0].tolist()}
        # ---- one packed H2D
        self.meta.copy_(self._ring[i], non_blocking=True)
        ev = self._ring_ev[i] or torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._ring_ev[i] = ev
        # ---- all layers
        par = 0
        if self.io is not None and self.io.prefetch:
            par = self._io_par                           # landing buffer the previous step filled
            self._io_par = 1 - par
        if any_trig and self.value_l2:
            self._vict_host.fill_(_VICT_PENDING)             # host write; the step's copy overwrites it
        if self.use_graph:
            self._graph(any_trig, par).replay()
            ops._launched(self._launches[any_trig])      # kernel nodes of the replayed graph
        else:
            self._launch_layers(any_trig, par)           # ops.decode_layer counts its own
        # ---- host work of the NEXT step while the device runs this one (matters when the caller
        # waits for every step's result, e.g. host-resident activations): valid when this step
        # changes no table (no eviction) and the next one appends no block
        if not any_trig and not bool((self.kv % bs == 0).any()):
            self._ahead = self._pack(self.seq + 1, self.kv + 1)
        if eng.score_exchange is not None:
            scored = []
            if self.value_l2:
                rs = np.nonzero(self.seq % bs == 0)[0]
                scored = tab[rs, nblk[rs] - 1].tolist()
            eng.score_exchange.publish(scored)
        # ---- layer-0 feedback (model_runner.py:1797-1802)
        rmv: Dict[int, Tuple[str, int]] = {}
        if any_trig:
            rows = np.nonzero(trig)[0]
            if self.log_victims and self.Lsel > 1:
                self.victim_log.append(self.d_victims.clone())
            if self.host_victims:                  # the host chose them: nothing to read back
                idx = host_vict[0][rows].astype(np.int64)
            elif self.value_l2:
                vict = self._vict_host.numpy()       # filled by the step's side-stream copy (_launch_layers)
                t0 = time.perf_counter()
                while (vict == _VICT_PENDING).any():
                    if time.perf_counter() - t0 > 30.0:
                        raise RuntimeError("victims of the trigger step did not arrive within 30 s")
                rows = rows[vict[rows] >= 0]
                idx = vict[rows].astype(np.int64)
            else:
                idx = np.ones(len(rows), dtype=np.int64)
            if len(rows):
                # llm_engine.py:1179-1193 -> scheduler.py:1879-1890, for all rows at once: the freed
                # ids go back to the allocator in row order (free = appendleft, so extendleft keeps
                # the reuse order of the per-row loop); the BlockTable objects are brought up to
                # date when the row next takes a block or by sync_host() (a per-row Python loop
                # cost 0.6 ms of a 1.6 ms trigger step at batch 512)
                rl, il = rows.tolist(), idx.tolist()
                rmv = dict(zip(rl, zip(self._req_ids[rows].tolist(), il)))
                late = rows[self.pend_del[rows] >= 0]
                for r in late.tolist():                           # (a second eviction before the row's next block: rare)
                    self._apply_pending_delete(r)
                eng.block_manager.allocator._free.extendleft(tab[rows, idx].tolist())
                self.pend_del[rows] = idx
            if len(rows):
                # delete column idx of every evicted row, shift left, 0-pad: all rows at once
                sel = tab[rows]
                shifted = np.concatenate([sel[:, 1:], np.zeros((len(rows), 1), dtype=sel.dtype)], axis=1)
                tab[rows] = np.where(self.cols[None, :] >= idx[:, None], shifted, sel)
                nblk[rows] -= 1
                self.kv[rows] -= bs
        if self.io is not None:
            torch.cuda.current_stream(self.device).synchronize()    # the caller owns out_host now
        return rmv
