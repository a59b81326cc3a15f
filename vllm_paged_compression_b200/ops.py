"""Op wrappers over the C ABI of libpagedevict_b200.so (torch tensors in, device work out).

Plays the role of vllm/_custom_ops.py:70-130,984-1023 (`paged_attention_v1/v2`,
`reshape_and_cache`, `copy_blocks`, `swap_blocks`) plus the fused ops the B200 path adds
(SURVEY.md 8b).  torch is plumbing here: data_ptr()/stride()/current stream only.  Every
wrapper raises RuntimeError with the native message on failure (the reference's
TORCH_CHECK convention) and there is no CPU / eager fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import MAX_PEERS, PcDecodeStep, PcEvictPolicy, PcScorePool, check

PC_BF16 = 0

# kernels launched through this module since import (bench.py reads it for `gpu_launches`)
launch_count = 0


def _launched(n: int = 1) -> None:
    global launch_count
    launch_count += n

# struct pc_evict_policy.decode_method: "random" = victim chosen by the host; the W2 variants
# evict like streamingLLM (table entry 1) / never (inverse_key_l2: the host never triggers)
_DECODE_METHOD = {"streamingLLM": 0, "value_l2": 1, "random": 2, "streamingLLM-1": 0, "inverse_key_l2": 0}


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _req_cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("paged-eviction ops need CUDA tensors: this package has no CPU fallback")


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype != torch.bfloat16:
        raise RuntimeError(f"only bf16 KV cache / activations are built, got {t.dtype}")
    return PC_BF16


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def make_policy(pol, block_size: int) -> PcEvictPolicy:
    """PagedEvictConfig (config.py) -> struct pc_evict_policy."""
    return PcEvictPolicy(
        prune_type=0 if pol.cache_prune_type == "budget" else 1,
        block_size=block_size,
        cache_budget=int(pol.cache_budget),
        evict_freq=int(pol.evict_freq),
        initial_blocks=int(pol.initial_blocks),
        num_block_merge=int(pol.num_block_merge),
        decode_method=_DECODE_METHOD[pol.decode_evict_method],
        topk_blocks=int(pol.topk_blocks) if pol.topk_blocks is not None else -1)


def make_score_pool(pool: torch.Tensor, head_offset: int = 0,
                    peer_ptrs: Optional[Sequence[int]] = None, local: int = 0) -> PcScorePool:
    """struct pc_score_pool for one layer.  pool fp32 [num_blocks, heads_total] on the device
    (this rank's copy); peer_ptrs = data pointers of EVERY rank's pool in rank order (symmetric
    memory, this rank's own at index `local`) or None for a single GPU."""
    if not pool.is_cuda or pool.dtype != torch.float32 or pool.dim() != 2 or not pool.is_contiguous():
        raise RuntimeError("score pool must be a contiguous fp32 CUDA tensor [num_blocks, heads_total]")
    sp = PcScorePool()
    ptrs = list(peer_ptrs) if peer_ptrs else [pool.data_ptr()]
    if len(ptrs) > MAX_PEERS:
        raise RuntimeError(f"at most {MAX_PEERS} score pools (tensor-parallel ranks) are supported")
    if peer_ptrs and ptrs[local] != pool.data_ptr():
        raise RuntimeError("peer_ptrs[local] must be this rank's own pool")
    for i, a in enumerate(ptrs):
        sp.pools[i] = a
    sp.n_pools, sp.local = len(ptrs), (local if peer_ptrs else 0)
    sp.heads_total, sp.head_offset = pool.shape[1], int(head_offset)
    sp._keepalive = pool
    return sp


def _sp(score: Optional[PcScorePool]):
    return None if score is None else C.byref(score)


def _cache_dims(key_cache: torch.Tensor, value_cache: torch.Tensor) -> Tuple[int, int, int, int]:
    """key_cache [nb, Hkv, D/8, bs, 8], value_cache [nb, Hkv, D, bs] (paged_attn.py:49-62)."""
    nb, Hkv, D, bs = value_cache.shape
    if key_cache.shape != (nb, Hkv, D // 8, bs, 8):
        raise RuntimeError(f"key_cache shape {tuple(key_cache.shape)} does not match value_cache "
                           f"{tuple(value_cache.shape)}")
    if value_cache.stride()[1:] != (D * bs, bs, 1) or key_cache.stride(0) != value_cache.stride(0):
        raise RuntimeError("key/value cache blocks must be dense [Hkv, D, bs] runs with equal block stride")
    return Hkv, D, bs, value_cache.stride(0)


# ------------------------------------------------------------------------------ W1
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str = "auto", k_scale: float = 1.0, v_scale: float = 1.0) -> None:
    """Same schema as _C_cache_ops::reshape_and_cache (vllm/_custom_ops.py:984-996).
    key/value [N, Hkv, D]; slot_mapping int64 [N], negative = skip."""
    _req_cuda(key, value, key_cache, value_cache, slot_mapping)
    if kv_cache_dtype != "auto" or k_scale != 1.0 or v_scale != 1.0:
        raise RuntimeError("only kv_cache_dtype='auto' with k_scale == v_scale == 1 is supported "
                           "(torch_cuda_sdpa.py:769-772,837)")
    Hkv, D, bs, blk_stride = _cache_dims(key_cache, value_cache)
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor")
    n = slot_mapping.numel()
    if key.shape[0] < n or value.shape[0] < n:
        raise RuntimeError("key/value have fewer rows than slot_mapping")
    for t in (key, value):
        if t.dim() == 3 and t.stride()[1:] != (D, 1) or t.dim() == 2 and t.stride(1) != 1:
            raise RuntimeError("key/value rows must be dense [Hkv, D]")
    check(_lib.load().pc_reshape_and_cache(
        key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
        slot_mapping.data_ptr(), n, Hkv, D, bs, key.stride(0), value.stride(0), blk_stride,
        _dtype_code(key), _stream(key)))
    _launched()


# ----------------------------------------------------------------------- R1+W1+S3
def prune_prompt_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                           value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                           req_meta: torch.Tensor, tile_offsets: torch.Tensor, total_tiles: int,
                           score: Optional[PcScorePool] = None) -> None:
    """Fused streamingLLM prompt prune + cache write + prompt block scoring.
    req_meta int32 [R,5] = {src_start, kept, first, pruned, dst_start}; tile_offsets int32
    [R+1]; score = the layer's score pool (optional)."""
    _req_cuda(key, value, key_cache, value_cache, slot_mapping, req_meta, tile_offsets)
    Hkv, D, bs, blk_stride = _cache_dims(key_cache, value_cache)
    if req_meta.dtype != torch.int32 or tile_offsets.dtype != torch.int32:
        raise RuntimeError("req_meta / tile_offsets must be int32")
    check(_lib.load().pc_prune_prompt_and_cache(
        key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
        slot_mapping.data_ptr(), req_meta.data_ptr(), tile_offsets.data_ptr(),
        req_meta.shape[0], int(total_tiles), Hkv, D, bs, key.stride(0), value.stride(0),
        blk_stride, _sp(score), _dtype_code(key), _stream(key)))
    _launched()


# ------------------------------------------------------------------- P3+S4+C1+C2
def evict_select_compact(block_tables: torch.Tensor, n_blocks: torch.Tensor,
                         seq_lens: torch.Tensor, kv_lens: Optional[torch.Tensor],
                         policy: PcEvictPolicy, num_kv_heads: int,
                         out_tables: torch.Tensor, out_kv_lens: Optional[torch.Tensor],
                         out_victim: torch.Tensor, score: Optional[PcScorePool] = None,
                         kv_lens_pre_adjusted: bool = True) -> None:
    """On-device decode trigger + victim argmin + table compaction (one launch)."""
    _req_cuda(block_tables, n_blocks, seq_lens, kv_lens, out_tables, out_kv_lens, out_victim)
    for t in (block_tables, n_blocks, seq_lens, out_tables, out_victim):
        if t.dtype != torch.int32:
            raise RuntimeError("block tables / lengths / victims must be int32")
    if block_tables.stride(-1) != 1 or out_tables.stride(-1) != 1:
        raise RuntimeError("block table rows must be contiguous")
    B = block_tables.shape[0]
    check(_lib.load().pc_evict_select_compact(
        block_tables.data_ptr(), block_tables.stride(0), n_blocks.data_ptr(), seq_lens.data_ptr(),
        _ptr(kv_lens), B, C.byref(policy), _sp(score), num_kv_heads, out_tables.data_ptr(),
        out_tables.stride(0), _ptr(out_kv_lens), out_victim.data_ptr(),
        1 if kv_lens_pre_adjusted else 0, _stream(block_tables)))
    _launched()


# ------------------------------------------------------------------------ A2 + S2
def paged_decode_workspace_bytes(batch: int, num_heads: int, num_kv_heads: int, head_size: int,
                                 block_size: int, max_kv_len: int) -> int:
    return int(_lib.load().pc_paged_decode_workspace_bytes(batch, num_heads, num_kv_heads,
                                                           head_size, block_size, max_kv_len))


_workspaces = {}


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch for the decode kernel's stream-K partials and arrival
    tickets; grown on demand, reused.  Zero-filled once: the kernel leaves its ticket
    region zero after every launch (include/pagedevict_b200.h)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def paged_decode_attn_score(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                            value_cache: torch.Tensor, block_tables: torch.Tensor,
                            kv_lens: torch.Tensor, scale: float, max_kv_len: int,
                            seq_lens: Optional[torch.Tensor] = None,
                            score: Optional[PcScorePool] = None,
                            workspace: Optional[torch.Tensor] = None) -> None:
    """Paged decode attention over (compacted) tables; rows whose last block fills this
    step (seq_lens % bs == 0) get that block's per-head value_l2 partial written to the
    score pool in the same pass.  query/out [B, Hq, D]."""
    _req_cuda(out, query, key_cache, value_cache, block_tables, kv_lens, seq_lens)
    Hkv, D, bs, blk_stride = _cache_dims(key_cache, value_cache)
    B, Hq, Dq = query.shape
    if Dq != D or out.shape != query.shape:
        raise RuntimeError("query/out must be [B, Hq, D] matching the cache head size")
    if query.stride()[1:] != (D, 1) or out.stride()[1:] != (D, 1):
        raise RuntimeError("query/out rows must be dense [Hq, D]")
    if block_tables.dtype != torch.int32 or kv_lens.dtype != torch.int32 or block_tables.stride(-1) != 1:
        raise RuntimeError("block_tables / kv_lens must be int32, table rows contiguous")
    need = paged_decode_workspace_bytes(B, Hq, Hkv, D, bs, max_kv_len)
    if workspace is None:
        workspace = _workspace(query.device, need)
    check(_lib.load().pc_paged_decode_attn_score(
        out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
        block_tables.data_ptr(), block_tables.stride(0), kv_lens.data_ptr(), B, Hq, Hkv, D, bs,
        int(max_kv_len), float(scale), query.stride(0), out.stride(0), blk_stride,
        _ptr(seq_lens) if score is not None else None, _sp(score), workspace.data_ptr(),
        workspace.numel(), _dtype_code(query), _stream(query)))
    _launched()


# ------------------------------------------------------- fused decode step of one layer
def make_decode_step(slot_mapping: torch.Tensor, block_tables: torch.Tensor, kv_lens: torch.Tensor,
                     seq_lens: torch.Tensor, batch: int, num_heads: int, num_kv_heads: int,
                     head_size: int, block_size: int, max_kv_len: int, scale: float,
                     q_stride: int, key_stride: int, value_stride: int, out_stride: int,
                     kv_block_stride: int, policy: Optional[PcEvictPolicy] = None,
                     n_blocks: Optional[torch.Tensor] = None,
                     compact_tables: Optional[torch.Tensor] = None, any_trigger: bool = False,
                     attend_uncompacted: bool = False, fuse_cache_write: bool = True,
                     plan: Optional[torch.Tensor] = None,
                     workspace: Optional[torch.Tensor] = None, pdl: bool = False) -> PcDecodeStep:
    """struct pc_decode_step: everything that is the same for every layer of one engine
    step (built once per step, passed by pointer to `decode_layer`)."""
    _req_cuda(slot_mapping, block_tables, kv_lens, seq_lens, n_blocks, compact_tables)
    dev = block_tables.device
    st = PcDecodeStep()
    st.slot_mapping, st.block_tables = slot_mapping.data_ptr(), block_tables.data_ptr()
    st.table_stride = block_tables.stride(0)
    st.kv_lens, st.seq_lens = kv_lens.data_ptr(), seq_lens.data_ptr()
    st.n_blocks = _ptr(n_blocks)
    st.compact_tables = _ptr(compact_tables)
    st.compact_stride = compact_tables.stride(0) if compact_tables is not None else 0
    st.batch, st.num_heads, st.num_kv_heads = batch, num_heads, num_kv_heads
    st.head_size, st.block_size, st.max_kv_len = head_size, block_size, int(max_kv_len)
    st.scale = float(scale)
    st.any_trigger, st.attend_uncompacted = int(any_trigger), int(attend_uncompacted)
    st.q_stride, st.key_stride, st.value_stride = q_stride, key_stride, value_stride
    st.out_stride, st.kv_block_stride = out_stride, kv_block_stride
    if policy is not None:
        st.policy = policy
    need = paged_decode_workspace_bytes(batch, num_heads, num_kv_heads, head_size, block_size, max_kv_len)
    ws = workspace if workspace is not None else _workspace(dev, need)
    if ws.numel() < need:
        raise RuntimeError(f"decode workspace too small: {ws.numel()} < {need} bytes")
    st.workspace, st.workspace_bytes = ws.data_ptr(), ws.numel()
    st.dtype = PC_BF16
    st.stream = torch.cuda.current_stream(dev).cuda_stream
    fused = bool(fuse_cache_write) and not (any_trigger and attend_uncompacted)
    st.fuse_cache_write = int(fused)
    _req_cuda(plan)
    st.plan = _ptr(plan)
    st.pdl = int(bool(pdl))
    st._keepalive = (slot_mapping, block_tables, kv_lens, seq_lens, n_blocks, compact_tables, ws, plan)
    st._launches = (1 if fused else 2) + (1 if any_trigger else 0)
    return st


def paged_decode_plan(batch: int, num_heads: int, num_kv_heads: int, head_size: int, block_size: int,
                      max_kv_len: int, num_sms: int = 0) -> Tuple[int, int]:
    """(chunk_pages, team) the host planner picks for this launch shape (pc_decode_step.plan);
    num_sms = 0: the current device's SM count."""
    out = (C.c_int32 * 2)()
    check(_lib.load().pc_paged_decode_plan(batch, num_heads, num_kv_heads, head_size, block_size,
                                           max_kv_len, int(num_sms), out))
    return int(out[0]), int(out[1])


def decode_layer(step: PcDecodeStep, out: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                 value: torch.Tensor, key_cache_ptr: int, value_cache_ptr: int,
                 score: Optional[PcScorePool] = None, out_victim: Optional[torch.Tensor] = None,
                 attend_tables: Optional[torch.Tensor] = None) -> None:
    """One layer's decode step in one native call (cache write -> [select + compact] ->
    attention + scoring).  Shapes / strides were validated when `step` was built.
    attend_tables [B, W] int32: this layer's tables from `evict_select_compact_layers` (the call
    then runs no selection of its own)."""
    check(_lib.load().pc_decode_layer(C.byref(step), out.data_ptr(), query.data_ptr(), key.data_ptr(),
                                      value.data_ptr(), key_cache_ptr, value_cache_ptr, _sp(score),
                                      _ptr(out_victim), _ptr(attend_tables),
                                      attend_tables.stride(0) if attend_tables is not None else 0))
    _launched(1 if attend_tables is not None else step._launches)


def evict_select_compact_layers(block_tables: torch.Tensor, n_blocks: torch.Tensor, seq_lens: torch.Tensor,
                                policy: PcEvictPolicy, pool_ptrs: Optional[torch.Tensor], heads_total: int,
                                out_tables: torch.Tensor, out_victims: torch.Tensor) -> None:
    """P3 + S4 + C1 of every layer in one launch.  pool_ptrs int64 [L] on the device (data
    pointers of this rank's score pool of each layer) or None (streamingLLM); out_tables int32
    [L, B, W]; out_victims int32 [L, B]."""
    _req_cuda(block_tables, n_blocks, seq_lens, pool_ptrs, out_tables, out_victims)
    L, B, W = out_tables.shape
    if out_victims.shape != (L, B) or not out_tables.is_contiguous() or not out_victims.is_contiguous():
        raise RuntimeError("out_tables must be contiguous [L, B, W], out_victims contiguous [L, B]")
    if pool_ptrs is not None and (pool_ptrs.dtype != torch.int64 or pool_ptrs.numel() != L):
        raise RuntimeError("pool_ptrs must be int64 [L]")
    check(_lib.load().pc_evict_select_compact_layers(
        block_tables.data_ptr(), block_tables.stride(0), n_blocks.data_ptr(), seq_lens.data_ptr(), B, L,
        C.byref(policy), _ptr(pool_ptrs), int(heads_total), out_tables.data_ptr(), W, B * W,
        out_victims.data_ptr(), _stream(block_tables)))
    _launched()


def _check_compat(kv_cache_dtype, k_scale, v_scale, alibi_slopes, blocksparse_vert_stride):
    if alibi_slopes is not None:
        raise RuntimeError("alibi_slopes are not supported by the paged-eviction backend")
    if kv_cache_dtype != "auto" or k_scale != 1.0 or v_scale != 1.0:
        raise RuntimeError("only kv_cache_dtype='auto' with k_scale == v_scale == 1 is supported")
    if blocksparse_vert_stride > 1:
        raise RuntimeError("blocksparse attention is not supported")


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                       seq_lens, block_size, max_seq_len, alibi_slopes=None, kv_cache_dtype="auto",
                       k_scale=1.0, v_scale=1.0, tp_rank=0, blocksparse_local_blocks=0,
                       blocksparse_vert_stride=0, blocksparse_block_size=64,
                       blocksparse_head_sliding_step=0) -> None:
    """Schema-compatible entry (vllm/_custom_ops.py:70-96) routed to the split-KV kernel."""
    _check_compat(kv_cache_dtype, k_scale, v_scale, alibi_slopes, blocksparse_vert_stride)
    paged_decode_attn_score(out, query, key_cache, value_cache, block_tables, seq_lens, scale,
                            max_seq_len)


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                       alibi_slopes=None, kv_cache_dtype="auto", k_scale=1.0, v_scale=1.0, tp_rank=0,
                       blocksparse_local_blocks=0, blocksparse_vert_stride=0,
                       blocksparse_block_size=64, blocksparse_head_sliding_step=0) -> None:
    """Schema-compatible entry (vllm/_custom_ops.py:99-130); exp_sums / max_logits /
    tmp_out are not touched (the kernel keeps its own split-KV workspace)."""
    _check_compat(kv_cache_dtype, k_scale, v_scale, alibi_slopes, blocksparse_vert_stride)
    paged_decode_attn_score(out, query, key_cache, value_cache, block_tables, seq_lens, scale,
                            max_seq_len)


def pdl_default(batch: int) -> bool:
    """Programmatic dependent launch between consecutive layers' decode kernels: the schedule
    prologue and the first ring of pages of layer l+1 load while layer l drains.  Measured on
    B200 at the bench shape and its tensor-parallel per-rank shapes (us per launch, PDL / plain):
    B 64 28.1 / 34.6, B 128 28.1 / 34.6, B 256 28.2 / 34.5, B 512 28.8 / 35.1
    (profiles/r02_tp_shape_pdl.txt).  Round 1 saw PDL lose above 128 rows; that was an O(B) scan
    for empty rows on CTA 0 inside the kernel, not PDL -- with the scan gone PDL wins everywhere."""
    return True


# ------------------------------------------------------------------------------ A1
def prefill_attn_topleft_causal(out: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                value: torch.Tensor, req_meta: torch.Tensor, max_q_len: int,
                                scale: float, req_meta_host=None) -> None:
    """query/out [T, Hq, D], key/value [T, Hkv, D] (un-pruned prompt rows); req_meta int32
    [R,6] = {q_start, q_len, kv_src_start, kv_len(kept), first, pruned}.  req_meta_host: the
    same rows as a contiguous int32 numpy array / CPU tensor -- lets the library validate them
    before the launch (RuntimeError instead of a device-side trap)."""
    _req_cuda(out, query, key, value, req_meta)
    host_ptr = None
    if req_meta_host is not None:
        h = req_meta_host if isinstance(req_meta_host, torch.Tensor) else torch.from_numpy(req_meta_host)
        if h.is_cuda or h.dtype != torch.int32 or not h.is_contiguous() or h.numel() != req_meta.numel():
            raise RuntimeError("req_meta_host must be a contiguous int32 CPU copy of req_meta")
        host_ptr = h.data_ptr()
    T, Hq, D = query.shape
    Hkv = key.shape[1]
    if req_meta.dtype != torch.int32 or not req_meta.is_contiguous():
        raise RuntimeError("req_meta must be contiguous int32 [R, 6]")
    for t in (query, out, key, value):
        if t.stride()[1:] != (D, 1):
            raise RuntimeError("q/k/v/out rows must be dense [H, D]")
    check(_lib.load().pc_prefill_attn_topleft_causal(
        out.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(), req_meta.data_ptr(),
        req_meta.shape[0], int(max_q_len), int(min(T, key.shape[0], value.shape[0])), Hq, Hkv, D, float(scale), query.stride(0), out.stride(0),
        key.stride(0), value.stride(0), host_ptr, _dtype_code(query), _stream(query)))
    _launched()


# ------------------------------------------------------------------------------ f3
_TOKEN_SCORE_KIND = {"token_value_l2": 0, "value_l2": 0, "token_inverse_key_l2": 1, "inverse_key_l2": 1}


def prompt_token_prune(key: torch.Tensor, value: torch.Tensor, out_key: torch.Tensor, out_value: torch.Tensor,
                       req_meta: torch.Tensor, mid_offsets: torch.Tensor, tok_offsets: torch.Tensor,
                       total_tokens: int, total_mid: int, kind: str, head_mixing_repack: bool = True,
                       workspace: Optional[torch.Tensor] = None) -> None:
    """Token-level prompt prune of local / global / inverse_key_l2 (kvcache_prunner.py:220-279) for a
    prefill batch.  key/value [T, Hkv, D] -> out_key/out_value rows [dst_start, dst_start + q_len -
    k_drop) per request.  req_meta int32 [R, 6] = {src_start, q_len, first, mid, k_drop, dst_start},
    mid_offsets / tok_offsets int32 [R + 1] (device)."""
    _req_cuda(key, value, out_key, out_value, req_meta, mid_offsets, tok_offsets)
    T, Hkv, D = key.shape
    for t in (key, value, out_key, out_value):
        if t.dim() != 3 or t.shape[1:] != (Hkv, D) or t.stride()[1:] != (D, 1):
            raise RuntimeError("key/value/out rows must be dense [Hkv, D]")
    for t in (req_meta, mid_offsets, tok_offsets):
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise RuntimeError("req_meta / mid_offsets / tok_offsets must be contiguous int32")
    lib = _lib.load()
    need = int(lib.pc_prompt_token_prune_workspace_bytes(int(total_mid), Hkv))
    if workspace is None:
        workspace = torch.empty(need, dtype=torch.uint8, device=key.device)
    check(lib.pc_prompt_token_prune(
        key.data_ptr(), value.data_ptr(), out_key.data_ptr(), out_value.data_ptr(), req_meta.data_ptr(),
        mid_offsets.data_ptr(), tok_offsets.data_ptr(), req_meta.shape[0], int(total_tokens), int(total_mid),
        Hkv, D, _TOKEN_SCORE_KIND[kind], 1 if head_mixing_repack else 0, key.stride(0), value.stride(0),
        out_key.stride(0), out_value.stride(0), workspace.data_ptr(), workspace.numel(), _dtype_code(key),
        _stream(key)))
    _launched(4 if total_mid > 0 else 1)


# ------------------------------------------------------------------------------ W2
def zero_cache_slots(key_cache: torch.Tensor, value_cache: torch.Tensor, slots: torch.Tensor) -> None:
    """Zero the K and V rows (all kv heads) at `slots` (int64 on the device, negative = skip): the
    zero-row write of streamingLLM-1 / inverse_key_l2 (torch_cuda_sdpa.py:919-933,962-975)."""
    _req_cuda(key_cache, value_cache, slots)
    if slots.numel() == 0:
        return
    Hkv, D, bs, blk_stride = _cache_dims(key_cache, value_cache)
    _dtype_code(key_cache)
    if slots.dtype != torch.int64 or not slots.is_contiguous():
        raise RuntimeError("slots must be a contiguous int64 tensor")
    check(_lib.load().pc_zero_cache_slots(key_cache.data_ptr(), value_cache.data_ptr(), slots.data_ptr(),
                                          slots.numel(), Hkv, D, bs, blk_stride, _stream(key_cache)))
    _launched()


def zero_cache_slots_layers(key_ptrs: torch.Tensor, value_ptrs: torch.Tensor, slots: torch.Tensor,
                            num_kv_heads: int, head_size: int, block_size: int, kv_block_stride: int) -> None:
    """The same for every layer of a step in one launch.  key_ptrs / value_ptrs int64 [L] on the
    device (pool base pointers per layer), slots int64 [L, n] on the device."""
    _req_cuda(key_ptrs, value_ptrs, slots)
    L, n = slots.shape
    if key_ptrs.dtype != torch.int64 or value_ptrs.dtype != torch.int64 or key_ptrs.numel() != L \
            or value_ptrs.numel() != L:
        raise RuntimeError("key_ptrs / value_ptrs must be int64 [L]")
    if slots.dtype != torch.int64 or slots.stride(1) != 1:
        raise RuntimeError("slots must be int64 [L, n] with contiguous rows")
    check(_lib.load().pc_zero_cache_slots_layers(key_ptrs.data_ptr(), value_ptrs.data_ptr(), slots.data_ptr(),
                                                 slots.stride(0), n, L, num_kv_heads, head_size, block_size,
                                                 kv_block_stride, _stream(slots)))
    _launched()


# ------------------------------------------------------------------- block maintenance
def copy_blocks(key_cache: torch.Tensor, value_cache: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """One layer of _C_cache_ops::copy_blocks (vllm/_custom_ops.py:1015-1018).
    block_mapping int64 [P, 2] on the device."""
    _req_cuda(key_cache, value_cache, block_mapping)
    if block_mapping.numel() == 0:
        return
    if block_mapping.dtype != torch.int64 or not block_mapping.is_contiguous():
        raise RuntimeError("block_mapping must be contiguous int64 [P, 2]")
    block_bytes = key_cache.stride(0) * key_cache.element_size()
    check(_lib.load().pc_copy_blocks(key_cache.data_ptr(), value_cache.data_ptr(),
                                     block_mapping.data_ptr(), block_mapping.shape[0], block_bytes,
                                     _stream(key_cache)))


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """swap_blocks (vllm/_custom_ops.py:1021-1023): src/dst [num_blocks, ...] pools on
    device or pinned host; block_mapping int64 [P, 2] on the HOST."""
    if block_mapping.is_cuda:
        raise RuntimeError("swap_blocks: block_mapping must be a CPU tensor")
    bm = block_mapping.to(torch.int64).contiguous()
    if bm.numel() == 0:
        return
    block_bytes = src.stride(0) * src.element_size()
    dev = src if src.is_cuda else dst
    _req_cuda(dev)
    check(_lib.load().pc_swap_blocks(src.data_ptr(), dst.data_ptr(), bm.data_ptr(), bm.shape[0],
                                     block_bytes, _stream(dev)))
