"""torch-CPU restatement of the two native ops + prefill SDPA.  TEST INFRASTRUCTURE ONLY.

Layouts (vllm/attention/ops/paged_attn.py:40-62):
  kv_cache    [2, num_blocks, block_size * Hkv * D]
  key_cache   [num_blocks, Hkv, D/x, block_size, x]   x = 16 / element_size
  value_cache [num_blocks, Hkv, D, block_size]
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch


def kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int):
    """PagedAttention.get_kv_cache_shape (paged_attn.py:40-46)."""
    return (2, num_blocks, block_size * num_kv_heads * head_size)


def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int, head_size: int):
    """PagedAttention.split_kv_cache (paged_attn.py:49-62)."""
    x = 16 // kv_cache.element_size()
    nb = kv_cache.shape[1]
    key_cache = kv_cache[0].view(nb, num_kv_heads, head_size // x, -1, x)
    value_cache = kv_cache[1].view(nb, num_kv_heads, head_size, -1)
    return key_cache, value_cache


def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
    """csrc/cache_kernels.cu:152-205 / csrc/cpu/cache.cpp:36-83 and the python scatter
    of tests/kernels/test_cache.py:180-189.  key/value [N, Hkv, D]; slot < 0 = skip."""
    nb, H, Dx, bs, x = key_cache.shape
    slots = slot_mapping.flatten().to(torch.long)
    keep = slots >= 0
    if not bool(keep.any()):
        return
    idx = torch.nonzero(keep).flatten()
    slots = slots[idx]
    blk = torch.div(slots, bs, rounding_mode="floor")
    off = slots % bs
    k = key[idx].reshape(-1, H, Dx, x)
    v = value[idx]
    # duplicate live slots are undefined on the GPU (racing stores); the block manager
    # never produces them, so the oracle refuses them.
    assert slots.unique().numel() == slots.numel(), "duplicate slots"
    key_cache[blk, :, :, off, :] = k
    value_cache[blk, :, :, off] = v


def gather_kv(key_cache: torch.Tensor, value_cache: torch.Tensor,
              block_table: Sequence[int], kv_len: int):
    """De-page one sequence: -> K [kv_len, Hkv, D], V [kv_len, Hkv, D]
    (tests/kernels/test_attention.py:86-98)."""
    nb, H, Dx, bs, x = key_cache.shape
    nblk = -(-kv_len // bs)
    bt = torch.as_tensor(list(block_table[:nblk]), dtype=torch.long)
    k = key_cache[bt]                                   # [nblk,H,Dx,bs,x]
    k = k.permute(0, 3, 1, 2, 4).reshape(nblk * bs, H, Dx * x)[:kv_len]
    v = value_cache[bt]                                 # [nblk,H,D,bs]
    v = v.permute(0, 3, 1, 2).reshape(nblk * bs, H, -1)[:kv_len]
    return k, v


def depage_v_block(value_cache: torch.Tensor, blk: int) -> torch.Tensor:
    """-> [bs, Hkv, D] (the TRUE layout; the reference's view at
    torch_cuda_sdpa.py:1521-1523 is wrong)."""
    return value_cache[blk].permute(2, 0, 1).contiguous()


def paged_attention(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                    block_tables, kv_lens, scale: float,
                    out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Single-query paged attention, numerics of csrc/attention/attention_kernels.cuh:
    fp32 logits (:295-301), softmax fp32 with inv = 1/(sum + 1e-6) (:366-384),
    probabilities cast to the cache dtype before P.V (:430-470), fp32 accumulate.
    Structure follows tests/kernels/test_attention.py:62-119.
    query [B, Hq, D] -> out [B, Hq, D]."""
    B, Hq, D = query.shape
    Hkv = value_cache.shape[1]
    G = Hq // Hkv
    out = torch.empty(B, Hq, D, dtype=torch.float32)
    bts = block_tables.tolist() if torch.is_tensor(block_tables) else block_tables
    lens = kv_lens.tolist() if torch.is_tensor(kv_lens) else kv_lens
    for b in range(B):
        L = int(lens[b])
        if L == 0:
            out[b] = 0
            continue
        k, v = gather_kv(key_cache, value_cache, bts[b], L)
        kf = k.float().repeat_interleave(G, dim=1)        # [L, Hq, D]
        vq = v.repeat_interleave(G, dim=1)
        logits = scale * torch.einsum("hd,lhd->hl", query[b].float(), kf)
        m = logits.max(dim=-1, keepdim=True).values
        e = torch.exp(logits - m)
        p = e * (1.0 / (e.sum(dim=-1, keepdim=True) + 1e-6))
        p = p.to(v.dtype).float()
        out[b] = torch.einsum("hl,lhd->hd", p, vq.float())
    return out.to(out_dtype or query.dtype)


def prefill_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                      seq_lens_q: Sequence[int], seq_lens_kv: Sequence[int],
                      scale: float, out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """_run_sdpa_forward (torch_cuda_sdpa.py:1069-1127): per request
    scaled_dot_product_attention(q[Lq], k[Lkv], v[Lkv], is_causal=True) with GQA by
    repeat_interleave.  torch's is_causal mask is TOP-LEFT aligned when Lq != Lkv
    (SURVEY.md 0.7): query i sees kv j <= i.  Evaluated here in fp32 with that mask
    written out explicitly.
    query [sum Lq, Hq, D]; key/value [sum Lkv, Hkv, D]."""
    Hq, Hkv = query.shape[1], key.shape[1]
    G = Hq // Hkv
    out = torch.empty(query.shape, dtype=torch.float32)
    sq = skv = 0
    for Lq, Lkv in zip(seq_lens_q, seq_lens_kv):
        q = query[sq:sq + Lq].float().transpose(0, 1)                       # [Hq,Lq,D]
        k = key[skv:skv + Lkv].float().repeat_interleave(G, 1).transpose(0, 1)
        v = value[skv:skv + Lkv].float().repeat_interleave(G, 1).transpose(0, 1)
        s = scale * (q @ k.transpose(1, 2))                                  # [Hq,Lq,Lkv]
        i = torch.arange(Lq)[:, None]
        j = torch.arange(Lkv)[None, :]
        s = s.masked_fill(j > i, float("-inf"))
        p = torch.softmax(s, dim=-1)
        out[sq:sq + Lq] = (p @ v).transpose(0, 1)
        sq += Lq
        skv += Lkv
    return out.to(out_dtype or query.dtype)
