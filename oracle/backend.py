"""Layer-level oracle: TorchCUDASDPABackendImpl.forward restated on CPU tensors.
TEST INFRASTRUCTURE ONLY.

Follows vllm/attention/backends/torch_cuda_sdpa.py:812-1064 (forward),
1222-1296 (prompt prune), 1375-1433 / 1478-1560 (decode prune), 1562-1586 (table
compaction) with the declared deviations of DESIGN.md:
  D1  value_l2 is computed in fp32, fixed order, on the TRUE V layout (oracle/scoring.py)
  D2  a block is scored when it FILLS, after this step's cache write (the reference
      scores before the write, so slot bs-1 is stale) and every full prompt block is
      scored unconditionally
  D3  a score belongs to its PHYSICAL block (dict block id -> per-head partials) instead of
      living in a per-request python list indexed like the block table.  For layer 0 --
      the only layer whose victims reach the block manager (model_runner.py:1797-1802) --
      this equals v1's "delete the victim's score with the victim"
      (torch_cuda_sdpa_v1.py:1541; HEAD never deletes).  For layers
      >= 1, v1's list drifts by one position between its own victim and layer 0's after
      every step on which the layers disagree (the list loses the layer's victim, the
      table loses layer 0's); here every block keeps its own score.
  D4  compaction is v1 / HEAD-row-0 semantics for every row
  D5  streamingLLM trigger steps attend over the COMPACTED table.  HEAD/v1 compute it
      but never use it (`is_pruned_decode` vs `is_prune_decode` typo,
      torch_cuda_sdpa.py:903,910 vs 858,1046); `head_compat_uncompacted=True`
      reproduces that for the golden traces.
State is per layer (one OracleLayer per attention layer), like BlockL2NormManager.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import paged_ops, policy, scoring, token_prune
from .policy import Policy


class OracleLayer:
    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int,
                 pol: Optional[Policy], head_compat_uncompacted: bool = False, native_ops=None,
                 rng=None, head_mixing_repack: bool = True):
        """native_ops: object providing reshape_and_cache / paged_attention with the
        signatures of oracle/paged_ops.py (default: that torch restatement;
        oracle/ref_cpu.RefCpuOps = the reference's own compiled CPU kernels)."""
        self.Hq, self.D, self.scale, self.Hkv = num_heads, head_size, float(scale), num_kv_heads
        self.pol = pol
        self.ops = native_ops if native_ops is not None else paged_ops
        self.compat = head_compat_uncompacted
        # HEAD's process-wide `random` (torch_cuda_sdpa.py:796): ONE random.Random(1234) shared by
        # every layer of the model, drawn from in layer order, rows in row order
        self.rng = rng
        self.repack = head_mixing_repack
        # physical block id -> per-head partial scores [Hkv] fp32 (D3)
        self.block_scores: Dict[int, np.ndarray] = {}
        self.last = {}

    # -------------------------------------------------------------- prefill
    def prefill(self, query, key, value, kv_cache, seq_lens: Sequence[int],
                block_tables: Sequence[Sequence[int]], req_ids: Sequence[str],
                compute_attention: bool = True) -> Optional[torch.Tensor]:
        """query [sumL, Hq*D], key/value [sumL, Hkv*D]; block_tables = post-prune host
        tables.  Returns out [sumL, Hq*D] (None when compute_attention is False: cache
        write + prompt scoring only, used to set up decode state for timing)."""
        p, bs = self.pol, self.pol.block_size
        q = query.view(-1, self.Hq, self.D)
        k = key.view(-1, self.Hkv, self.D)
        v = value.view(-1, self.Hkv, self.D)
        kc, vc = paged_ops.split_kv_cache(kv_cache, self.Hkv, self.D)
        kept, slots, kps, vps = [], [], [], []
        s = 0
        tok_kind = {"token_value_l2": "value_l2", "token_inverse_key_l2": "inverse_key_l2"}.get(p.prompt_evict_method)
        for L, tab in zip(seq_lens, block_tables):
            if tok_kind is not None:      # f3: per-head token drop + the reference's re-pack
                pk, pv = token_prune.prune_prompt(k[s:s + L].float().numpy(), v[s:s + L].float().numpy(),
                                                  p.cache_budget, bs, tok_kind, repack=self.repack)
                kps.append(torch.from_numpy(pk).to(k.dtype))
                vps.append(torch.from_numpy(pv).to(v.dtype))
            else:
                gi = torch.from_numpy(policy.prompt_gather_index(L, p)) + s
                kps.append(k[gi])
                vps.append(v[gi])
            kept.append(kps[-1].shape[0])
            slots.extend(policy.prefill_slot_mapping(list(tab), L, bs))
            s += L
        kp, vp = torch.cat(kps), torch.cat(vps)
        assert len(slots) == kp.shape[0]
        self.ops.reshape_and_cache(kp, vp, kc, vc, torch.tensor(slots, dtype=torch.long))
        # scores of every FULL cached prompt block (D2)
        s = 0
        for L, tab in zip(kept, block_tables) if p.decode_evict_method == "value_l2" else ():
            vr = vp[s:s + L].float().numpy()
            for j in range(L // bs):
                self.block_scores[int(tab[j])] = scoring.per_head_block_score(vr[j * bs:(j + 1) * bs])
            s += L
        self.last = {"slots": slots, "kept": kept}
        if not compute_attention:
            return None
        if p.disable_evict_prefill:
            out = paged_ops.prefill_attention(q, k, v, seq_lens, seq_lens, self.scale)
        else:
            out = paged_ops.prefill_attention(q, kp, vp, seq_lens, kept, self.scale)
        return out.view(-1, self.Hq * self.D)

    # -------------------------------------------------------------- decode
    def decode(self, query, key, value, kv_cache, seq_lens: Sequence[int],
               kv_lens: Sequence[int], block_tables: Sequence[Sequence[int]],
               req_ids: Sequence[str]):
        """One decode step.  seq_lens = logical lengths incl. this token; kv_lens =
        physical lengths as the host passes them (already -= bs on trigger rows,
        model_runner.py:540-543); block_tables = host tables (not compacted).
        -> (out [B, Hq*D], rmv {row: (req_id, idx)}, attn_tables list-of-lists)"""
        p, bs = self.pol, self.pol.block_size
        B = len(seq_lens)
        kc, vc = paged_ops.split_kv_cache(kv_cache, self.Hkv, self.D)
        rmv: Dict[int, Tuple[str, int]] = {}
        attn_tables: List[List[int]] = []
        slots = []
        zero_slots = []
        stream_like = p.decode_evict_method in ("streamingLLM", "streamingLLM-1")
        for r in range(B):
            tab = list(block_tables[r])
            n = len(tab)
            slots.append(policy.decode_slot_mapping(tab, seq_lens[r] - 1, bs))
            zs = policy.zero_slot(self.rng, seq_lens[r], tab, p)          # W2
            if zs is not None:
                zero_slots.append(zs)
            if policy.decode_trigger(seq_lens[r], n * bs, p):
                lo, hi = policy.victim_candidates(n, p)
                if stream_like:
                    idx = 1
                elif p.decode_evict_method == "random":
                    idx = policy.random_victim(self.rng, n, p)
                else:
                    sc = scoring.combine_heads(np.stack([self.block_scores[int(t)] for t in tab[:n - 1]]))
                    idx = policy.argmin_lowest_index(sc, lo, hi)
                rmv[r] = (req_ids[r], idx)
                if not (self.compat and stream_like):
                    del tab[idx]
            attn_tables.append(tab)
        k = key.view(-1, self.Hkv, self.D)
        v = value.view(-1, self.Hkv, self.D)
        if zero_slots:     # torch_cuda_sdpa.py:919-933: zero rows appended to the step's K/V
            z = torch.zeros(len(zero_slots), self.Hkv, self.D, dtype=k.dtype)
            k = torch.cat([k, z])
            v = torch.cat([v, z])
        self.ops.reshape_and_cache(k, v, kc, vc, torch.tensor(slots + zero_slots, dtype=torch.long))
        out = self.ops.paged_attention(query.view(-1, self.Hq, self.D), kc, vc,
                                        attn_tables, list(kv_lens), self.scale)
        for r in range(B):                                                # D2
            if p.decode_evict_method == "value_l2" and policy.score_trigger(seq_lens[r], p):
                last_blk = block_tables[r][-1]
                vb = paged_ops.depage_v_block(vc, last_blk).float().numpy()
                self.block_scores[int(last_blk)] = scoring.per_head_block_score(vb)
        self.last = {"slots": slots, "zero_slots": zero_slots}
        return out.view(-1, self.Hq * self.D), rmv, attn_tables

    def scores_of(self, table: Sequence[int], n_full: int) -> np.ndarray:
        """[n_full, Hkv] per-head partials of the first n_full (full) blocks of a table."""
        if n_full == 0:
            return np.zeros((0, self.Hkv), np.float32)
        return np.stack([self.block_scores[int(t)] for t in table[:n_full]])
