"""Tensor parallelism of the paged-eviction path: KV heads are sharded, block ids are not.

The reference shards attention by KV heads under TP (vllm/model_executor/models/llama.py:
117-130; per-rank KV pool [2, nb, bs * (Hkv/TP) * D], vllm/worker/cache_engine.py:35-39) and
broadcasts ONE set of block tables / slot mappings to every worker, so cache writes, table
compaction and attention need no communication.  The only cross-rank dependency the
eviction path adds is the value_l2 block score: it is a mean over ALL kv heads
(vllm/attention/kvcache_prunner.py:51-56), and every rank must pick the same victim.  In
the reference each rank would pick from its local heads and only rank 0 / layer 0 informs
the host (a latent divergence); here every rank holds every head's partials:

  score pool [num_blocks, Hkv_total] fp32 per layer and rank; rank r computes columns
  [r * Hkv_local, (r + 1) * Hkv_local) inside its own cache-write / decode kernels.

Two ways to fill the other columns (`mode`):
  "p2p"   the pools live in symmetric memory (torch.distributed._symmetric_memory, CUDA
          VMM handles exchanged once at allocation); the scoring kernels store each partial
          straight into EVERY rank's pool through NVLink peer addresses
          (csrc/score_common.cuh score_store) -- no extra launch, no NCCL call on the data
          path.  `publish()` is one device-side barrier so that a rank cannot run a whole
          fill interval ahead of its peers (see DESIGN.md "Multi-GPU" for the race argument).
  "nccl"  plain local pools; `publish()` all-gathers the rows scored this step (all layers
          in one collective) and scatters them into the local pools.  Baseline / fallback,
          and the path the gloo CPU tests exercise.
Victim selection sums heads in ascending GLOBAL order and divides by Hkv_total, so the
chosen victims are bit-identical to TP = 1 on every rank (tests/test_parallel.py).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_heads(num_heads: int, num_kv_heads: int, tp_rank: int, tp_size: int) -> Tuple[range, range]:
    """(q heads, kv heads) owned by `tp_rank`: contiguous kv heads and the q heads of those
    groups (llama.py:117-130).  KV-head replication (tp_size > num_kv_heads) is not built."""
    if num_kv_heads % tp_size != 0:
        raise ValueError(f"tensor parallel size {tp_size} must divide {num_kv_heads} kv heads")
    g = num_heads // num_kv_heads
    hl = num_kv_heads // tp_size
    kv = range(tp_rank * hl, (tp_rank + 1) * hl)
    return range(kv.start * g, kv.stop * g), kv


class ScoreExchange:
    """Owns the score pools of all layers of one rank and makes every rank's partials
    visible in every rank's pools."""

    def __init__(self, num_layers: int, num_kv_heads_local: int, group=None, mode: str = "p2p") -> None:
        if mode not in ("p2p", "nccl"):
            raise ValueError(f"unknown score exchange mode {mode!r}")
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.mode = mode
        self.num_layers = num_layers
        self.hl = num_kv_heads_local
        self.big: Optional[torch.Tensor] = None     # [L, num_blocks, Hkv_total]
        self._handle = None
        self._peer_base: Optional[List[int]] = None
        self._next_layer = 0

    # ------------------------------------------------------------------ allocation
    def _allocate(self, num_blocks: int, heads_total: int, device) -> None:
        shape = (self.num_layers, num_blocks, heads_total)
        if self.mode == "p2p":
            import torch.distributed._symmetric_memory as symm_mem
            t = symm_mem.empty(shape, dtype=torch.float32, device=device)
            t.zero_()
            name = self.group.group_name if hasattr(self.group, "group_name") else self.group
            self._handle = symm_mem.rendezvous(t, name)
            self._peer_base = [int(p) for p in self._handle.buffer_ptrs]
            if self._peer_base[self.rank] != t.data_ptr():
                raise RuntimeError("symmetric memory: local buffer pointer mismatch")
            self.big = t
            torch.cuda.synchronize(device)
            self._handle.barrier()
        else:
            self.big = torch.zeros(shape, dtype=torch.float32, device=device)

    def alloc_pool(self, num_blocks: int, heads_total: int, device):
        """-> (pool [num_blocks, heads_total] of the next layer, peer pointers or None).
        Called once per layer, in layer order, by TorchCUDASDPABackendImpl._scores."""
        if heads_total != self.hl * self.world:
            raise ValueError("heads_total must be num_kv_heads_local * world_size")
        if self.big is None:
            self._allocate(num_blocks, heads_total, device)
        if self._next_layer >= self.num_layers or self.big.shape[1] != num_blocks:
            raise RuntimeError("ScoreExchange: more pools requested than layers, or pool size changed")
        l = self._next_layer
        self._next_layer += 1
        pool = self.big[l]
        if self.mode == "p2p":
            off = l * num_blocks * heads_total * 4
            return pool, [b + off for b in self._peer_base]
        return pool, None

    # ------------------------------------------------------------------ per-step hook
    def publish(self, block_ids: Sequence[int]) -> None:
        """Called once per engine step after the last layer, with the physical ids of the
        blocks scored during the step (host-known: rows whose last block filled, or the
        kept full blocks of a prefill batch).  No-op when nothing was scored."""
        if self.big is None or len(block_ids) == 0 or self.world == 1:
            return
        if self.mode == "p2p":
            # the partials are already in every pool; order the ranks so that nobody enters
            # the next fill interval before all peers have finished this one
            self._handle.barrier()
            return
        ids = torch.as_tensor(list(block_ids), dtype=torch.long, device=self.big.device)
        c0 = self.rank * self.hl
        mine = self.big[:, ids, c0:c0 + self.hl].contiguous()              # [L, n, hl]
        gathered = torch.empty((self.world * mine.shape[0],) + tuple(mine.shape[1:]), dtype=mine.dtype,
                               device=mine.device)               # concatenated along dim 0 (gloo and nccl)
        dist.all_gather_into_tensor(gathered, mine, group=self.group)
        gathered = gathered.view((self.world,) + tuple(mine.shape))
        # [W, L, n, hl] -> [L, n, W * hl]: rank-major == ascending global head order
        self.big[:, ids, :] = gathered.permute(1, 2, 0, 3).reshape(mine.shape[0], mine.shape[1], -1)
