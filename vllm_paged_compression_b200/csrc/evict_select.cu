// On-device decode eviction: trigger test, victim selection (argmin of value_l2 block
// scores, lowest index wins ties), block-table compaction, kv-length update and
// score-row compaction -- one CTA per batch row, no host round trip.
//
// Replaces the per-layer Python loops of
//   _prune_decode_reqs_streamingLLM (vllm/attention/backends/torch_cuda_sdpa.py:1375-1433)
//   _prune_decode_reqs_topk         (:1478-1560, argmin :1536-1541)
//   _create_updated_block_tables    (:1562-1586; semantics of torch_cuda_sdpa_v1.py /
//                                    HEAD row 0)
//   BlockL2NormManager.get_updated_block_tables(is_rmv_l2norm=True)
//                                   (vllm/attention/block_l2norm_manager.py:95-106)
//   kv-len adjust                   (vllm/worker/model_runner.py:540-543)
// Integer / fp32-compare domain: bit-exact against oracle/policy.py.
#include <float.h>

#include "pc_common.cuh"
#include "pc_api_internal.h"

namespace pc {

struct EvictParams {
  const int* __restrict__ block_tables;
  long long table_stride;
  const int* __restrict__ n_blocks;
  const int* __restrict__ seq_lens;
  const int* __restrict__ kv_lens;
  pc_evict_policy pol;
  ScorePool score;
  int* out_tables;
  long long out_stride;
  int* out_kv_lens;
  int* out_victim;
  int pre_adjusted;
};

constexpr int kEvictThreads = 256;

__device__ __forceinline__ bool decode_trigger(const pc_evict_policy& pol, int seq_len, int n) {
  const int bs = pol.block_size;
  if (pol.prune_type == 0)  // budget: torch_cuda_sdpa.py:1390-1407,1491-1500
    return seq_len > pol.cache_budget && (seq_len % bs) == 0;
  // percentage / evict_freq: torch_cuda_sdpa.py:1310 (commented) + page_evict_kv_util_bak.py:69
  return (seq_len % (bs * pol.evict_freq)) == 0 &&
         n * bs > (pol.initial_blocks + pol.num_block_merge) * bs;
}

// One batch row: trigger test, victim, compacted row.  All threads of the CTA take part;
// returns the victim (-1 = row did not trigger).
// `given`: decode_method 2 only -- the victim the HOST chose for this row (HEAD as shipped draws it
// from Python's `random`, torch_cuda_sdpa.py:802,1504-1513); used iff the row triggers.
__device__ __forceinline__ int evict_row(const pc_evict_policy& pol, const int* __restrict__ row, int n, int seq_len,
                                         const float* __restrict__ pool, int HT, int* orow, long long out_stride,
                                         int given) {
  __shared__ float best_v[kEvictThreads / 32];
  __shared__ int best_i[kEvictThreads / 32];
  __shared__ int victim_s;
  const int tid = threadIdx.x;
  const bool trig = (n >= 3) && decode_trigger(pol, seq_len, n);

  int victim = -1;
  if (trig) {
    if (pol.decode_method == 0) {
      victim = 1;  // streamingLLM: always the block after the sink (torch_cuda_sdpa.py:1404-1405)
    } else if (pol.decode_method == 2) {
      victim = (given >= 1 && given < n) ? given : -1;   // host-chosen; an index outside the row evicts nothing
    } else {
      // candidates [1, hi): global -> n-1 ; local -> min(1 + topk, n-1)   (:1528-1534)
      const int lo = 1;
      int hi = n - 1;
      if (pol.topk_blocks > 0) hi = min(hi, 1 + pol.topk_blocks);
      float bv = FLT_MAX;
      int bi = 0x7fffffff;
      for (int i = lo + tid; i < hi; i += kEvictThreads) {
        const float* srow = pool + (size_t)row[i] * HT;  // scores live with the physical block
        float acc = 0.f;  // heads ascending, then / Hkv: oracle/scoring.py combine_heads
        for (int hh = 0; hh < HT; ++hh) acc = __fadd_rn(acc, __ldcg(srow + hh));
        const float sc = __fdiv_rn(acc, (float)HT);
        if (bi == 0x7fffffff || sc < bv) { bv = sc; bi = i; }  // strict <: lowest index wins
      }
      // (value, index) min-reduction; ties -> lower index
#pragma unroll
      for (int sh = 16; sh >= 1; sh >>= 1) {
        const float ov = __shfl_xor_sync(kFull, bv, sh);
        const int oi = __shfl_xor_sync(kFull, bi, sh);
        if (oi != 0x7fffffff && (bi == 0x7fffffff || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
      }
      if ((tid & 31) == 0) { best_v[tid >> 5] = bv; best_i[tid >> 5] = bi; }
      __syncthreads();
      if (tid == 0) {
        for (int w = 1; w < kEvictThreads / 32; ++w) {
          const float ov = best_v[w];
          const int oi = best_i[w];
          if (oi != 0x7fffffff && (bi == 0x7fffffff || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
        }
        victim_s = (bi == 0x7fffffff) ? lo : bi;
      }
      __syncthreads();
      victim = victim_s;
    }
  }

  // ---- table compaction: delete column `victim`, shift left, 0-pad
  const int n_after = n - (victim >= 0 ? 1 : 0);
  for (int i = tid; i < out_stride; i += kEvictThreads) {
    int v = 0;
    if (i < n_after) v = (victim >= 0 && i >= victim) ? row[i + 1] : row[i];
    orow[i] = v;
  }
  return victim;
}

__global__ void __launch_bounds__(kEvictThreads) evict_select_compact_kernel(const EvictParams p) {
  const int b = blockIdx.x;
  const int given = p.pol.decode_method == 2 ? p.out_victim[b] : -1;
  const int victim = evict_row(p.pol, p.block_tables + (size_t)b * p.table_stride, p.n_blocks[b], p.seq_lens[b],
                               p.score.n_pools > 0 ? p.score.pools[p.score.local] : nullptr, p.score.heads_total,
                               p.out_tables + (size_t)b * p.out_stride, p.out_stride, given);
  __syncthreads();   // every thread has read `given` before thread 0 overwrites it
  if (threadIdx.x == 0) {
    p.out_victim[b] = victim;
    if (p.out_kv_lens)
      p.out_kv_lens[b] = p.kv_lens[b] - ((victim >= 0 && !p.pre_adjusted) ? p.pol.block_size : 0);
  }
}

// Every layer of an engine step in ONE launch: grid (batch, layers).  A layer's victims depend
// only on the block scores written on EARLIER steps and on the step's tables, never on this
// step's attention, so the selections of all layers can run before the first layer's attention
// -- one launch instead of one per layer, and the attention kernels of a trigger step keep
// their programmatic dependent launch chain.
struct EvictLayersParams {
  const int* __restrict__ block_tables;
  long long table_stride;
  const int* __restrict__ n_blocks;
  const int* __restrict__ seq_lens;
  pc_evict_policy pol;
  const float* const* __restrict__ pools;   // [layers] this rank's score pool of every layer
  int heads_total;
  int* out_tables;                          // [layers][batch][out_stride]
  long long out_stride, out_layer_stride;
  int* out_victims;                         // [layers][batch]
};

__global__ void __launch_bounds__(kEvictThreads) evict_select_compact_layers_kernel(const EvictLayersParams p) {
  const int b = blockIdx.x, layer = blockIdx.y;
  const float* pool = p.pools != nullptr ? p.pools[layer] : nullptr;
  const int given = p.pol.decode_method == 2 ? p.out_victims[(size_t)layer * gridDim.x + b] : -1;
  const int victim = evict_row(p.pol, p.block_tables + (size_t)b * p.table_stride, p.n_blocks[b], p.seq_lens[b], pool,
                               p.heads_total, p.out_tables + (size_t)layer * p.out_layer_stride + (size_t)b * p.out_stride,
                               p.out_stride, given);
  __syncthreads();   // every thread has read `given` before thread 0 overwrites it
  if (threadIdx.x == 0) p.out_victims[(size_t)layer * gridDim.x + b] = victim;
}

int evict_select_compact_layers(const EvictLayersArgs& a, cudaStream_t stream) {
  if (a.B <= 0 || a.L <= 0) return 0;
  if (a.pol.block_size <= 0) return set_error(PC_ERR_INVALID, "policy.block_size must be > 0");
  if (a.pol.decode_method == 1 && (a.pools == nullptr || a.heads_total <= 0))
    return set_error(PC_ERR_INVALID, "value_l2 eviction needs the per-layer score pools");
  if (a.L > 65535) return set_error(PC_ERR_INVALID, "too many layers for one launch");
  EvictLayersParams p;
  p.block_tables = a.block_tables; p.table_stride = a.table_stride; p.n_blocks = a.n_blocks;
  p.seq_lens = a.seq_lens; p.pol = a.pol; p.pools = a.pools; p.heads_total = a.heads_total;
  p.out_tables = a.out_tables; p.out_stride = a.out_stride; p.out_layer_stride = a.out_layer_stride;
  p.out_victims = a.out_victims;
  evict_select_compact_layers_kernel<<<dim3(a.B, a.L), kEvictThreads, 0, stream>>>(p);
  return check_launch("evict_select_compact_layers_kernel");
}

int evict_select_compact(const EvictSelectArgs& a, cudaStream_t stream) {
  if (a.B <= 0) return 0;
  if (a.pol.block_size <= 0) return set_error(PC_ERR_INVALID, "policy.block_size must be > 0");
  if (a.pol.decode_method == 1 && a.score.n_pools <= 0)
    return set_error(PC_ERR_INVALID, "value_l2 eviction needs a score pool");
  EvictParams p;
  p.block_tables = a.block_tables;
  p.table_stride = a.table_stride;
  p.n_blocks = a.n_blocks;
  p.seq_lens = a.seq_lens;
  p.kv_lens = a.kv_lens;
  p.pol = a.pol;
  p.score = a.score;
  p.out_tables = a.out_tables;
  p.out_stride = a.out_stride;
  p.out_kv_lens = a.out_kv_lens;
  p.out_victim = a.out_victim;
  p.pre_adjusted = a.pre_adjusted;
  evict_select_compact_kernel<<<a.B, kEvictThreads, 0, stream>>>(p);
  return check_launch("evict_select_compact_kernel");
}

}  // namespace pc
