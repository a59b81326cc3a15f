// value_l2 block score, per kv head, with the FIXED fp32 reduction order that
// oracle/scoring.py restates bit-for-bit.
//
// Reference formula: KVCachePruner.get_block_score, "value_l2"
// (vllm/attention/kvcache_prunner.py:51-56): sum_t ||V[t,h,:]||_2 per head, then the
// mean over heads (done by the consumer of the per-head partials, evict_select.cu).
//
//   lane_j = sum_{i<D/8} v[8i+j]^2   sequential in i, fmaf(v,v,acc), acc0 = 0
//   sumsq  = ((l0+l1)+(l2+l3)) + ((l4+l5)+(l6+l7))          (xor-1, xor-2, xor-4 butterfly)
//   norm_t = sqrt_rn(sumsq)
//   head   = sum_{t<BS} norm_t       sequential in t, acc0 = 0
//
// One warp computes one (block, head).  `get(t, d)` returns V[t, d] as float.
#pragma once
#include "pc_common.cuh"
#include "pc_api_internal.h"

namespace pc {

template <int D, int BS, typename Get>
__device__ __forceinline__ float warp_value_l2_head_score(Get get) {
  const int lane = threadIdx.x & 31;
  const int j = lane & 7;
  const int tq = lane >> 3;  // 0..3
  constexpr int TPL = BS / 4;  // tokens per lane
  float norm[TPL];
#pragma unroll
  for (int k = 0; k < TPL; ++k) {
    const int t = tq + 4 * k;
    float acc = 0.f;
#pragma unroll 4
    for (int i = 0; i < D / 8; ++i) {
      const float v = get(t, 8 * i + j);
      acc = __fmaf_rn(v, v, acc);
    }
    acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 1));
    acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 2));
    acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 4));
    norm[k] = __fsqrt_rn(acc);
  }
  float head = 0.f;
#pragma unroll
  for (int t = 0; t < BS; ++t) {
    const float nt = __shfl_sync(kFull, norm[t / 4], (t % 4) * 8);
    head = __fadd_rn(head, nt);
  }
  return head;  // identical in every lane
}

// Publish one per-head partial to every rank's pool (peer stores over NVLink when
// n_pools > 1; they become visible to the peers at the next inter-rank barrier).
__device__ __forceinline__ void score_store(const ScorePool& sp, long long block_id, int h_local, float v) {
  const size_t idx = (size_t)block_id * sp.heads_total + sp.head_offset + h_local;
#pragma unroll
  for (int i = 0; i < PC_MAX_PEERS; ++i)
    if (i < sp.n_pools) sp.pools[i][idx] = v;
}

}  // namespace pc
