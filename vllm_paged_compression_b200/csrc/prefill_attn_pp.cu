// Prefill attention (A1), second tcgen05 kernel: ONE CTA per SM works on TWO query tiles that
// share their K/V tiles ("ping-pong"), kv tiles are 128 rows wide.
//
// prefill_attn_tc.cu runs two independent CTAs per SM, each one 128-query tile over 64-wide kv
// tiles: two dependent chains (S MMA -> softmax -> P.V) whose per-tile fixed work (barrier round
// trips, TMEM loads / stores, MMA issue) is paid every 64 columns, and nothing is saturated (ncu:
// tensor pipe 42 %, MUFU 44 %).  Here the two chains live in one CTA:
//   * work item = (request, kv head, query tile, PAIR of q heads of that kv head): both tiles have
//     the same rows, the same causal mask and the same K/V tiles, which are loaded once;
//   * kv tile = 128 rows: half the per-tile fixed work, M128 x N128 MMAs;
//   * TMEM (512 columns): S0 | S1 | O0 | O1, 128 columns each (D = 128); P_i (bf16) is written
//     over the first 64 columns of S_i and read by P.V as a TMEM operand;
//   * warpgroup 0 / 1 = softmax of tile 0 / 1 (thread = query row, 128 columns per step),
//     warpgroup 2 = TMA producer warp + MMA issuer warp, warpgroup 3 = epilogue (O / l -> global);
//     registers follow the roles (setmaxnreg);
//   * the MMA issuer interleaves the two tiles: S0(j) S1(j) | PV0(j) S0(j+1) | PV1(j) S1(j+1) | ...
//     so that the tensor pipe works for tile 1 while warpgroup 0 is in its softmax and vice versa.
// Semantics, numerics, masks, the TMA box gather of the streamingLLM prune, the lazy rescale and
// the item order are those of prefill_attn_tc.cu (same tests).
#include <cuda.h>

#include "pc_api_internal.h"
#include "tcgen05.cuh"

namespace pc {

namespace {

constexpr int kTM = 128;        // query rows per tile
constexpr int kTN = 128;        // kv rows per tile
constexpr int kPpThreads = 512;
constexpr int kRegsSoftmaxPp = 176, kRegsRolesPp = 88, kRegsEpiPp = 72;   // 128 x (176 + 176 + 88 + 72) = 64 K
constexpr float kRescaleThresholdPp = 8.f;
constexpr int kBoxRowsPp = 16;

struct PpParams {
  __nv_bfloat16* __restrict__ out;
  const int* __restrict__ req_meta;  // [R][6] {q_start, q_len, kv_src_start, kv_len, first, pruned}
  int Hq, Hkv;
  long long out_stride;
  float scale_log2;
};

template <int D>
struct PpSmem {
  static constexpr int kQ = kTM * D * 2;         // one Q tile
  static constexpr int kKV = kTN * D * 2;        // one K or V stage
  static constexpr int oQ = 0;                   // 2 tiles
  static constexpr int oK = oQ + 2 * kQ;         // 2 stages
  static constexpr int oV = oK + 2 * kKV;        // 2 stages
  static constexpr int oBar = oV + 2 * kKV;
  static constexpr int oLinv = oBar + 32 * 8;    // [item parity][tile][kTM] floats
  static constexpr int kBytes = oLinv + 2 * 2 * kTM * 4;
  static constexpr int kTmemCols = 512;
};

enum {
  P_QFULL = 0, P_KFULL = 1, P_VFULL = 3, P_KFREE = 5, P_VFREE = 7, P_SFULL = 9, P_PFULL = 11, P_PVDONE = 13,
  P_LFULL = 15 /* [tile][item parity] */, P_OFREE = 19, P_COUNT = 21
};

struct PpItem {
  int q_start, q_len, kv_src, kv_len, first, pruned, i0, hq0, hkv, ntiles;
};

template <int D>
__global__ void __launch_bounds__(kPpThreads, 1)
prefill_attn_pp_kernel(const PpParams p, const __grid_constant__ CUtensorMap tm_q,
                       const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                       const int q_tiles, const int num_reqs, const int chunk_reqs) {
  using L = PpSmem<D>;
  constexpr int KS = D / 16;       // k-steps of S = Q K^T
  constexpr int PK = kTN / 16;     // k-steps of O += P V
  constexpr uint32_t kIdescS = t5::instr_desc_bf16(kTM, kTN, false);
  constexpr uint32_t kIdescPV = t5::instr_desc_bf16(kTM, D, true);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + P_COUNT);
  float* linv = reinterpret_cast<float*>(smem + L::oLinv);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = p.Hq / p.Hkv;
  const int pairs = p.Hq / 2;                      // items per (request, query tile)
  const int total = q_tiles * pairs * num_reqs;

  // item k -> (chunk of requests, query tile descending, request, head pair): the order of
  // prefill_attn_tc.cu with pairs of q heads in the place of q heads
  auto get_item = [&](int k) {
    PpItem it;
    const int per_chunk = q_tiles * pairs * chunk_reqs;
    const int c = k / per_chunk, k2 = k - c * per_chunk;
    const int r0 = c * chunk_reqs;
    const int pq = pairs * min(chunk_reqs, num_reqs - r0);
    const int qt = q_tiles - 1 - k2 / pq;
    const int rem = k2 % pq;
    const int r = r0 + rem / pairs;
    it.hq0 = 2 * (rem % pairs);
    it.hkv = it.hq0 / G;
    const int* m = p.req_meta + r * 6;
    it.q_start = m[0]; it.q_len = m[1]; it.kv_src = m[2]; it.kv_len = m[3]; it.first = m[4]; it.pruned = m[5];
    it.i0 = qt * kTM;
    it.ntiles = it.i0 < it.q_len ? (min(it.kv_len, it.i0 + kTM) + kTN - 1) / kTN : 0;
    return it;
  };
  auto issue_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* full, const PpItem& it, int t) {
    mbar_expect_tx(full, L::kKV);
#pragma unroll
    for (int g = 0; g < kTN / kBoxRowsPp; ++g) {
      const int j = t * kTN + g * kBoxRowsPp;
      const int src = it.kv_src + (j < it.first ? j : j + it.pruned);
#pragma unroll
      for (int b = 0; b < D / 64; ++b)
        t5::tma_load_2d(dst + b * (kTN * 128) + g * (kBoxRowsPp * 128), tm, it.hkv * D + b * 64, src, full);
    }
  };

  if (tid == 0) {
    if ((smem_u32(smem) & 1023u) != 0u) __trap();
    mbar_init(&bar[P_QFULL], 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bar[P_KFULL + s], 1); mbar_init(&bar[P_VFULL + s], 1);
      mbar_init(&bar[P_KFREE + s], 1); mbar_init(&bar[P_VFREE + s], 1);
      mbar_init(&bar[P_SFULL + s], 1); mbar_init(&bar[P_PFULL + s], kTM); mbar_init(&bar[P_PVDONE + s], 1);
      mbar_init(&bar[P_LFULL + 2 * s], kTM); mbar_init(&bar[P_LFULL + 2 * s + 1], kTM);
      mbar_init(&bar[P_OFREE + s], kTM);
    }
    mbar_fence_init();
  }
  if (warp == 9) t5::tmem_alloc<L::kTmemCols>(tmem_slot);
  t5::fence_before_sync();
  __syncthreads();
  t5::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_s0 = tmem;                    // S_i at tmem + i * kTN
  const uint32_t tmem_o0 = tmem + 2 * kTN;          // O_i at tmem_o0 + i * D

  if (warp >= 8 && warp < 12) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsRolesPp));
    if (warp == 8) {
      // ================================================= TMA producer (one lane): Q pair, K and V tiles
      if (lane == 0) {
        int g = 0, n_item = 0;
        for (int k = blockIdx.x; k < total; k += gridDim.x) {
          const PpItem it = get_item(k);
          if (it.ntiles == 0) continue;
          if (it.pruned != 0 && (((it.first | it.pruned) & 15) != 0)) __trap();
          // both Q buffers are free once the previous item's last S1 MMA has completed (= k_free of its stage)
          if (g > 0) t5::mbar_wait_relaxed(&bar[P_KFREE + ((g - 1) & 1)], ((g - 1) >> 1) & 1);
          mbar_expect_tx(&bar[P_QFULL], 2 * L::kQ);
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int b = 0; b < D / 64; ++b)
              t5::tma_load_2d(smem + L::oQ + i * L::kQ + b * (kTM * 128), &tm_q, (it.hq0 + i) * D + b * 64,
                              it.q_start + it.i0, &bar[P_QFULL]);
          for (int t = 0; t < it.ntiles; ++t, ++g) {
            const int st = g & 1;
            if (g >= 2) t5::mbar_wait_relaxed(&bar[P_KFREE + st], ((g - 2) >> 1) & 1);
            issue_kv(&tm_k, smem + L::oK + st * L::kKV, &bar[P_KFULL + st], it, t);
            if (g >= 2) t5::mbar_wait_relaxed(&bar[P_VFREE + st], ((g - 2) >> 1) & 1);
            issue_kv(&tm_v, smem + L::oV + st * L::kKV, &bar[P_VFULL + st], it, t);
          }
          ++n_item;
        }
      }
      __syncwarp();
    } else if (warp == 9) {
      // ================================================= MMA issuer (one lane)
      if (lane == 0) {
        const uint32_t sQ = smem_u32(smem + L::oQ), sK = smem_u32(smem + L::oK), sV = smem_u32(smem + L::oV);
        auto issue_s = [&](int i, int st) {   // S_i = Q_i K^T of the K tile in stage st
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            const uint64_t a = t5::smem_desc_sw128(sQ + i * L::kQ + (ks >> 2) * (kTM * 128) + (ks & 3) * 32, 16, 1024);
            const uint64_t b = t5::smem_desc_sw128(sK + st * L::kKV + (ks >> 2) * (kTN * 128) + (ks & 3) * 32, 16, 1024);
            t5::mma_f16_ss(tmem_s0 + i * kTN, a, b, kIdescS, ks > 0);
          }
          t5::commit(&bar[P_SFULL + i]);
        };
        auto issue_pv = [&](int i, int st, bool acc) {   // O_i (+)= P_i V of the V tile in stage st
#pragma unroll
          for (int kk = 0; kk < PK; ++kk) {
            const uint64_t b = t5::smem_desc_sw128(sV + st * L::kKV + kk * (16 * 128), kTN * 128, 1024);
            t5::mma_f16_ts(tmem_o0 + i * D, tmem_s0 + i * kTN + kk * 8, b, kIdescPV, (acc || kk > 0));
          }
          t5::commit(&bar[P_PVDONE + i]);
        };
        int g = 0, n_item = 0;
        for (int k = blockIdx.x; k < total; k += gridDim.x) {
          const PpItem it = get_item(k);
          if (it.ntiles == 0) continue;
          mbar_wait(&bar[P_QFULL], n_item & 1);
          ++n_item;
          // first tile of the item: both S
          mbar_wait(&bar[P_KFULL + (g & 1)], (g >> 1) & 1);
          t5::fence_after_sync();
          issue_s(0, g & 1);
          issue_s(1, g & 1);
          t5::commit(&bar[P_KFREE + (g & 1)]);
          for (int j = 0; j < it.ntiles; ++j, ++g) {
            const int st = g & 1;
            const bool more = j + 1 < it.ntiles;
            mbar_wait(&bar[P_VFULL + st], (g >> 1) & 1);
            if (more) mbar_wait(&bar[P_KFULL + (st ^ 1)], ((g + 1) >> 1) & 1);
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              mbar_wait(&bar[P_PFULL + i], g & 1);
              // the first P.V of an item overwrites O_i: the epilogue must have read the previous item's
              if (j == 0 && n_item > 1) mbar_wait(&bar[P_OFREE + i], (n_item - 2) & 1);
              t5::fence_after_sync();
              issue_pv(i, st, j > 0);
              if (i == 1) t5::commit(&bar[P_VFREE + st]);
              if (more) {
                issue_s(i, st ^ 1);       // S_i(j+1) follows P.V_i(j) in the pipe: it may overwrite P_i(j)
                if (i == 1) t5::commit(&bar[P_KFREE + (st ^ 1)]);
              }
            }
          }
        }
      }
      __syncwarp();
    }
  } else if (warp < 8) {
    // ================================================= softmax warpgroups (tile i = warp / 4, thread = query row)
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmaxPp));
    const int i = warp >> 2;
    const int rt = tid & 127;                          // row of the tile = TMEM lane
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tmem_s = tmem_s0 + i * kTN + lane_base;
    const uint32_t tmem_o = tmem_o0 + i * D + lane_base;
    int g = 0, n_item = 0;
    for (int k = blockIdx.x; k < total; k += gridDim.x) {
      const PpItem it = get_item(k);
      if (it.ntiles == 0) continue;
      const int row = it.i0 + rt;
      float m_use = -INFINITY, l_run = 0.f;
      for (int j = 0; j < it.ntiles; ++j, ++g) {
        const int j0 = j * kTN;
        mbar_wait(&bar[P_SFULL + i], g & 1);
        t5::fence_after_sync();
        uint32_t sr[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) t5::tmem_ld32(sr[c], tmem_s + c * 32);
        t5::tmem_ld_wait();
        const bool need_mask = (j0 + kTN - 1 > it.i0) || (j0 + kTN > it.kv_len);
        if (need_mask) {
          const int lim = min(row, it.kv_len - 1) - j0;  // columns > lim are masked
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (__all_sync(kFull, lim >= c * 32 + 31)) continue;
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (c * 32 + e > lim) sr[c][e] = 0xff800000u;  // -inf
          }
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e) mx[e & 3] = fmaxf(mx[e & 3], __uint_as_float(sr[c][e]));
        const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
        float alpha = 1.f;
        const bool grow = m_tile > m_use + kRescaleThresholdPp || m_use == -INFINITY;
        if (grow) {
          alpha = (m_use == -INFINITY) ? 0.f : fast_exp2(m_use - m_tile);
          m_use = m_tile;
          l_run *= alpha;
        }
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        const float neg_m = -m_use;
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            const float p0 = fast_exp2(fmaf(__uint_as_float(sr[c][e]), p.scale_log2, neg_m));
            const float p1 = fast_exp2(fmaf(__uint_as_float(sr[c][e + 1]), p.scale_log2, neg_m));
            sum[(e >> 1) & 3] += p0 + p1;
            sr[c][e >> 1] = pack_bf16x2(p0, p1);
          }
        l_run += (sum[0] + sum[1]) + (sum[2] + sum[3]);
        // only a rescale of O has to wait for the previous P.V of this tile
        if (j > 0 && __any_sync(kFull, grow)) {
          mbar_wait(&bar[P_PVDONE + i], (g - 1) & 1);
          t5::fence_after_sync();
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            t5::tmem_ld32(o, tmem_o + c * 32);
            t5::tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
            t5::tmem_st32(tmem_o + c * 32, o);
          }
          t5::tmem_st_wait();
        }
        // packed P row -> TMEM over the scores it came from: 64 columns = 128 bf16
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int e = 0; e < 16; ++e) { pk[e] = sr[2 * h][e]; pk[16 + e] = sr[2 * h + 1][e]; }
          t5::tmem_st32(tmem_s + h * 32, pk);
        }
        t5::tmem_st_wait();
        t5::fence_before_sync();
        t5::mbar_arrive(&bar[P_PFULL + i]);
      }
      // l_full is double-buffered by item parity: a one-tile item lets this warpgroup finish item
      // n + 1 before the epilogue has looked at item n, and a single barrier would then be two
      // phases ahead of its waiter (parity aliasing)
      linv[((n_item & 1) * 2 + i) * kTM + rt] = 1.f / l_run;
      t5::mbar_arrive(&bar[P_LFULL + 2 * i + (n_item & 1)]);
      ++n_item;
    }
  } else {
    // ================================================= epilogue warpgroup: O_i / l -> bf16 -> global, i = 0, 1
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsEpiPp));
    const int rt = tid & 127;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    int g = 0, n_item = 0;
    for (int k = blockIdx.x; k < total; k += gridDim.x) {
      const PpItem it = get_item(k);
      if (it.ntiles == 0) continue;
      g += it.ntiles;
      const int row = it.i0 + rt;
#pragma unroll 1
      for (int i = 0; i < 2; ++i) {
        t5::mbar_wait_sleep(&bar[P_LFULL + 2 * i + (n_item & 1)], (n_item >> 1) & 1, 128);
        const float inv = linv[((n_item & 1) * 2 + i) * kTM + rt];
        t5::mbar_wait_sleep(&bar[P_PVDONE + i], (g - 1) & 1, 64);
        t5::fence_after_sync();
        __nv_bfloat16* dst = p.out + (size_t)(it.q_start + (row < it.q_len ? row : 0)) * p.out_stride +
                             (size_t)(it.hq0 + i) * D;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t o[32];
          t5::tmem_ld32(o, tmem_o0 + i * D + lane_base + c * 32);
          t5::tmem_ld_wait();
          if (c == D / 32 - 1) {          // O_i is in registers: the next item's first P.V may overwrite it
            t5::fence_before_sync();
            t5::mbar_arrive(&bar[P_OFREE + i]);
          }
          if (row < it.q_len) {
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              uint4 val;
              val.x = pack_bf16x2(__uint_as_float(o[q4 * 8 + 0]) * inv, __uint_as_float(o[q4 * 8 + 1]) * inv);
              val.y = pack_bf16x2(__uint_as_float(o[q4 * 8 + 2]) * inv, __uint_as_float(o[q4 * 8 + 3]) * inv);
              val.z = pack_bf16x2(__uint_as_float(o[q4 * 8 + 4]) * inv, __uint_as_float(o[q4 * 8 + 5]) * inv);
              val.w = pack_bf16x2(__uint_as_float(o[q4 * 8 + 6]) * inv, __uint_as_float(o[q4 * 8 + 7]) * inv);
              *reinterpret_cast<uint4*>(dst + c * 32 + q4 * 8) = val;
            }
          }
        }
      }
      ++n_item;
    }
  }

  // ---- teardown
  t5::fence_before_sync();
  __syncthreads();
  if (warp == 9) {
    t5::fence_after_sync();
    t5::tmem_dealloc<L::kTmemCols>(tmem);
  }
}

}  // namespace

// ---- host
namespace {

typedef CUresult (*EncodeTiledFnPp)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);
EncodeTiledFnPp encode_tiled_fn_pp() {
  static EncodeTiledFnPp fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnPp>(sym);
  }
  return fn;
}
// [rows, heads*D] bf16 matrix, box = 64 columns x box_rows, SWIZZLE_128B, out-of-range rows read as zero
int make_rows_map_pp(CUtensorMap* m, const void* base, long long rows, int inner, long long stride, int box_rows) {
  EncodeTiledFnPp enc = encode_tiled_fn_pp();
  if (enc == nullptr) return set_error(PC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)stride * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(PC_ERR_CUDA, "cuTensorMapEncodeTiled failed (prefill q/k/v map)");
  return 0;
}

template <int D>
int launch_prefill_pp(const PpParams& p, const PrefillArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  if (int rc = make_rows_map_pp(&tq, a.q, a.num_tokens, a.Hq * D, a.q_stride, kTM)) return rc;
  if (int rc = make_rows_map_pp(&tk, a.k, a.num_tokens, a.Hkv * D, a.key_stride, kBoxRowsPp)) return rc;
  if (int rc = make_rows_map_pp(&tv, a.v, a.num_tokens, a.Hkv * D, a.value_stride, kBoxRowsPp)) return rc;
  const int smem = PpSmem<D>::kBytes;
  auto kern = prefill_attn_pp_kernel<D>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(prefill_pp)", e);
  const int q_tiles = (a.max_q_len + kTM - 1) / kTM;
  const long long total = (long long)q_tiles * (a.Hq / 2) * a.num_reqs;
  if (total > 0x7fffffffLL) return set_error(PC_ERR_INVALID, "prefill: too many query tiles");
  const int sms = device_sm_count();
  if (sms <= 0) return PC_ERR_CUDA;
  const int ctas = (int)(total < sms ? total : sms);
  // requests per chunk of the heavy-first order (see prefill_attn_tc.cu): K + V of a chunk in ~40 MB of L2
  const long long kv_req = (long long)a.max_q_len * a.Hkv * D * 4;
  long long chunk = (40LL << 20) / (kv_req > 0 ? kv_req : 1);
  if (chunk < 1) chunk = 1;
  if (chunk > a.num_reqs || kv_req * a.num_reqs <= (80LL << 20)) chunk = a.num_reqs;
  kern<<<ctas, kPpThreads, smem, stream>>>(p, tq, tk, tv, q_tiles, a.num_reqs, (int)chunk);
  return check_launch("prefill_attn_pp_kernel");
}

}  // namespace

bool prefill_attn_pp_supported(const PrefillArgs& a) {
  return (a.D == 64 || a.D == 128) && a.Hkv > 0 && a.Hq % a.Hkv == 0 && ((a.Hq / a.Hkv) % 2) == 0;
}

// arguments were validated by prefill_attn() (prefill_attn_tc.cu)
int prefill_attn_pp(const PrefillArgs& a, cudaStream_t stream) {
  PpParams p;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.req_meta = a.req_meta;
  p.Hq = a.Hq; p.Hkv = a.Hkv;
  p.out_stride = a.out_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  if (a.D == 128) return launch_prefill_pp<128>(p, a, stream);
  return launch_prefill_pp<64>(p, a, stream);
}

}  // namespace pc
