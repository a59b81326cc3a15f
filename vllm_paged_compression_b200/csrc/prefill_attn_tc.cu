// Prefill attention (A1) on the 5th-gen tensor cores: tcgen05.mma with TMEM accumulators.
//
// Replaces _run_sdpa_forward (vllm/attention/backends/torch_cuda_sdpa.py:1069-1127): causal
// SDPA per request over the streamingLLM-pruned prompt with torch's TOP-LEFT aligned mask
// (query i sees kept kv j <= i), GQA by index arithmetic instead of repeat_interleave
// (:1078-1080), pruned K/V rows gathered on the fly (kvcache_prunner.py:203-215).
//
// One CTA = one 128-query tile of one (request, q head); 7 warps, warp-specialised:
//   warps 0-3  softmax: thread t owns query row t = TMEM lane t.  S row out of TMEM
//              (tcgen05.ld), online softmax in the exp2 domain, P (bf16) into shared
//              memory as the A operand of P.V, lazy rescale of O in TMEM, epilogue.
//   warp 4     K producer: Q tile once, then K tiles, cp.async gathers into the
//              no-swizzle core-matrix layout ([16-byte chunk][row], chunk-major).
//   warp 5     V producer (same layout; V is the MN-major B operand of P.V).
//   warp 6     MMA issuer (one lane): S_j = Q K_j^T into one of two TMEM S buffers,
//              O += P_j V_j into the TMEM O accumulator; tcgen05.commit -> mbarriers.
// S_{j+1} is issued before P_j.V_j so the tensor pipe works on the next scores while the
// softmax warps are busy with tile j.  TMEM: 2 x 128 cols (S) + D cols (O) of 512.
//
// O stays in TMEM for the whole row of tiles (hardware accumulate).  When the running max
// of a row grows by more than 2^8 the warp rescales its 32 O rows in TMEM (ld, mul, st);
// smaller increases keep the old reference max -- P and l stay consistent, the result is
// mathematically identical and P stays <= 256 (exact in fp32, same relative bf16 rounding).
#include "pc_api_internal.h"
#include "tcgen05.cuh"

namespace pc {

namespace {

constexpr int kTM = 128;       // query rows per CTA (UMMA M)
constexpr int kTcThreads = 224;
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct PrefillTcParams {
  const __nv_bfloat16* __restrict__ q;
  const __nv_bfloat16* __restrict__ k;
  const __nv_bfloat16* __restrict__ v;
  __nv_bfloat16* __restrict__ out;
  const int* __restrict__ req_meta;  // [R][6] {q_start, q_len, kv_src_start, kv_len, first, pruned}
  int Hq, Hkv;
  long long q_stride, out_stride, key_stride, value_stride;
  float scale_log2;
};

// TN = kv rows per tile (UMMA N of S, K extent of P.V).  TN = 64 fits two CTAs per SM
// (112 KB of shared memory and 256 TMEM columns each), TN = 128 one.
template <int D, int TN>
struct TcSmem {
  static constexpr int kTN = TN;
  static constexpr int kTmemCols = (2 * TN + D) <= 256 ? 256 : 512;
  static constexpr int kQ = kTM * D * 2;
  static constexpr int kKV = kTN * D * 2;      // one K or V stage
  static constexpr int kP = kTM * kTN * 2;
  static constexpr int oQ = 0;
  static constexpr int oK = oQ + kQ;           // 2 stages
  static constexpr int oV = oK + 2 * kKV;      // 2 stages
  static constexpr int oP = oV + 2 * kKV;
  static constexpr int oBar = oP + kP;
  // barriers: q_full, k_full[2], v_full[2], s_full[2], p_full, pv_done  (9) + tmem slot
  static constexpr int kBytes = oBar + 16 * 8;
};

enum { B_QFULL = 0, B_KFULL = 1, B_VFULL = 3, B_SFULL = 5, B_PFULL = 7, B_PVDONE = 8, B_COUNT = 9 };

// One warp gathers a [rows x D] bf16 tile into chunk-major core-matrix layout:
//   byte(row, chunk) = chunk * (ROWS * 16) + row * 16      (chunk = 8 elements along D)
// Lane mapping: the L1TEX tag stage coalesces only ADJACENT lanes (ncu: the first mapping,
// lane -> (row % 8, chunk / 8), cost 32 tag wavefronts per LDGSTS and made the kernel
// LSU-bound at 87 % of l1tex__t_output_wavefronts).  kGatherMap 1: lane -> (chunk % 4, row)
// = 8 rows x 64 B per instruction; 2: lane -> (chunk % 8, row) = 4 rows x 128 B.
#ifndef PC_GATHER_MAP
#define PC_GATHER_MAP 2
#endif
template <int D, int ROWS, typename RowSrc>
__device__ __forceinline__ void gather_tile(uint8_t* dst, int lane, RowSrc row_src) {
  constexpr int CH = D / 8;
#if PC_GATHER_MAP == 0
  constexpr int RPI = 8, CPI = 4;
  const int rl = lane & 7, cl = lane >> 3;
#elif PC_GATHER_MAP == 1
  constexpr int RPI = 8, CPI = 4;
  const int rl = lane >> 2, cl = lane & 3;
#else
  constexpr int RPI = 4, CPI = 8;
  const int rl = lane >> 3, cl = lane & 7;
#endif
#pragma unroll 4
  for (int rg = 0; rg < ROWS / RPI; ++rg) {
    const int row = rg * RPI + rl;
    bool ok;
    const __nv_bfloat16* src = row_src(row, ok);
#pragma unroll
    for (int cq = 0; cq < CH / CPI; ++cq) {
      const int c = cq * CPI + cl;
      cp_async_16_zfill(dst + c * (ROWS * 16) + row * 16, src + c * 8, ok);
    }
  }
}

}  // namespace

namespace {

template <int D, int TN>
__global__ void __launch_bounds__(kTcThreads, (TN == 64 ? 2 : 1))
prefill_attn_tc_kernel(const PrefillTcParams p) {
  using L = TcSmem<D, TN>;
  constexpr int kTN = TN;
  constexpr int kTmemCols = L::kTmemCols;
  constexpr int SC = TN / 32;       // 32-column chunks of an S row
  constexpr int KS = D / 16;        // k-steps of S = Q K^T
  constexpr int PK = kTN / 16;      // k-steps of O += P V
  constexpr uint32_t kIdescS = t5::instr_desc_bf16(kTM, kTN, false);
  constexpr uint32_t kIdescPV = t5::instr_desc_bf16(kTM, D, true);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + B_COUNT);

  const int r = blockIdx.z;
  const int hq = blockIdx.y;
  const int* m = p.req_meta + r * 6;
  const int q_start = m[0], q_len = m[1], kv_src = m[2], kv_len = m[3], first = m[4], pruned = m[5];
  // heaviest query tiles (most visible kv) are scheduled first
  const int i0 = (int)(gridDim.x - 1 - blockIdx.x) * kTM;
  if (i0 >= q_len) return;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kv_hi = min(kv_len, i0 + kTM);
  const int ntiles = (kv_hi + kTN - 1) / kTN;

  if (tid == 0) {
    mbar_init(&bar[B_QFULL], 32);
    mbar_init(&bar[B_KFULL], 32); mbar_init(&bar[B_KFULL + 1], 32);
    mbar_init(&bar[B_VFULL], 32); mbar_init(&bar[B_VFULL + 1], 32);
    mbar_init(&bar[B_SFULL], 1);  mbar_init(&bar[B_SFULL + 1], 1);
    mbar_init(&bar[B_PFULL], kTM);
    mbar_init(&bar[B_PVDONE], 1); mbar_init(&bar[B_PVDONE + 1], 1);
    mbar_fence_init();
  }
  if (warp == 6) t5::tmem_alloc<kTmemCols>(tmem_slot);
  t5::fence_before_sync();
  __syncthreads();
  t5::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_o = tmem + 2 * kTN;

  if (warp == 4) {
    // ================================================= Q + K producer
    gather_tile<D, kTM>(smem + L::oQ, lane, [&](int row, bool& ok) {
      ok = (i0 + row) < q_len;
      return p.q + (size_t)(q_start + (ok ? i0 + row : 0)) * p.q_stride + (size_t)hq * D;
    });
    t5::cp_async_arrive_noinc(&bar[B_QFULL]);
    for (int t = 0; t < ntiles; ++t) {
      const int st = t & 1;
      if (t >= 2) t5::mbar_wait_relaxed(&bar[B_SFULL + st], ((t - 2) >> 1) & 1);  // S_{t-2} done
      const int j0 = t * kTN;
      gather_tile<D, kTN>(smem + L::oK + st * L::kKV, lane, [&](int row, bool& ok) {
        const int j = j0 + row;
        ok = j < kv_len;
        const long long src = (long long)kv_src + (ok ? (j < first ? j : j + pruned) : 0);
        return p.k + (size_t)src * p.key_stride + (size_t)hkv * D;
      });
      t5::cp_async_arrive_noinc(&bar[B_KFULL + st]);
    }
  } else if (warp == 5) {
    // ================================================= V producer
    for (int t = 0; t < ntiles; ++t) {
      const int st = t & 1;
      if (t >= 2) t5::mbar_wait_relaxed(&bar[B_PVDONE + st], ((t - 2) >> 1) & 1);  // P.V_{t-2} done
      const int j0 = t * kTN;
      gather_tile<D, kTN>(smem + L::oV + st * L::kKV, lane, [&](int row, bool& ok) {
        const int j = j0 + row;
        ok = j < kv_len;
        const long long src = (long long)kv_src + (ok ? (j < first ? j : j + pruned) : 0);
        return p.v + (size_t)src * p.value_stride + (size_t)hkv * D;
      });
      t5::cp_async_arrive_noinc(&bar[B_VFULL + st]);
    }
  } else if (warp == 6) {
    // ================================================= MMA issuer (one lane)
    if (lane == 0) {
      const uint32_t sQ = smem_u32(smem + L::oQ), sK = smem_u32(smem + L::oK);
      const uint32_t sV = smem_u32(smem + L::oV), sP = smem_u32(smem + L::oP);
      auto issue_s = [&](int t) {
        const int st = t & 1;
        mbar_wait(&bar[B_KFULL + st], (t >> 1) & 1);
        fence_proxy_async();
        t5::fence_after_sync();
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          const uint64_t a = t5::smem_desc(sQ + ks * 2 * (kTM * 16), kTM * 16, 128);
          const uint64_t b = t5::smem_desc(sK + st * L::kKV + ks * 2 * (kTN * 16), kTN * 16, 128);
          t5::mma_f16_ss(tmem + st * kTN, a, b, kIdescS, ks > 0);
        }
        t5::commit(&bar[B_SFULL + st]);
      };
      mbar_wait(&bar[B_QFULL], 0);
      issue_s(0);
      for (int j = 0; j < ntiles; ++j) {
        const int st = j & 1;
        if (j + 1 < ntiles) issue_s(j + 1);
        mbar_wait(&bar[B_VFULL + st], (j >> 1) & 1);
        mbar_wait(&bar[B_PFULL], j & 1);
        fence_proxy_async();
        t5::fence_after_sync();
#pragma unroll
        for (int kk = 0; kk < PK; ++kk) {
          const uint64_t a = t5::smem_desc(sP + kk * 2 * (kTM * 16), kTM * 16, 128);
          const uint64_t b = t5::smem_desc(sV + st * L::kKV + kk * 256, 128, kTN * 16);
          t5::mma_f16_ss(tmem_o, a, b, kIdescPV, (j > 0 || kk > 0));
        }
        t5::commit(&bar[B_PVDONE + st]);
      }
    }
    __syncwarp();
  } else {
    // ================================================= softmax warps (thread = query row)
    const int row = i0 + tid;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    uint8_t* sProw = smem + L::oP + tid * 16;
    float m_use = -INFINITY, l_run = 0.f;
    for (int j = 0; j < ntiles; ++j) {
      const int sb = j & 1;
      const int j0 = j * kTN;
      mbar_wait(&bar[B_SFULL + sb], (j >> 1) & 1);
      t5::fence_after_sync();
      uint32_t sr[SC][32];
#pragma unroll
      for (int c = 0; c < SC; ++c) t5::tmem_ld32(sr[c], tmem + lane_base + sb * kTN + c * 32);
      t5::tmem_ld_wait();
      const bool need_mask = (j0 + kTN - 1 > i0) || (j0 + kTN > kv_len);
      if (need_mask) {
        const int lim = min(row, kv_len - 1) - j0;  // columns > lim are masked
#pragma unroll
        for (int c = 0; c < SC; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i > lim) sr[c][i] = 0xff800000u;  // -inf
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int c = 0; c < SC; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(sr[c][i]));
      const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
      // lazy rescale: move the reference max only when it grows by more than the threshold
      float alpha = 1.f;
      const bool grow = m_tile > m_use + kRescaleThreshold || m_use == -INFINITY;
      if (grow) {
        alpha = (m_use == -INFINITY) ? 0.f : fast_exp2(m_use - m_tile);
        m_use = m_tile;
        l_run *= alpha;
      }
      float sum[4] = {0.f, 0.f, 0.f, 0.f};
      const float neg_m = -m_use;
#pragma unroll
      for (int c = 0; c < SC; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float p0 = fast_exp2(fmaf(__uint_as_float(sr[c][i]), p.scale_log2, neg_m));
          const float p1 = fast_exp2(fmaf(__uint_as_float(sr[c][i + 1]), p.scale_log2, neg_m));
          sum[(i >> 1) & 3] += p0 + p1;
          sr[c][i >> 1] = pack_bf16x2(p0, p1);  // packed P overwrites the consumed scores
        }
      l_run += (sum[0] + sum[1]) + (sum[2] + sum[3]);

      if (j > 0) {
        mbar_wait(&bar[B_PVDONE + ((j - 1) & 1)], ((j - 1) >> 1) & 1);  // P buffer free, O settled
        t5::fence_after_sync();
        if (__any_sync(kFull, grow)) {
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            t5::tmem_ld32(o, tmem_o + lane_base + c * 32);
            t5::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            t5::tmem_st32(tmem_o + lane_base + c * 32, o);
          }
          t5::tmem_st_wait();
        }
      }
      // P row -> smem, chunk-major: 8 kv columns (16 bytes) per chunk
#pragma unroll
      for (int c = 0; c < SC; ++c)
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const uint4 val = make_uint4(sr[c][q4 * 4], sr[c][q4 * 4 + 1], sr[c][q4 * 4 + 2], sr[c][q4 * 4 + 3]);
          *reinterpret_cast<uint4*>(sProw + (c * 4 + q4) * (kTM * 16)) = val;
        }
      fence_proxy_async();
      t5::fence_before_sync();
      t5::mbar_arrive(&bar[B_PFULL]);
    }
    // ---- epilogue: O / l -> bf16 -> global
    mbar_wait(&bar[B_PVDONE + ((ntiles - 1) & 1)], ((ntiles - 1) >> 1) & 1);
    t5::fence_after_sync();
    const float inv = 1.f / l_run;
    __nv_bfloat16* dst = p.out + (size_t)(q_start + (row < q_len ? row : 0)) * p.out_stride + (size_t)hq * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t o[32];
      t5::tmem_ld32(o, tmem_o + lane_base + c * 32);
      t5::tmem_ld_wait();
      if (row < q_len) {
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

  // ---- teardown
  t5::fence_before_sync();
  __syncthreads();
  if (warp == 6) {
    t5::fence_after_sync();
    t5::tmem_dealloc<kTmemCols>(tmem);
  }
}

}  // namespace

template <int D, int TN>
static int launch_prefill_tc(const PrefillTcParams& p, const PrefillArgs& a, cudaStream_t stream) {
  const int smem = TcSmem<D, TN>::kBytes;
  auto kern = prefill_attn_tc_kernel<D, TN>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(prefill_tc)", e);
  dim3 grid((a.max_q_len + kTM - 1) / kTM, a.Hq, a.num_reqs);
  kern<<<grid, kTcThreads, smem, stream>>>(p);
  return check_launch("prefill_attn_tc_kernel");
}

static int g_tc_tile_n = 64;
void set_prefill_tc_tile(int tn) { g_tc_tile_n = (tn == 128) ? 128 : 64; }

int prefill_attn_tc(const PrefillArgs& a, cudaStream_t stream) {
  PrefillTcParams p;
  p.q = static_cast<const __nv_bfloat16*>(a.q);
  p.k = static_cast<const __nv_bfloat16*>(a.k);
  p.v = static_cast<const __nv_bfloat16*>(a.v);
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.req_meta = a.req_meta;
  p.Hq = a.Hq; p.Hkv = a.Hkv;
  p.q_stride = a.q_stride; p.out_stride = a.out_stride;
  p.key_stride = a.key_stride; p.value_stride = a.value_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  if (a.D == 128)
    return g_tc_tile_n == 128 ? launch_prefill_tc<128, 128>(p, a, stream)
                              : launch_prefill_tc<128, 64>(p, a, stream);
  if (a.D == 64)
    return g_tc_tile_n == 128 ? launch_prefill_tc<64, 128>(p, a, stream)
                              : launch_prefill_tc<64, 64>(p, a, stream);
  return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size (supported: 64|128)");
}

}  // namespace pc
