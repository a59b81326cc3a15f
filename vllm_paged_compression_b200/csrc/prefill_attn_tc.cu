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
//   warp 4     K producer (one lane): Q tile once, then K tiles, TMA tensor-map loads
//              (cp.async.bulk.tensor.2d, SWIZZLE_128B) -- the streamingLLM gather is one
//              16-row box per kept 16-token run, so the pruned K/V never exists in memory.
//   warp 5     V producer (same; V is the MN-major B operand of P.V).
//              (v1 gathered with cp.async: LDGSTS is served one thread per cycle by the
//              L1TEX t-stage = 16 B/clk/SM, 87 % busy at 22 % tensor-pipe activity --
//              profiles/r01_prefill_tc_v1_ncu_summary.txt.)
//   warp 6     MMA issuer (one lane): S_j = Q K_j^T into one of two TMEM S buffers,
//              O += P_j V_j into the TMEM O accumulator; tcgen05.commit -> mbarriers.
// S_{j+1} is issued before P_j.V_j so the tensor pipe works on the next scores while the
// softmax warps are busy with tile j.  TMEM: 2 x 128 cols (S) + D cols (O) of 512.
//
// O stays in TMEM for the whole row of tiles (hardware accumulate).  When the running max
// of a row grows by more than 2^8 the warp rescales its 32 O rows in TMEM (ld, mul, st);
// smaller increases keep the old reference max -- P and l stay consistent, the result is
// mathematically identical and P stays <= 256 (exact in fp32, same relative bf16 rounding).
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through the runtime)

#include "pc_api_internal.h"
#include "tcgen05.cuh"

namespace pc {

namespace {

constexpr int kTM = 128;       // query rows per CTA (UMMA M)
constexpr int kTcThreads = 384;   // warpgroup 0 softmax, 1 = K / V producers + MMA issuer (+ 1 idle warp), 2 = epilogue
// register budgets per thread after setmaxnreg (two CTAs per SM: 2 x 128 x (128 + 40 + 72) = 60 K of 64 K; the budgets of a CTA must add up to its launch allocation, 384 x 80: setmaxnreg.inc takes from the CTA's own pool and blocks forever otherwise)
constexpr int kRegsSoftmax = 128, kRegsRoles = 40, kRegsEpilogue = 72;
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct PrefillTcParams {
  __nv_bfloat16* __restrict__ out;
  const int* __restrict__ req_meta;  // [R][6] {q_start, q_len, kv_src_start, kv_len, first, pruned}
  int Hq, Hkv;
  long long out_stride;
  float scale_log2;
};
#ifndef PC_P_TMEM
#define PC_P_TMEM 1   // P as a TMEM A operand (0: P through shared memory, the first version)
#endif
// K/V TMA box rows (template parameter BOX) = one run of kept tokens: 16 (first, pruned multiples of
// 16: block sizes 16 / 32) or 8 (block size 8; a box of 8 rows is exactly one 1024-byte SW128 atom)

// TN = kv rows per tile (UMMA N of S, K extent of P.V).  TN = 64 fits two CTAs per SM
// (112 KB of shared memory and 256 TMEM columns each), TN = 128 one.
template <int D, int TN>
struct TcSmem {
  static constexpr int kTN = TN;
  static constexpr int kTmemCols = (2 * TN + D) <= 256 ? 256 : 512;
  static constexpr int kQ = kTM * D * 2;
  static constexpr int kKV = kTN * D * 2;      // one K or V stage
  static constexpr int kP = PC_P_TMEM ? 0 : kTM * kTN * 2;   // P staging (only when P goes through smem)
  static constexpr int oQ = 0;
  static constexpr int oK = oQ + kQ;           // 2 stages
  static constexpr int oV = oK + 2 * kKV;      // 2 stages
  static constexpr int oP = oV + 2 * kKV;
  static constexpr int oBar = oP + kP;
  // barriers: q_full, k_full[2], v_full[2], s_full[2], p_full[2], pv_done[2], l_full[2], o_free (14) + tmem slot
  static constexpr int oLinv = oBar + 16 * 8;   // 1 / l of the item's rows, softmax -> epilogue warps, 2 x kTM floats
  static constexpr int kBytes = oLinv + 2 * kTM * 4;
};

// p_full is double-buffered like the S / P buffers: the softmax warps may finish tile g + 1 while
// the MMA thread still waits for V_g, and a single barrier would then be two phases ahead of its
// waiter (parity aliasing -> deadlock).
enum { B_QFULL = 0, B_KFULL = 1, B_VFULL = 3, B_SFULL = 5, B_PFULL = 7, B_PVDONE = 9, B_LFULL = 11 /* [item parity] */,
       B_OFREE = 13, B_COUNT = 14 };

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------
// A CTA walks work items (query tile, q head, request) k = blockIdx.x, + gridDim.x, ...;
// heaviest query tiles first.  Launched either with one CTA per item (the hardware CTA
// scheduler balances the uneven items: long prompts) or persistently with 2 CTAs per SM.  TMEM allocation, barrier setup and the
// role dispatch happen once; the mbarrier phases simply keep counting over a CTA-wide tile
// counter g.  The next item's Q tile is requested as soon as the last S MMA of the current
// item has read Q, i.e. one tile before the item ends, so its latency hides behind the last
// softmax + P.V + epilogue (the one-shot kernel pays launch, setup, first-load latency and
// teardown per 128-query tile: 33 % tensor-pipe activity on 1024-token prompts, ncu).
struct TcItem {
  int q_start, q_len, kv_src, kv_len, first, pruned, i0, hq, hkv, ntiles;
};

template <int D, int TN, int BOX>
__global__ void __launch_bounds__(kTcThreads, (TN == 64 ? 2 : 1))
prefill_attn_tc_kernel(const PrefillTcParams p, const __grid_constant__ CUtensorMap tm_q,
                               const __grid_constant__ CUtensorMap tm_k,
                               const __grid_constant__ CUtensorMap tm_v, const int q_tiles,
                               const int num_reqs, const int heavy_major, const int chunk_reqs,
                               const int xitem) {
  using L = TcSmem<D, TN>;
  constexpr int kTN = TN;
  constexpr int kTmemCols = L::kTmemCols;
  constexpr int SC = TN / 32;
  constexpr int KS = D / 16;
  constexpr int PK = kTN / 16;
  constexpr uint32_t kIdescS = t5::instr_desc_bf16(kTM, kTN, false);
  constexpr uint32_t kIdescPV = t5::instr_desc_bf16(kTM, D, true);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + B_COUNT);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int per_qt = p.Hq * num_reqs;
  const int total = q_tiles * per_qt;
  const int G = p.Hq / p.Hkv;

  // item k -> (query tile descending, request, q head); ntiles == 0: nothing to do
  auto get_item = [&](int k) {
    TcItem it;
    // heavy_major: all (request, head) pairs of the heaviest query tile first (persistent walk,
    // short prompts: static round-robin needs the global order).  Otherwise the query tiles of one
    // (request, head) are adjacent, heaviest first, so the CTAs in flight share few K/V streams
    // (4 x 8192 tokens: 134 MB of K/V > L2 when every stream is live at once: 1061 vs 1104 TFLOP/s).
    // The heavy-first order runs inside CHUNKS of `chunk_reqs` requests whose K/V fit L2 together:
    // every query-tile level of a chunk re-reads the chunk's K/V, and with all 64 prompts of the
    // bench batch in one level-major sweep (142 MB of K/V) those re-reads went to DRAM (ncu: 1.44 GB
    // read for 0.68 GB algorithmic).
    int qt, rem, r0 = 0;
    if (heavy_major) {
      const int per_chunk = q_tiles * p.Hq * chunk_reqs;
      const int c = k / per_chunk, k2 = k - c * per_chunk;
      r0 = c * chunk_reqs;
      const int pq = p.Hq * min(chunk_reqs, num_reqs - r0);
      qt = q_tiles - 1 - k2 / pq;
      rem = k2 % pq;
    } else {
      rem = k / q_tiles;
      qt = q_tiles - 1 - (k - rem * q_tiles);
    }
    const int r = r0 + rem / p.Hq;
    it.hq = rem % p.Hq;
    it.hkv = it.hq / G;
    const int* m = p.req_meta + r * 6;
    it.q_start = m[0]; it.q_len = m[1]; it.kv_src = m[2]; it.kv_len = m[3]; it.first = m[4]; it.pruned = m[5];
    it.i0 = qt * kTM;
    it.ntiles = it.i0 < it.q_len ? (min(it.kv_len, it.i0 + kTM) + kTN - 1) / kTN : 0;
    return it;
  };
  auto issue_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* full, const TcItem& it, int t) {
    mbar_expect_tx(full, L::kKV);
#pragma unroll
    for (int g = 0; g < kTN / BOX; ++g) {
      const int j = t * kTN + g * BOX;
      const int src = it.kv_src + (j < it.first ? j : j + it.pruned);
#pragma unroll
      for (int b = 0; b < D / 64; ++b)
        t5::tma_load_2d(dst + b * (kTN * 128) + g * (BOX * 128), tm, it.hkv * D + b * 64, src, full);
    }
  };

  if (tid == 0) {
    if ((smem_u32(smem) & 1023u) != 0u) __trap();
    mbar_init(&bar[B_QFULL], 1);
    mbar_init(&bar[B_KFULL], 1); mbar_init(&bar[B_KFULL + 1], 1);
    mbar_init(&bar[B_VFULL], 1); mbar_init(&bar[B_VFULL + 1], 1);
    mbar_init(&bar[B_SFULL], 1);  mbar_init(&bar[B_SFULL + 1], 1);
    mbar_init(&bar[B_PFULL], kTM); mbar_init(&bar[B_PFULL + 1], kTM);
    mbar_init(&bar[B_PVDONE], 1); mbar_init(&bar[B_PVDONE + 1], 1);
    mbar_init(&bar[B_LFULL], kTM); mbar_init(&bar[B_LFULL + 1], kTM); mbar_init(&bar[B_OFREE], kTM);
    mbar_fence_init();
  }
  if (warp == 6) t5::tmem_alloc<kTmemCols>(tmem_slot);
  t5::fence_before_sync();
  __syncthreads();
  t5::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_o = tmem + 2 * kTN;
  float* linv = reinterpret_cast<float*>(smem + L::oLinv);
  // registers follow the roles (two CTAs per SM share the file): the softmax rows hold a whole
  // 64-column tile per thread, the single-lane roles and the epilogue need few.  Each role branch
  // starts with its setmaxnreg so that the allocator sees the budget of the code it dominates.
  constexpr bool kRealloc = (TN == 64);

  // the one thread of a single-lane role: elect.sync lets the compiler keep tcgen05.mma / TMA operands in uniform
  // registers (tcgen05.cuh: elect_one); head size 64 measured 2 % slower with it (profiles/r02_prefill_elect_ab.jsonl)
  // and keeps lane 0
  auto role_lane = [&]() {
    if constexpr (D == 64) return lane == 0; else return t5::elect_one();
  };
  if (warp >= 4 && warp < 8) {
  if constexpr (kRealloc) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsRoles));
  if (warp == 4) {
    // ================================================= Q + K producer (one lane, TMA)
    if (role_lane()) {
      int g = 0;   // tiles issued so far by this CTA (all items)
      for (int k = blockIdx.x; k < total; k += gridDim.x) {
        const TcItem it = get_item(k);
        if (it.ntiles == 0) continue;
        if (it.pruned != 0 && (((it.first | it.pruned) & (BOX - 1)) != 0)) __trap();
        // K of the item's first tile goes out first (its stage only needs S_{g-2}); the Q buffer is free once
        // the previous item's last S MMA (tile g - 1) has completed
        if (g >= 2) t5::mbar_wait_relaxed(&bar[B_SFULL + (g & 1)], ((g - 2) >> 1) & 1);
        issue_kv(&tm_k, smem + L::oK + (g & 1) * L::kKV, &bar[B_KFULL + (g & 1)], it, 0);
        if (g > 0) t5::mbar_wait_relaxed(&bar[B_SFULL + ((g - 1) & 1)], ((g - 1) >> 1) & 1);
        mbar_expect_tx(&bar[B_QFULL], L::kQ);
#pragma unroll
        for (int b = 0; b < D / 64; ++b)
          t5::tma_load_2d(smem + L::oQ + b * (kTM * 128), &tm_q, it.hq * D + b * 64, it.q_start + it.i0,
                          &bar[B_QFULL]);
        // the NEXT item's Q tile (read exactly once: always a DRAM miss) and its first K / V tile go
        // to L2 now, a whole item ahead of their TMA loads
        for (int kn = k + gridDim.x; kn < total; kn += gridDim.x) {
          const TcItem nx = get_item(kn);
          if (nx.ntiles == 0) continue;
#pragma unroll
          for (int b = 0; b < D / 64; ++b) t5::tma_prefetch_l2_2d(&tm_q, nx.hq * D + b * 64, nx.q_start + nx.i0);
#pragma unroll
          for (int gg = 0; gg < kTN / BOX; ++gg) {
            const int jn = gg * BOX;
            const int src = nx.kv_src + (jn < nx.first ? jn : jn + nx.pruned);
#pragma unroll
            for (int b = 0; b < D / 64; ++b) {
              t5::tma_prefetch_l2_2d(&tm_k, nx.hkv * D + b * 64, src);
              t5::tma_prefetch_l2_2d(&tm_v, nx.hkv * D + b * 64, src);
            }
          }
          break;
        }
        ++g;
        for (int t = 1; t < it.ntiles; ++t, ++g) {
          const int st = g & 1;
          if (g >= 2) t5::mbar_wait_relaxed(&bar[B_SFULL + st], ((g - 2) >> 1) & 1);   // S_{g-2} done
          issue_kv(&tm_k, smem + L::oK + st * L::kKV, &bar[B_KFULL + st], it, t);
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ================================================= V producer (one lane, TMA)
    if (role_lane()) {
      int g = 0;
      for (int k = blockIdx.x; k < total; k += gridDim.x) {
        const TcItem it = get_item(k);
        for (int t = 0; t < it.ntiles; ++t, ++g) {
          const int st = g & 1;
          if (g >= 2) t5::mbar_wait_relaxed(&bar[B_PVDONE + st], ((g - 2) >> 1) & 1);  // P.V_{g-2} done
          issue_kv(&tm_v, smem + L::oV + st * L::kKV, &bar[B_VFULL + st], it, t);
        }
      }
    }
    __syncwarp();
  } else if (warp == 6) {
    // ================================================= MMA issuer (one lane)
    if (role_lane()) {
      const uint32_t sQ = smem_u32(smem + L::oQ), sK = smem_u32(smem + L::oK);
      const uint32_t sV = smem_u32(smem + L::oV), sP = smem_u32(smem + L::oP);
      auto issue_s = [&](int g) {
        const int st = g & 1;
        mbar_wait(&bar[B_KFULL + st], (g >> 1) & 1);
        t5::fence_after_sync();
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          const uint64_t a = t5::smem_desc_sw128(sQ + (ks >> 2) * (kTM * 128) + (ks & 3) * 32, 16, 1024);
          const uint64_t b = t5::smem_desc_sw128(sK + st * L::kKV + (ks >> 2) * (kTN * 128) + (ks & 3) * 32, 16, 1024);
          t5::mma_f16_ss(tmem + st * kTN, a, b, kIdescS, ks > 0);
        }
        t5::commit(&bar[B_SFULL + st]);
      };
      // The tile stream runs ACROSS items: the first S MMA of the next item (new Q, its K tile 0) is issued
      // before the last P.V of the current one, exactly like S_{j+1} inside an item, so the softmax warps
      // find the next item's scores ready when they hand over the last P (the item boundary used to cost
      // them one S MMA + commit latency: the MMA thread waited for the last P before it looked at Q).
      auto next_item = [&](int& kk) {     // tiles of the next non-empty item at or after kk (0: none left)
        for (; kk < total; kk += gridDim.x) {
          const int nt = get_item(kk).ntiles;
          if (nt != 0) return nt;
        }
        return 0;
      };
      int g = 0, n_item = 0, n_q = 0, k = blockIdx.x;
      int ntiles = next_item(k);
      bool have = ntiles != 0;
      if (have) {
        mbar_wait(&bar[B_QFULL], 0);
        n_q = 1;
        issue_s(0);
      }
      while (have) {
        int kn = k + gridDim.x, nt_nx = 0;
        bool have_nx = false;
        for (int j = 0; j < ntiles; ++j, ++g) {
          const int st = g & 1;
          if (j + 1 < ntiles) {
            issue_s(g + 1);
          } else {
            nt_nx = next_item(kn);
            have_nx = nt_nx != 0;
            if (have_nx && xitem) {
              mbar_wait(&bar[B_QFULL], n_q & 1);
              ++n_q;
              issue_s(g + 1);
            }
          }
          mbar_wait(&bar[B_VFULL + st], (g >> 1) & 1);
          mbar_wait(&bar[B_PFULL + st], (g >> 1) & 1);
          // the first P.V of an item overwrites O: the epilogue warps must have read the previous item's
          if (j == 0 && n_item > 0) mbar_wait(&bar[B_OFREE], (n_item - 1) & 1);
          fence_proxy_async();
          t5::fence_after_sync();
#pragma unroll
          for (int kk = 0; kk < PK; ++kk) {
            const uint64_t b = t5::smem_desc_sw128(sV + st * L::kKV + kk * (16 * 128), kTN * 128, 1024);
#if PC_P_TMEM
            // P_g lives in the first TN / 2 columns of the S buffer it was computed from
            t5::mma_f16_ts(tmem_o, tmem + st * kTN + kk * 8, b, kIdescPV, (j > 0 || kk > 0));
#else
            const uint64_t a = t5::smem_desc(sP + kk * 2 * (kTM * 16), kTM * 16, 128);
            t5::mma_f16_ss(tmem_o, a, b, kIdescPV, (j > 0 || kk > 0));
#endif
          }
          t5::commit(&bar[B_PVDONE + st]);
        }
        ++n_item;
        if (have_nx && !xitem) {
          mbar_wait(&bar[B_QFULL], n_q & 1);
          ++n_q;
          issue_s(g);
        }
        ntiles = nt_nx; k = kn; have = have_nx;
      }
    }
    __syncwarp();
  }
  } else if (warp < 4) {
    // ================================================= softmax warps (thread = query row)
    if constexpr (kRealloc) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    uint8_t* sProw = smem + L::oP + tid * 16;
    int g = 0, n_item = 0;
    for (int k = blockIdx.x; k < total; k += gridDim.x) {
      const TcItem it = get_item(k);
      if (it.ntiles == 0) continue;
      const int row = it.i0 + tid;
      float m_use = -INFINITY, l_run = 0.f;
      for (int j = 0; j < it.ntiles; ++j, ++g) {
        const int sb = g & 1;
        const int j0 = j * kTN;
        mbar_wait(&bar[B_SFULL + sb], (g >> 1) & 1);
        t5::fence_after_sync();
        uint32_t sr[SC][32];
#pragma unroll
        for (int c = 0; c < SC; ++c) t5::tmem_ld32(sr[c], tmem + lane_base + sb * kTN + c * 32);
        t5::tmem_ld_wait();
        const bool need_mask = (j0 + kTN - 1 > it.i0) || (j0 + kTN > it.kv_len);
        if (need_mask) {
          const int lim = min(row, it.kv_len - 1) - j0;  // columns > lim are masked
#pragma unroll
          for (int c = 0; c < SC; ++c) {
            // a 32-column chunk no row of the warp masks is skipped (warp-uniform test): on a
            // diagonal tile only one warp in four has a partial chunk
            if (__all_sync(kFull, lim >= c * 32 + 31)) continue;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i > lim) sr[c][i] = 0xff800000u;  // -inf
          }
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < SC; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(sr[c][i]));
        const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
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
            sr[c][i >> 1] = pack_bf16x2(p0, p1);
          }
        l_run += (sum[0] + sum[1]) + (sum[2] + sum[3]);

#if PC_P_TMEM
        // P_g has its own buffer (the S buffer it came from): only a rescale of O has to wait
        // for the previous P.V
        if (j > 0 && __any_sync(kFull, grow)) {
          mbar_wait(&bar[B_PVDONE + ((g - 1) & 1)], ((g - 1) >> 1) & 1);  // O settled
          t5::fence_after_sync();
          {
#else
        if (j > 0) {   // (j == 0: the previous item's epilogue already waited for its last P.V)
          mbar_wait(&bar[B_PVDONE + ((g - 1) & 1)], ((g - 1) >> 1) & 1);  // P buffer free, O settled
          t5::fence_after_sync();
          if (__any_sync(kFull, grow)) {
#endif
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
#if PC_P_TMEM
        // packed P row -> TMEM, over the scores it came from (A operand of P.V): 32 columns
        // (= 64 bf16) per store
#pragma unroll
        for (int h = 0; h < SC / 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int i = 0; i < 16; ++i) { pk[i] = sr[2 * h][i]; pk[16 + i] = sr[2 * h + 1][i]; }
          t5::tmem_st32(tmem + lane_base + sb * kTN + h * 32, pk);
        }
        t5::tmem_st_wait();
#else
#pragma unroll
        for (int c = 0; c < SC; ++c)
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const uint4 val = make_uint4(sr[c][q4 * 4], sr[c][q4 * 4 + 1], sr[c][q4 * 4 + 2], sr[c][q4 * 4 + 3]);
            *reinterpret_cast<uint4*>(sProw + (c * 4 + q4) * (kTM * 16)) = val;
          }
        fence_proxy_async();
#endif
        t5::fence_before_sync();
        t5::mbar_arrive(&bar[B_PFULL + sb]);
      }
      // ---- hand the row's 1 / l to the epilogue warps and move on to the next item: the softmax
      // warps are the critical resource (MUFU), they neither wait for the last P.V nor drain O
      // (l_full is double-buffered by item parity: after a one-tile item the softmax warps can finish
      // item n + 1 before the epilogue has looked at item n; a single barrier would then be two
      // phases ahead of its waiter)
      linv[(n_item & 1) * kTM + tid] = 1.f / l_run;
      t5::mbar_arrive(&bar[B_LFULL + (n_item & 1)]);
      ++n_item;
    }
  }
  else {
    // ================================================= epilogue warps: O / l -> bf16 -> global
    if constexpr (kRealloc) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsEpilogue));
    // (TMEM lane quadrant = warp % 4; O is read as soon as the item's last P.V has completed and
    // released to the MMA warp -- o_free -- before the conversion and the stores)
    const int etid = tid - 256;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    int g = 0, n_item = 0;
    for (int k = blockIdx.x; k < total; k += gridDim.x) {
      const TcItem it = get_item(k);
      if (it.ntiles == 0) continue;
      g += it.ntiles;
      const int row = it.i0 + etid;
      // (polling with a sleep: these warps idle most of an item and must not take issue slots
      // from the softmax warps that share their schedulers)
      t5::mbar_wait_sleep(&bar[B_LFULL + (n_item & 1)], (n_item >> 1) & 1, 128);
      const float inv = linv[(n_item & 1) * kTM + etid];
      t5::mbar_wait_sleep(&bar[B_PVDONE + ((g - 1) & 1)], ((g - 1) >> 1) & 1, 64);
      t5::fence_after_sync();
      __nv_bfloat16* dst = p.out + (size_t)(it.q_start + (row < it.q_len ? row : 0)) * p.out_stride + (size_t)it.hq * D;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t o[32];
        t5::tmem_ld32(o, tmem_o + lane_base + c * 32);
        t5::tmem_ld_wait();
        if (c == D / 32 - 1) {          // O is in registers: the next item's first P.V may overwrite it
          t5::fence_before_sync();
          t5::mbar_arrive(&bar[B_OFREE]);
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
      ++n_item;
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

// One CTA per item or persistent CTAs.  B200, TFLOP/s, one CTA per item / persistent (P in TMEM):
//   16 x 1024 q (544 kept) 642 / 728    64 x 1024 690 / 718    16 x 1024, D 64 380 / 444
//   16 x 2048 (1056 kept)  879 / 884     8 x 4096 1026 / 1019   4 x 8192 (no eviction) 1133 / 1082
// The persistent walk saves the per-item launch / setup / first-load latency that dominates short
// prompts; on long ones the hardware CTA scheduler balances the uneven items better than the
// static round-robin, and the per-(request, head) item order keeps the live K/V streams in L2.
// Rule: persistent for max_q_len <= 2048.  (PC_DEBUG builds can override: -1 = rule, 0 = one CTA
// per item, 1 = persistent, 2 = persistent with the one-shot item order.)
#ifdef PC_DEBUG
static int g_tc_persistent = -1;
void set_prefill_tc_persistent(int v) { g_tc_persistent = v; }
static int g_tc_tile_n = 64;
static int g_tc_chunk_reqs = 0;
void set_prefill_tc_tile(int tn) {   // debug knob reused: values other than 64 / 128 set the requests per L2 chunk
  if (tn == 64 || tn == 128) g_tc_tile_n = tn; else g_tc_chunk_reqs = tn;
}
static int g_prefill_pp = -1;
void set_prefill_pp(int v) { g_prefill_pp = v; }
static int g_prefill_dq = -1;
void set_prefill_dq(int v) { g_prefill_dq = v; }   // 1 = always (when supported), 0 = never, -1 = rule
static int g_tc_xitem = 1;
void set_prefill_tc_xitem(int v) { g_tc_xitem = v; }
#else
constexpr int g_tc_xitem = 1;
constexpr int g_prefill_dq = -1;
constexpr int g_tc_persistent = -1;
constexpr int g_tc_tile_n = 64;
#endif

// ---- host: tensor maps.  The encoder is a driver entry point fetched through the runtime
// (no link against libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}
// [rows, heads*D] bf16 matrix with row stride `stride` elements; box = 64 columns x box_rows,
// SWIZZLE_128B (the UMMA SW128 canonical layouts), out-of-range rows read as zero.
int prefill_make_rows_map(CUtensorMap* m, const void* base, long long rows, int inner, long long stride,
                          int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
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

template <int D, int TN, int BOX>
static int launch_prefill_tc(const PrefillTcParams& p, const PrefillArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  if (int rc = prefill_make_rows_map(&tq, a.q, a.num_tokens, a.Hq * D, a.q_stride, kTM)) return rc;
  if (int rc = prefill_make_rows_map(&tk, a.k, a.num_tokens, a.Hkv * D, a.key_stride, BOX)) return rc;
  if (int rc = prefill_make_rows_map(&tv, a.v, a.num_tokens, a.Hkv * D, a.value_stride, BOX)) return rc;
  const int smem = TcSmem<D, TN>::kBytes;
  auto kern = prefill_attn_tc_kernel<D, TN, BOX>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(prefill_tc)", e);
  const bool persist = g_tc_persistent < 0 ? (a.max_q_len <= 2048) : (g_tc_persistent != 0);
  const int q_tiles = (a.max_q_len + kTM - 1) / kTM;
  const long long total = (long long)q_tiles * a.Hq * a.num_reqs;
  const long long slots = (long long)device_sm_count() * (TN == 64 ? 2 : 1);
  if (total > 0x7fffffffLL) return set_error(PC_ERR_INVALID, "prefill: too many query tiles");
  const int ctas = (int)((persist && total > slots) ? slots : total);
  // requests per chunk of the heavy-first order: K + V of a chunk (bounded by the longest prompt) in ~40 MB of L2
  const long long kv_req = (long long)a.max_q_len * a.Hkv * D * 4;
  long long chunk = (40LL << 20) / (kv_req > 0 ? kv_req : 1);
  if (chunk < 1) chunk = 1;
  // (a batch whose K/V fit L2 anyway keeps the single heavy-first sweep: 16 x 1024 measured 753
  // TFLOP/s in one sweep against 717 in chunks of 6, the chunk ends add light / heavy transitions)
  if (chunk > a.num_reqs || kv_req * a.num_reqs <= (80LL << 20)) chunk = a.num_reqs;
#ifdef PC_DEBUG
  if (g_tc_chunk_reqs > 0) chunk = g_tc_chunk_reqs < a.num_reqs ? g_tc_chunk_reqs : a.num_reqs;
#endif
  kern<<<ctas, kTcThreads, smem, stream>>>(p, tq, tk, tv, q_tiles, a.num_reqs,
                                            (persist && g_tc_persistent != 2) ? 1 : 0,   // 2: debug, persistent + local order
                                            (int)chunk, g_tc_xitem);
  return check_launch("prefill_attn_tc_kernel");
}

// A1 entry: argument checks + launch.  The streamingLLM gather runs as TMA box coordinates, so
// `first` and `pruned` of every pruned request must be multiples of 8 (whole blocks of 8 / 16 / 32;
// multiples of 16 take 16-row boxes, otherwise 8-row boxes: twice the TMA instructions).  The check needs the HOST copy
// of req_meta; without one (req_meta_host == NULL) a violating request reaches the device-side
// guard, which is a trap -- callers that cannot vouch for the alignment pass the host copy.
int prefill_attn(const PrefillArgs& a, cudaStream_t stream) {
  if (a.num_reqs <= 0 || a.max_q_len <= 0) return 0;
  if (a.Hkv <= 0 || a.Hq % a.Hkv) return set_error(PC_ERR_INVALID, "num_heads must be a multiple of num_kv_heads");
  if ((a.q_stride | a.out_stride | a.key_stride | a.value_stride) % 8)
    return set_error(PC_ERR_INVALID, "q/out/key/value row strides must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) |
       reinterpret_cast<uintptr_t>(a.v) | reinterpret_cast<uintptr_t>(a.out)) & 15)
    return set_error(PC_ERR_INVALID, "q/k/v/out pointers must be 16-byte aligned");
  if (a.num_tokens <= 0) return set_error(PC_ERR_INVALID, "prefill: num_tokens must be the row count of q/k/v");
  bool box8 = false;
  if (a.req_meta_host != nullptr) {
    for (int r = 0; r < a.num_reqs; ++r) {
      const int* m = a.req_meta_host + r * 6;
      if (m[1] < 0 || m[3] < 0 || m[3] > m[1] || m[0] < 0 || m[2] < 0 || m[4] < 0 || m[5] < 0 ||
          (long long)m[0] + m[1] > a.num_tokens || (long long)m[2] + m[3] + m[5] > a.num_tokens)
        return set_error(PC_ERR_INVALID, "prefill: req_meta row out of range (q_len >= kv_len >= 0, rows inside num_tokens)");
      if (m[5] != 0 && (((m[4] | m[5]) & 7) != 0))
        return set_error(PC_ERR_INVALID,
                         "prefill: first / pruned must be multiples of 8 tokens (whole blocks of 8 / 16 / 32)");
      if (m[5] != 0 && (((m[4] | m[5]) & 15) != 0)) box8 = true;     // block size 8: 8-row TMA boxes
    }
  }
  // Dual-query-tile kernel (prefill_attn_dq.cu: two q heads of a kv head per CTA over one shared K/V ring).
  // Rule from the same-box A/Bs in profiles/r02_prefill_dq_ab.jsonl and r02_prefill_dq_threshold_ab.jsonl: head
  // size 128 and prompts of at most 6144 tokens (+26 % at 256 tokens, +19 % at 512, +11 % at 1024, +5-6 % at
  // 2048 / 4096 / 6144; -2 % at 8192 without eviction and at head size 64, which stay with the kernel of this file).
  if (prefill_attn_dq_supported(a) &&
      (g_prefill_dq > 0 || (g_prefill_dq < 0 && a.D == 128 && a.max_q_len <= kDqMaxQLen)))
    return prefill_attn_dq(a, box8, stream);
#ifdef PC_DEBUG
  // A/B runs: the ping-pong kernel (two q heads of a kv head per CTA, 128-wide kv tiles; prefill_attn_pp.cu).
  // Measured 1-8 % slower than this kernel at every shape (profiles/r02_prefill_experiments.txt), so it is
  // built into the PC_DEBUG library only.
  if (g_prefill_pp > 0 && prefill_attn_pp_supported(a)) return prefill_attn_pp(a, stream);
#endif
  PrefillTcParams p;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.req_meta = a.req_meta;
  p.Hq = a.Hq; p.Hkv = a.Hkv;
  p.out_stride = a.out_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
#ifdef PC_DEBUG
  if (g_tc_tile_n == 128) {
    if (a.D == 128) return launch_prefill_tc<128, 128, 16>(p, a, stream);
    if (a.D == 64) return launch_prefill_tc<64, 128, 16>(p, a, stream);
  }
#endif
  if (box8) {
    if (a.D == 128) return launch_prefill_tc<128, 64, 8>(p, a, stream);
    if (a.D == 64) return launch_prefill_tc<64, 64, 8>(p, a, stream);
  }
  if (a.D == 128) return launch_prefill_tc<128, 64, 16>(p, a, stream);
  if (a.D == 64) return launch_prefill_tc<64, 64, 16>(p, a, stream);
  return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size (supported: 64|128)");
}

}  // namespace pc
