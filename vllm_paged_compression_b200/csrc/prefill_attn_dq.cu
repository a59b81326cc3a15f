// Prefill attention (A1), "dual query tile" kernel: ONE CTA per SM runs TWO streams of
// prefill_attn_tc.cu (two q heads of one kv head, the same 128 query rows) over ONE shared K/V ring.
//
// Why (profiles/r02_prefill_experiments.txt, clock trace of one CTA of prefill_attn_tc_kernel): a K or V
// tile lands ~2600-3100 clk after its TMA request under load.  The two-CTAs-per-SM kernel has 2 K + 2 V
// stages per CTA (96 KB each), so V_{g} is requested when P.V_{g-2} completes and two tiles cannot take
// less than that latency: the MMA thread waits for V (and K), the softmax warps for S.  The two CTAs of
// an SM mostly work on neighbouring q heads of the SAME kv head and each loads the same K/V tiles.
// Here the two streams share the tiles: one ring of K and V stages feeds both, at half the L2 -> shared-memory
// traffic per flop.  Shared memory (head size 128, 226 KB): TWO Q pairs (4 x 32 KB, by item parity) + 3 K + 3 V
// stages (16 KB each: two tiles of look-ahead).  A first version had one Q pair and 5 + 5 stages; a fit of its
// launch times over 256-4096-token prompts gave ~2000 clk per kv tile plus ~6000 clk PER ITEM (31 % of the time at
// 1024-token prompts): both streams reach an item boundary together, and the next Q pair could only be requested
// when the item's last S MMAs had completed (TMA ~3000 clk + S ~1300 clk before the softmax warps see scores
// again, with nothing else on the SM to fill the gap).  With two Q sets the next item's pair lands during the
// current item and the first S of the next item is issued like any other tile's.
//   * TMEM (512 columns): stream i at i * 256: S buffers [0, 64) [64, 128) (P_g bf16 over the first half
//     of the S buffer it came from), O at [128, 256);
//   * warpgroups 0 / 1 = softmax of stream 0 / 1 (thread = query row, 64 columns per tile),
//     warpgroup 2 = K/V producer warp, MMA issuer warp of stream 0, Q producer warp, MMA issuer warp of
//     stream 1, warpgroup 3 = epilogue; registers follow the roles (setmaxnreg 176 / 176 / 88 / 72);
//   * EACH STREAM HAS ITS OWN MMA ISSUER THREAD: a first version with one issuer for both streams
//     (S0 S1 | PV0 PV1 in program order) was 1-25 % slower than the two-CTA kernel -- the issuer sat in
//     stream 0's wait for P while stream 1's MMAs were ready.  The ring barriers count both streams
//     (k_free / v_free: 2 commits);
//   * the K/V producer runs ahead across work items (the Q pair has its own producer), each issuer
//     treats the tiles of consecutive items as one stream: S of tile g + 1 (the next item's first tile
//     included) is issued before P.V of tile g.  K is requested one tile ahead of V (K0 K1 V0 K2 V1 ...);
//   * the single-lane roles run under elect.sync (t5::elect_one), which keeps the tcgen05.mma / TMA operands in
//     uniform registers: under `lane == 0` every such instruction sat in an ELECT / R2UR / BRA.U.ANY loop and the
//     issuer thread's loop WAS the tile period (clock trace, profiles/r02_prefill_dq_trace_1024*.txt);
//   * work items are walked with an ItemIter (no divisions) and every role asks for the next item's descriptor at
//     the top of the current item.
// Semantics, numerics, masks, the TMA box gather of the streamingLLM prune, the lazy rescale and the
// item order are those of prefill_attn_tc.cu (same tests).  Needs an even number of q heads per kv
// head; everything else goes to prefill_attn_tc.cu.
#include <cuda.h>

#include "pc_api_internal.h"
#include "tcgen05.cuh"

namespace pc {

namespace {

constexpr int kTM = 128;        // query rows per tile (UMMA M)
constexpr int kTN = 64;         // kv rows per tile
constexpr int kStages = 3;      // K stages = V stages
constexpr int kQSets = 2;       // Q pairs in shared memory (by item parity): the next item's pair lands during this item
constexpr int kDqThreads = 512;
constexpr int kRegsSoftmaxDq = 176, kRegsRolesDq = 88, kRegsEpiDq = 72;   // 128 x (2 x 176 + 88 + 72) = 64 K
constexpr float kRescaleThresholdDq = 8.f;   // log2 domain

struct DqParams {
  __nv_bfloat16* __restrict__ out;
  const int* __restrict__ req_meta;  // [R][6] {q_start, q_len, kv_src_start, kv_len, first, pruned}
  int Hq, Hkv;
  long long out_stride;
  float scale_log2;
};

template <int D>
struct DqSmem {
  static constexpr int kQ = kTM * D * 2;         // one Q tile
  static constexpr int kKV = kTN * D * 2;        // one K or V stage
  static constexpr int oQ = 0;                   // kQSets x 2 tiles
  static constexpr int oK = oQ + kQSets * 2 * kQ;  // kStages
  static constexpr int oV = oK + kStages * kKV;  // kStages
  static constexpr int oBar = oV + kStages * kKV;
  static constexpr int oLinv = oBar + 48 * 8;    // [item parity][stream][kTM] floats
  static constexpr int kBytes = oLinv + 2 * 2 * kTM * 4;
  static constexpr int kTmemCols = 512;
};

// every barrier a waiter can fall behind on is indexed so that the waiter is at most one phase late
// (S / P / P.V by tile parity, l by item parity, the ring by stage)
enum {
  Q_QFULL = 0,                       // [item parity]
  Q_KFULL = 2,                       // [stage]
  Q_VFULL = Q_KFULL + kStages,
  Q_KFREE = Q_VFULL + kStages,       // S of both streams done with the stage (2 commits)
  Q_VFREE = Q_KFREE + kStages,       // P.V of both streams done with the stage (2 commits)
  Q_SFULL = Q_VFREE + kStages,       // [stream][tile parity]
  Q_PFULL = Q_SFULL + 4,
  Q_PVDONE = Q_PFULL + 4,
  Q_LFULL = Q_PVDONE + 4,            // [stream][item parity]
  Q_OFREE = Q_LFULL + 4,             // [stream]
  Q_QFREE = Q_OFREE + 2,             // [item parity]: the last S MMAs of both streams have read the Q pair (2 commits)
  Q_COUNT = Q_QFREE + 2
};
static_assert(Q_COUNT + 1 <= 48, "barrier area");

struct DqItem {
  int q_start, q_len, kv_src, kv_len, first, pruned, i0, hq0, hkv, ntiles;
};
struct ItemIter {   // position of a CTA in the item order
  int k;            // item index (>= total: past the end)
  int qt;           // query tile (descending inside a chunk of requests)
  int r0, nreq;     // the chunk: first request, requests in it
  int rr, pr;       // request inside the chunk, pair of q heads
};

#ifdef PC_DEBUG
// per-role clock64 stamps of CTA 0 (profiles/tools_trace_prefill_dq.py): [role][kDqTraceCap] words of
// code << 56 | (tile & 0xffff) << 40 | clock (40 bits)
constexpr int kDqTraceRoles = 8, kDqTraceCap = 8192;
__device__ unsigned long long g_dq_trace[kDqTraceRoles * kDqTraceCap];
__device__ int g_dq_trace_on = 0;
struct DqTracer {
  unsigned long long* base;
  int n;
  __device__ __forceinline__ void init(int role, bool on) {
    base = on ? g_dq_trace + role * kDqTraceCap : nullptr;
    n = 0;
  }
  __device__ __forceinline__ void operator()(int code, int g) {
    if (base != nullptr && n < kDqTraceCap)
      base[n++] = ((unsigned long long)code << 56) | ((unsigned long long)(g & 0xffff) << 40) |
                  ((unsigned long long)clock64() & 0xffffffffffull);
  }
};
#define DQ_TRACER(name, role, on) DqTracer name; name.init(role, (on) && blockIdx.x == 0 && g_dq_trace_on != 0)
#define DQ_TR(name, code, g) name(code, g)
#else
#define DQ_TRACER(name, role, on)
#define DQ_TR(name, code, g)
#endif

// a load the compiler leaves where it is written (volatile asm keeps its order among the mbarrier / tcgen05 asm)
__device__ __forceinline__ int ld_early(const int* p) {
  int v;
  asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <int D, int BOX>
__global__ void __launch_bounds__(kDqThreads, 1)
prefill_attn_dq_kernel(const DqParams p, const __grid_constant__ CUtensorMap tm_q,
                       const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                       const int q_tiles, const int num_reqs, const int chunk_reqs) {
  using L = DqSmem<D>;
  constexpr int SC = kTN / 32;     // 32-column chunks of a score row
  constexpr int KS = D / 16;       // k-steps of S = Q K^T
  constexpr int PK = kTN / 16;     // k-steps of O += P V
  constexpr uint32_t kIdescS = t5::instr_desc_bf16(kTM, kTN, false);
  constexpr uint32_t kIdescPV = t5::instr_desc_bf16(kTM, D, true);
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + Q_COUNT);
  float* linv = reinterpret_cast<float*>(smem + L::oLinv);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = p.Hq / p.Hkv;
  const int pairs = p.Hq / 2;                      // items per (request, query tile)
  const int total = q_tiles * pairs * num_reqs;

  // item k -> (chunk of requests, query tile descending, request, head pair): the order of
  // prefill_attn_tc.cu with pairs of q heads in the place of q heads.  A CTA visits k = blockIdx.x,
  // + gridDim.x, ...: the position is kept as an ItemIter and stepped WITHOUT divisions -- the six
  // integer divisions of the closed form went through the XU pipe (I2F / RCP / F2I) behind the softmax
  // warps' exponentials and cost the softmax warps 1200-2100 clk between items (clock trace,
  // profiles/r02_prefill_dq_trace_1024_elect.txt).
  const int step_rr = (int)gridDim.x / pairs, step_pr = (int)gridDim.x - step_rr * pairs;
  auto it_begin = [&]() {
    ItemIter s;
    s.k = blockIdx.x; s.qt = 0; s.r0 = 0; s.nreq = 1; s.rr = 0; s.pr = 0;
    if (s.k < total) {
      const int per_chunk = q_tiles * pairs * chunk_reqs;
      const int c = s.k / per_chunk, k2 = s.k - c * per_chunk;
      s.r0 = c * chunk_reqs;
      s.nreq = min(chunk_reqs, num_reqs - s.r0);
      const int pq = pairs * s.nreq, lv = k2 / pq, rem = k2 - lv * pq;
      s.qt = q_tiles - 1 - lv;
      s.rr = rem / pairs;
      s.pr = rem - s.rr * pairs;
    }
    return s;
  };
  auto it_next = [&](ItemIter& s) {    // k += gridDim.x
    s.k += gridDim.x;
    if (s.k >= total) return;
    s.pr += step_pr; s.rr += step_rr;
    if (s.pr >= pairs) { s.pr -= pairs; ++s.rr; }
    while (s.rr >= s.nreq) {           // past the chunk's requests at this query-tile level
      s.rr -= s.nreq;
      if (--s.qt < 0) { s.qt = q_tiles - 1; s.r0 += s.nreq; s.nreq = min(chunk_reqs, num_reqs - s.r0); }
    }
  };
  // EARLY: the six words are requested by volatile loads, which stay where they are written (a role asks for the
  // NEXT item's descriptor at the top of an item and uses it at the end: no global-memory latency between items)
  auto get_item = [&](const ItemIter& s, bool early = false) {
    DqItem it;
    it.hq0 = 2 * s.pr;
    it.hkv = it.hq0 / G;               // (dead code except in the K/V producer)
    const int* m = p.req_meta + (s.r0 + s.rr) * 6;
    if (early) {
      it.q_start = ld_early(m); it.q_len = ld_early(m + 1); it.kv_src = ld_early(m + 2);
      it.kv_len = ld_early(m + 3); it.first = ld_early(m + 4); it.pruned = ld_early(m + 5);
    } else {
      it.q_start = m[0]; it.q_len = m[1]; it.kv_src = m[2]; it.kv_len = m[3]; it.first = m[4]; it.pruned = m[5];
    }
    it.i0 = s.qt * kTM;
    it.ntiles = it.i0 < it.q_len ? (min(it.kv_len, it.i0 + kTM) + kTN - 1) / kTN : 0;
    return it;
  };
  auto peek_item = [&](ItemIter& s) {  // steps s once; that item, requested now (ntiles 0: past the end, or empty)
    DqItem it;
    it.ntiles = 0;
    it_next(s);
    if (s.k < total) it = get_item(s, true);
    return it;
  };
  auto seek_item = [&](ItemIter& s) {  // the first non-empty item at or after s (ntiles 0: none left)
    DqItem it;
    it.ntiles = 0;
    for (; s.k < total; it_next(s)) {
      it = get_item(s);
      if (it.ntiles != 0) break;
    }
    return it;
  };
  auto issue_kv = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* full, const DqItem& it, int t) {
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
    for (int s = 0; s < kQSets; ++s) { mbar_init(&bar[Q_QFULL + s], 1); mbar_init(&bar[Q_QFREE + s], 2); }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&bar[Q_KFULL + s], 1); mbar_init(&bar[Q_VFULL + s], 1);
      mbar_init(&bar[Q_KFREE + s], 2); mbar_init(&bar[Q_VFREE + s], 2);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(&bar[Q_SFULL + s], 1); mbar_init(&bar[Q_PFULL + s], kTM); mbar_init(&bar[Q_PVDONE + s], 1);
      mbar_init(&bar[Q_LFULL + s], kTM);
    }
    mbar_init(&bar[Q_OFREE], kTM); mbar_init(&bar[Q_OFREE + 1], kTM);
    mbar_fence_init();
  }
  if (warp == 9) t5::tmem_alloc<L::kTmemCols>(tmem_slot);
  t5::fence_before_sync();
  __syncthreads();
  t5::fence_after_sync();
  const uint32_t tmem = *tmem_slot;   // stream i: S buffers at tmem + i * 256 + sb * kTN, O at tmem + i * 256 + 128

  if (warp >= 8 && warp < 12) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsRolesDq));
    if (warp == 8) {
      // ================================================= K / V producer (one lane): runs ahead of the MMAs
      // by up to kStages tiles, across item boundaries
      if (t5::elect_one()) {
        DQ_TRACER(tr, 5, true);
        // K runs ONE TILE AHEAD of V: K0 K1 V0 K2 V1 ...  With K_g V_g K_g+1 in turn, K_g+1 waited behind V_g's
        // stage (free when P.V_g-3 is done, a tile later than S_g-2), which left K less than two tile periods
        // of lead against ~3000 clk of TMA latency under load.
        struct Cur { ItemIter pos, npos; DqItem it, nx; int t, g; };
        auto cur_begin = [&](Cur& c) {
          c.pos = it_begin(); c.it = seek_item(c.pos); c.t = 0; c.g = 0;
          c.npos = c.pos; c.nx.ntiles = 0;
          if (c.it.ntiles != 0) c.nx = peek_item(c.npos);        // the next item's descriptor, requested a whole item ahead
        };
        auto cur_step = [&](Cur& c) {                             // next tile (of the next non-empty item after an item's last)
          ++c.g;
          if (++c.t == c.it.ntiles) {
            if (c.nx.ntiles == 0 && c.npos.k < total) { it_next(c.npos); c.nx = seek_item(c.npos); }
            c.it = c.nx; c.pos = c.npos; c.t = 0;
            if (c.it.ntiles != 0) c.nx = peek_item(c.npos);
          }
        };
        Cur ck, cv;
        cur_begin(ck);
        cur_begin(cv);
        auto put_k = [&]() {
          const int st = ck.g % kStages, use = ck.g / kStages;
          if (ck.t == 0 && ck.it.pruned != 0 && (((ck.it.first | ck.it.pruned) & (BOX - 1)) != 0)) __trap();
          if (use > 0) t5::mbar_wait_relaxed(&bar[Q_KFREE + st], (use - 1) & 1);
          DQ_TR(tr, 30, ck.g);
          issue_kv(&tm_k, smem + L::oK + st * L::kKV, &bar[Q_KFULL + st], ck.it, ck.t);
          cur_step(ck);
        };
        if (ck.it.ntiles != 0) put_k();
        while (cv.it.ntiles != 0) {
          if (ck.it.ntiles != 0) put_k();
          const int st = cv.g % kStages, use = cv.g / kStages;
          if (use > 0) t5::mbar_wait_relaxed(&bar[Q_VFREE + st], (use - 1) & 1);
          DQ_TR(tr, 32, cv.g);
          issue_kv(&tm_v, smem + L::oV + st * L::kKV, &bar[Q_VFULL + st], cv.it, cv.t);
          cur_step(cv);
        }
      }
      __syncwarp();
    } else if (warp == 10) {
      // ================================================= Q producer (one lane): the pair of Q tiles of an item
      if (t5::elect_one()) {
        DQ_TRACER(tr, 6, true);
        int n = 0;           // non-empty items before this one
        ItemIter cur = it_begin();
        DqItem it = seek_item(cur);
        while (it.ntiles != 0) {
          it_next(cur);
          const DqItem nx = seek_item(cur);            // the next non-empty item (ntiles 0: none)
          // Q set n & 1 is free once the last S MMAs of item n - 2 (both streams) have completed.  The issuers
          // cannot start item n before this load, so q_free[n & 1] is never more than one phase ahead of this wait.
          const int qs = n & 1;
          if (n >= kQSets) t5::mbar_wait_relaxed(&bar[Q_QFREE + qs], ((n >> 1) - 1) & 1);
          DQ_TR(tr, 40, n);
          mbar_expect_tx(&bar[Q_QFULL + qs], 2 * L::kQ);
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int b = 0; b < D / 64; ++b)
              t5::tma_load_2d(smem + L::oQ + (qs * 2 + i) * L::kQ + b * (kTM * 128), &tm_q, (it.hq0 + i) * D + b * 64,
                              it.q_start + it.i0, &bar[Q_QFULL + qs]);
          // the NEXT item's Q pair (read exactly once: always a DRAM miss) goes to L2 now
          if (nx.ntiles != 0) {
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
              for (int b = 0; b < D / 64; ++b)
                t5::tma_prefetch_l2_2d(&tm_q, (nx.hq0 + i) * D + b * 64, nx.q_start + nx.i0);
            // (the next item's first K / V tile prefetched here as well measured 0.4-1.4 % SLOWER:
            // profiles/r02_prefill_dq_kvprefetch_ab.jsonl)
          }
          it = nx;
          ++n;
        }
      }
      __syncwarp();
    } else {
      // ================================================= MMA issuer of stream i (one lane; warp 9 -> 0, warp 11 -> 1)
      if (t5::elect_one()) {
        const int i = (warp == 9) ? 0 : 1;
        DQ_TRACER(tr, 2 + i, true);
        const uint32_t sQ0 = smem_u32(smem + L::oQ) + i * L::kQ, sK = smem_u32(smem + L::oK), sV = smem_u32(smem + L::oV);
        const uint32_t tmem_i = tmem + i * 256;
        // S_i(g) = Q_i K_g^T with the Q pair of set qs; the K stage is released when both streams' S MMAs have
        // completed (2 commits), the Q set likewise after an item's last S
        auto issue_s = [&](int g, int qs, bool last_of_item) {
          const int st = g % kStages;
          const uint32_t sQ = sQ0 + qs * (2 * L::kQ);
          mbar_wait(&bar[Q_KFULL + st], (g / kStages) & 1);
          DQ_TR(tr, 10, g);
          t5::fence_after_sync();
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            const uint64_t a = t5::smem_desc_sw128(sQ + (ks >> 2) * (kTM * 128) + (ks & 3) * 32, 16, 1024);
            const uint64_t b = t5::smem_desc_sw128(sK + st * L::kKV + (ks >> 2) * (kTN * 128) + (ks & 3) * 32, 16, 1024);
            t5::mma_f16_ss(tmem_i + (g & 1) * kTN, a, b, kIdescS, ks > 0);
          }
          t5::commit(&bar[Q_SFULL + i * 2 + (g & 1)]);
          t5::commit(&bar[Q_KFREE + st]);
          if (last_of_item) t5::commit(&bar[Q_QFREE + qs]);
        };
        int g = 0, n_item = 0;
        ItemIter cur = it_begin();
        int ntiles = seek_item(cur).ntiles;
        bool have = ntiles != 0;
        if (have) {
          mbar_wait(&bar[Q_QFULL], 0);
          issue_s(0, 0, ntiles == 1);
        }
        while (have) {
          int nt_nx = 0;
          bool have_nx = false;
          ItemIter nxi = cur;
          const int nt_peek = peek_item(nxi).ntiles;  // requested now, used at this item's last tile
          for (int j = 0; j < ntiles; ++j, ++g) {
            const int st = g % kStages, sb = g & 1;
            if (j + 1 < ntiles) {
              issue_s(g + 1, n_item & 1, j + 2 == ntiles);
            } else {
              nt_nx = nt_peek;
              if (nt_nx == 0 && nxi.k < total) {  // an empty item (a query tile past a short prompt): look further
                it_next(nxi);
                nt_nx = seek_item(nxi).ntiles;
              }
              have_nx = nt_nx != 0;
              if (have_nx) {                      // the next item's first tile: the other Q set, loaded during this item
                DQ_TR(tr, 15, g);
                mbar_wait(&bar[Q_QFULL + ((n_item + 1) & 1)], ((n_item + 1) >> 1) & 1);
                DQ_TR(tr, 16, g);
                issue_s(g + 1, (n_item + 1) & 1, nt_nx == 1);
              }
            }
            DQ_TR(tr, 11, g);
            mbar_wait(&bar[Q_VFULL + st], (g / kStages) & 1);
            DQ_TR(tr, 12, g);
            mbar_wait(&bar[Q_PFULL + i * 2 + sb], (g >> 1) & 1);
            DQ_TR(tr, 13, g);
            // the first P.V of an item overwrites O_i: the epilogue must have read the previous item's
            if (j == 0 && n_item > 0) mbar_wait(&bar[Q_OFREE + i], (n_item - 1) & 1);
            DQ_TR(tr, 14, g);
            t5::fence_after_sync();
#pragma unroll
            for (int kk = 0; kk < PK; ++kk) {
              const uint64_t b = t5::smem_desc_sw128(sV + st * L::kKV + kk * (16 * 128), kTN * 128, 1024);
              // P_i(g) lives in the first kTN / 2 columns of the S buffer it was computed from
              t5::mma_f16_ts(tmem_i + 128, tmem_i + sb * kTN + kk * 8, b, kIdescPV, (j > 0 || kk > 0));
            }
            t5::commit(&bar[Q_PVDONE + i * 2 + sb]);
            t5::commit(&bar[Q_VFREE + st]);
          }
          ++n_item;
          ntiles = nt_nx; cur = nxi; have = have_nx;
        }
      }
      __syncwarp();
    }
  } else if (warp < 8) {
    // ================================================= softmax warpgroups (stream i = warp / 4, thread = query row)
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmaxDq));
    const int i = warp >> 2;
    const int rt = tid & 127;                          // row of the tile = TMEM lane
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tmem_s = tmem + i * 256 + lane_base;
    const uint32_t tmem_o = tmem + i * 256 + 128 + lane_base;
    int g = 0, n_item = 0;
    DQ_TRACER(tr, i, (warp & 3) == 0 && lane == 0);
    ItemIter cur = it_begin();
    DqItem it = seek_item(cur);
    while (it.ntiles != 0) {
      ItemIter nxi = cur;
      DqItem nx = peek_item(nxi);                      // requested now, used when this item is done
      const int row = it.i0 + rt;
      float m_use = -INFINITY, l_run = 0.f;
      for (int j = 0; j < it.ntiles; ++j, ++g) {
        const int sb = g & 1;
        const int j0 = j * kTN;
        DQ_TR(tr, j == 0 ? 1 : 2, g);
        mbar_wait(&bar[Q_SFULL + i * 2 + sb], (g >> 1) & 1);
        DQ_TR(tr, 3, g);
        t5::fence_after_sync();
        uint32_t sr[SC][32];
#pragma unroll
        for (int c = 0; c < SC; ++c) t5::tmem_ld32(sr[c], tmem_s + sb * kTN + c * 32);
        t5::tmem_ld_wait();
        const bool need_mask = (j0 + kTN - 1 > it.i0) || (j0 + kTN > it.kv_len);
        const int lim = need_mask ? min(row, it.kv_len - 1) - j0 : kTN;  // columns > lim are masked
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < SC; ++c) {
            if (__all_sync(kFull, lim >= c * 32 + 31)) continue;
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (c * 32 + e > lim) sr[c][e] = 0xff800000u;  // -inf
          }
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < SC; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e) mx[e & 3] = fmaxf(mx[e & 3], __uint_as_float(sr[c][e]));
        const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
        float alpha = 1.f;
        const bool grow = m_tile > m_use + kRescaleThresholdDq || m_use == -INFINITY;
        if (grow) {
          alpha = (m_use == -INFINITY) ? 0.f : fast_exp2(m_use - m_tile);
          m_use = m_tile;
          l_run *= alpha;
        }
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        const float neg_m = -m_use;
#pragma unroll
        for (int c = 0; c < SC; ++c) {
          // a chunk that is masked for every row of the warp (the upper rows of an item's diagonal tiles): P = 0
          // without the exponentials (exp2(-inf) = 0: the same bits)
          if (need_mask && __all_sync(kFull, lim < c * 32)) {
#pragma unroll
            for (int e = 0; e < 16; ++e) sr[c][e] = 0u;
            continue;
          }
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            const float p0 = fast_exp2(fmaf(__uint_as_float(sr[c][e]), p.scale_log2, neg_m));
            const float p1 = fast_exp2(fmaf(__uint_as_float(sr[c][e + 1]), p.scale_log2, neg_m));
            sum[(e >> 1) & 3] += p0 + p1;
            sr[c][e >> 1] = pack_bf16x2(p0, p1);
          }
        }
        l_run += (sum[0] + sum[1]) + (sum[2] + sum[3]);
        // P_g has its own buffer (the S buffer it came from): only a rescale of O has to wait for the
        // previous P.V of this stream
        if (j > 0 && __any_sync(kFull, grow)) {
          mbar_wait(&bar[Q_PVDONE + i * 2 + ((g - 1) & 1)], ((g - 1) >> 1) & 1);
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
        // packed P row -> TMEM over the scores it came from: 32 columns = 64 bf16 per store
#pragma unroll
        for (int h = 0; h < SC / 2; ++h) {
          uint32_t pk[32];
#pragma unroll
          for (int e = 0; e < 16; ++e) { pk[e] = sr[2 * h][e]; pk[16 + e] = sr[2 * h + 1][e]; }
          t5::tmem_st32(tmem_s + sb * kTN + h * 32, pk);
        }
        t5::tmem_st_wait();
        t5::fence_before_sync();
        t5::mbar_arrive(&bar[Q_PFULL + i * 2 + sb]);
        DQ_TR(tr, 4, g);
      }
      // 1 / l to the epilogue warps (l_full is double-buffered by item parity: a one-tile item lets this
      // warpgroup finish item n + 1 before the epilogue has looked at item n)
      linv[((n_item & 1) * 2 + i) * kTM + rt] = 1.f / l_run;
      t5::mbar_arrive(&bar[Q_LFULL + i * 2 + (n_item & 1)]);
      ++n_item;
      if (nx.ntiles == 0 && nxi.k < total) {           // an empty item: look further
        it_next(nxi);
        nx = seek_item(nxi);
      }
      it = nx;
      cur = nxi;
    }
  } else {
    // ================================================= epilogue warpgroup: O_i / l -> bf16 -> global, i = 0, 1
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsEpiDq));
    const int rt = tid & 127;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    int g = 0, n_item = 0;
    DQ_TRACER(tr, 4, (warp & 3) == 0 && lane == 0);
    for (ItemIter cur = it_begin(); cur.k < total; it_next(cur)) {
      const DqItem it = get_item(cur);
      if (it.ntiles == 0) continue;
      g += it.ntiles;
      const int row = it.i0 + rt;
#pragma unroll 1
      for (int i = 0; i < 2; ++i) {
        DQ_TR(tr, 20 + i, g - 1);
        t5::mbar_wait_sleep(&bar[Q_LFULL + i * 2 + (n_item & 1)], (n_item >> 1) & 1, 128);
        DQ_TR(tr, 22 + i, g - 1);
        const float inv = linv[((n_item & 1) * 2 + i) * kTM + rt];
        t5::mbar_wait_sleep(&bar[Q_PVDONE + i * 2 + ((g - 1) & 1)], ((g - 1) >> 1) & 1, 64);
        DQ_TR(tr, 24 + i, g - 1);
        t5::fence_after_sync();
        __nv_bfloat16* dst = p.out + (size_t)(it.q_start + (row < it.q_len ? row : 0)) * p.out_stride +
                             (size_t)(it.hq0 + i) * D;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t o[32];
          t5::tmem_ld32(o, tmem + i * 256 + 128 + lane_base + c * 32);
          t5::tmem_ld_wait();
          if (c == D / 32 - 1) {          // O_i is in registers: the next item's first P.V may overwrite it
            t5::fence_before_sync();
            t5::mbar_arrive(&bar[Q_OFREE + i]);
            DQ_TR(tr, 26 + i, g - 1);
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

// tensor maps: prefill_attn_tc.cu
int prefill_make_rows_map(CUtensorMap* m, const void* base, long long rows, int inner, long long stride, int box_rows);

namespace {

template <int D, int BOX>
int launch_prefill_dq(const DqParams& p, const PrefillArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  if (int rc = prefill_make_rows_map(&tq, a.q, a.num_tokens, a.Hq * D, a.q_stride, kTM)) return rc;
  if (int rc = prefill_make_rows_map(&tk, a.k, a.num_tokens, a.Hkv * D, a.key_stride, BOX)) return rc;
  if (int rc = prefill_make_rows_map(&tv, a.v, a.num_tokens, a.Hkv * D, a.value_stride, BOX)) return rc;
  const int smem = DqSmem<D>::kBytes;
  auto kern = prefill_attn_dq_kernel<D, BOX>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(prefill_dq)", e);
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
  kern<<<ctas, kDqThreads, smem, stream>>>(p, tq, tk, tv, q_tiles, a.num_reqs, (int)chunk);
  return check_launch("prefill_attn_dq_kernel");
}

}  // namespace

#ifdef PC_DEBUG
int prefill_dq_trace(int on, void* host_out, size_t bytes) {   // on: set the flag; host_out != NULL: copy the stamps out
  cudaError_t e = cudaMemcpyToSymbol(g_dq_trace_on, &on, sizeof(int));
  if (e == cudaSuccess && host_out != nullptr) {
    const size_t all = sizeof(unsigned long long) * kDqTraceRoles * kDqTraceCap;
    e = cudaMemcpyFromSymbol(host_out, g_dq_trace, bytes < all ? bytes : all);
  }
  return e == cudaSuccess ? 0 : set_cuda_error("prefill_dq_trace", e);
}
#endif

bool prefill_attn_dq_supported(const PrefillArgs& a) {
  return (a.D == 64 || a.D == 128) && a.Hkv > 0 && a.Hq % a.Hkv == 0 && ((a.Hq / a.Hkv) % 2) == 0;
}

// arguments were validated by prefill_attn() (prefill_attn_tc.cu); box8: 8-row TMA boxes (block size 8)
int prefill_attn_dq(const PrefillArgs& a, bool box8, cudaStream_t stream) {
  DqParams p;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.req_meta = a.req_meta;
  p.Hq = a.Hq; p.Hkv = a.Hkv;
  p.out_stride = a.out_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  if (a.D == 128) return box8 ? launch_prefill_dq<128, 8>(p, a, stream) : launch_prefill_dq<128, 16>(p, a, stream);
  return box8 ? launch_prefill_dq<64, 8>(p, a, stream) : launch_prefill_dq<64, 16>(p, a, stream);
}

}  // namespace pc
