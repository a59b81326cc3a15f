// Paged decode attention over (compacted) block tables with value_l2 block scoring
// fused into the same KV read.  sm_100a.
//
// Replaces csrc/attention/attention_kernels.cuh:86-668 (paged_attention_v1/v2 + reduce)
// and the extra V read of torch_cuda_sdpa.py:1518-1526 (last-block scoring).
//
// Design (HBM-bound; see DESIGN.md "decode kernel"):
//   * unit of work = one PAGE = (sequence b, kv head h [, 8-q-head group hg], block):
//     two contiguous chunks of the reference cache layout (K [D/8][BS][8], V [D][BS],
//     vllm/attention/ops/paged_attn.py:49-62), fetched with two TMA 1-D bulk copies
//     (cp.async.bulk -> UBLKCP) that complete on one mbarrier.
//   * stream-K schedule: all pages of the launch are linearised (b, h, hg, block) and
//     cut into W equal contiguous ranges, one per WARP (W = SMs x kWarps, interleaved
//     over SMs so every SM carries the same bytes).  A warp walks its range with a
//     kStages-deep page ring that never drains at sequence boundaries; there is no work
//     counter and no __syncthreads after the prologue.  kv_lens live on the device, so
//     every CTA derives the same schedule from a prefix sum of blocks per row.
//   * the whole GQA group (<= 8 q heads) is served from ONE read of the page:
//     S^T[16 tok x 8 heads] = K_tile(A, ldmatrix straight from the cache layout)
//     x Q^T(B, registers) with mma.sync m16n8k16 bf16/fp32; online softmax with warp
//     shuffles in the exp2 domain; P^T is transposed in registers (movmatrix) and
//     O^T[D x 8 heads] += V^T_tile(A, ldmatrix straight from the cache layout) x P^T.
//     Tensor cores are used because 2*G FMAs per KV byte exceed the SIMT fp32 budget
//     at HBM speed for G >= 4; the legacy HMMA path is far from saturated (HBM-bound).
//   * numerics follow attention_kernels.cuh: fp32 logits, probabilities rounded to
//     bf16 before P.V, fp32 accumulation, final 1/(sum + 1e-6).
//   * a (sequence, head) whose pages straddle ranges is finished stream-K style: every
//     contributor publishes (m, l, O) fp32 to its own workspace slot (<= 2 per warp) and
//     the LAST one to arrive combines them in range order (deterministic), warp-parallel
//     over heads x dims.  Arrival tickets are reset by the combiner, so the op is a
//     single launch with no memset.
//   * scoring: when seq_lens[b] % BS == 0 the row's last block just filled; the warp
//     that owns that page computes the per-head value_l2 partial from the V page it
//     already has in shared memory (zero extra HBM bytes) and stores it at
//     score_table[slot][nblk-1][h].
#include "pc_common.cuh"
#include "score_common.cuh"
#include "pc_api_internal.h"

namespace pc {

struct DecodeParams {
  const __nv_bfloat16* __restrict__ q;
  const __nv_bfloat16* __restrict__ k_cache;
  const __nv_bfloat16* __restrict__ v_cache;
  const int* __restrict__ block_tables;
  const int* __restrict__ kv_lens;
  __nv_bfloat16* __restrict__ out;
  float* part_o;    // [2 * max_workers][8][D]
  float* part_ml;   // [2 * max_workers][8][2]
  int* tickets;     // [B * GH] arrival counters, zero between launches
  const int* __restrict__ seq_lens;     // nullptr: no scoring
  const int* __restrict__ score_slots;  // [B]
  float* score_table;                   // [slots][score_max_blocks][Hkv]
  int score_max_blocks;
  int B, Hq, Hkv, G, HG, GH;  // G = Hq/Hkv, HG = ceil(G/8), GH = Hkv*HG groups per row
  long long q_stride;         // elements between sequences in q
  long long out_stride;       // elements between sequences in out
  long long table_stride;     // ints between rows
  long long kv_block_stride;  // elements between blocks in k/v cache
  float scale_log2;
  int min_quantum;            // fewest pages a warp is given (tiny launches use fewer warps)
  unsigned long long* trace;  // debug: [warps][4] globaltimer stamps (nullptr = off)
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

template <int D, int BS, int kWarps, int kStages>
struct DecodeTraits {
  static constexpr int kPageBytes = BS * D * 2;        // K or V page of one (block, head)
  static constexpr int kStageBytes = 2 * kPageBytes;   // K + V
  static constexpr int kRingBytes = kWarps * kStages * kStageBytes;
  static constexpr int kBarBytes = kWarps * kStages * 8;
  static constexpr int kThreads = kWarps * 32;
  static size_t smem_bytes(int B) { return (size_t)kRingBytes + kBarBytes + ((size_t)B + 1) * 4; }
};

// Position of one page in the (row, group, block) linearisation.
struct PageCursor {
  int b, gh, blk, nblk;
};

__device__ __forceinline__ void cursor_seek(PageCursor& c, long long page, const int* prefix, int B,
                                            int GH) {
  const int key = (int)(page / GH);
  int lo = 0, hi = B;  // largest b with prefix[b] <= key  (prefix[B] > key)
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (prefix[mid] <= key) lo = mid; else hi = mid;
  }
  c.b = lo;
  c.nblk = prefix[lo + 1] - prefix[lo];
  const int rem = (int)(page - (long long)GH * prefix[lo]);
  c.gh = rem / c.nblk;
  c.blk = rem - c.gh * c.nblk;
}

__device__ __forceinline__ void cursor_next(PageCursor& c, const int* prefix, int B, int GH) {
  if (++c.blk < c.nblk) return;
  c.blk = 0;
  if (++c.gh < GH) return;
  c.gh = 0;
  do { ++c.b; } while (c.b < B && prefix[c.b + 1] == prefix[c.b]);
  c.nblk = (c.b < B) ? prefix[c.b + 1] - prefix[c.b] : 1;
}

// first page of worker w when P pages are cut into W ranges
__device__ __forceinline__ long long range_lo(long long w, long long P, long long W) { return w * P / W; }
// worker whose range holds page p
__device__ __forceinline__ int worker_of(long long p, long long P, long long W) {
  return (int)(((p + 1) * W + P - 1) / P - 1);
}

template <int D, int BS, int kWarps, int kStages>
__global__ void __launch_bounds__(kWarps * 32, 1)
paged_decode_attn_score_kernel(const DecodeParams p) {
  using T = DecodeTraits<D, BS, kWarps, kStages>;
  constexpr int KSTEPS = D / 16;   // k-steps of the QK^T contraction
  constexpr int MT = D / 16;       // m-tiles (16 dims) of O^T
  constexpr int TT = BS / 16;      // 16-token tiles per page

  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;  // fragment row group
  const int t = lane & 3;   // fragment column pair

  const unsigned long long t_start = p.trace ? gtimer() : 0ull;
  uint8_t* my_stages = smem + (size_t)warp * kStages * T::kStageBytes;
  uint64_t* my_bars = reinterpret_cast<uint64_t*>(smem + T::kRingBytes) + warp * kStages;
  int* prefix = reinterpret_cast<int*>(smem + T::kRingBytes + T::kBarBytes);  // [B + 1]

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&my_bars[s], 1);
  }
  // ---- schedule: blocks per row -> exclusive prefix sum (same in every CTA)
  if (warp == 0) {
    int carry = 0;
    if (lane == 0) prefix[0] = 0;
    for (int base = 0; base < p.B; base += 32) {
      const int r = base + lane;
      const int len = (r < p.B) ? p.kv_lens[r] : 0;
      int n = (len + BS - 1) / BS;
      if (r < p.B && len == 0 && blockIdx.x == 0) {  // empty sequence: defined output (zeros)
        __nv_bfloat16* o = p.out + (size_t)r * p.out_stride;
        for (int i = 0; i < p.Hq * D; ++i) o[i] = __float2bfloat16(0.f);
      }
#pragma unroll
      for (int sh = 1; sh < 32; sh <<= 1) {
        const int v = __shfl_up_sync(kFull, n, sh);
        if (lane >= sh) n += v;
      }
      if (r < p.B) prefix[r + 1] = carry + n;
      carry += __shfl_sync(kFull, n, 31);
    }
  }
  mbar_fence_init();
  __syncthreads();

  const long long P = (long long)prefix[p.B] * p.GH;  // pages in this launch
  if (P == 0) return;
  const long long Wmax = (long long)gridDim.x * kWarps;
  long long W = (P + p.min_quantum - 1) / p.min_quantum;
  if (W > Wmax) W = Wmax;
  const long long w = (long long)warp * gridDim.x + blockIdx.x;  // interleaved over SMs
  if (w >= W) return;
  const long long lo = range_lo(w, P, W);
  const long long hi = range_lo(w + 1, P, W);
  const int npages = (int)(hi - lo);
  if (npages <= 0) return;

  PageCursor ci, cc;  // issue / consume cursors
  cursor_seek(ci, lo, prefix, p.B, p.GH);
  cc = ci;

  auto table_entry = [&](const PageCursor& c) {  // uniform address: one broadcast load
    return __ldg(p.block_tables + (size_t)c.b * p.table_stride + c.blk);
  };
  auto issue = [&](int slot, const PageCursor& c, int phys) {  // lane 0 only
    const int h = c.gh / p.HG;
    uint8_t* dst = my_stages + (size_t)slot * T::kStageBytes;
    const size_t off = ((size_t)phys * p.kv_block_stride) + (size_t)h * (D * BS);
    mbar_expect_tx(&my_bars[slot], T::kStageBytes);
    tma_load_1d(dst, p.k_cache + off, T::kPageBytes, &my_bars[slot]);
    tma_load_1d(dst + T::kPageBytes, p.v_cache + off, T::kPageBytes, &my_bars[slot]);
  };

  // prologue: fill the ring (slot of page i = i % kStages, parity (i / kStages) & 1)
  // The table entry of the NEXT page to issue is always loaded one page ahead.
  int issued = 0;
  int phys_next = table_entry(ci);
#pragma unroll
  for (int s = 0; s < kStages; ++s) {
    if (s < npages) {
      if (lane == 0) issue(s, ci, phys_next);
      cursor_next(ci, prefix, p.B, p.GH);
      if (++issued < npages) phys_next = table_entry(ci);
    }
  }

  // Q^T as the mma B operand: b0 = Q[head g][16j + 2t..], b1 = Q[head g][16j + 8 + 2t..]
  auto load_q = [&](uint32_t (&qf)[KSTEPS][2], int b, int gh) {
    const int h = gh / p.HG, hg = gh - h * p.HG;
    const int head0 = h * p.G + hg * 8;
    const int nheads = min(8, p.G - hg * 8);
    const bool valid = g < nheads;
    const __nv_bfloat16* qrow = p.q + (size_t)b * p.q_stride + (size_t)(head0 + (valid ? g : 0)) * D;
#pragma unroll
    for (int j = 0; j < KSTEPS; ++j) {
      qf[j][0] = valid ? __ldg(reinterpret_cast<const unsigned int*>(qrow + 16 * j + 2 * t)) : 0u;
      qf[j][1] = valid ? __ldg(reinterpret_cast<const unsigned int*>(qrow + 16 * j + 8 + 2 * t)) : 0u;
    }
  };

  uint32_t qf[KSTEPS][2];
  load_q(qf, cc.b, cc.gh);
  if (p.trace && lane == 0) p.trace[w * 4 + 0] = t_start, p.trace[w * 4 + 1] = gtimer();

  float o[MT][4];
  float m_run[2], l_run[2];
  int seg_begin = cc.blk;      // first block of the current segment
  bool first_seg = true;       // the worker's first segment (decides its workspace slot)
  auto reset_acc = [&]() {
#pragma unroll
    for (int m = 0; m < MT; ++m) o[m][0] = o[m][1] = o[m][2] = o[m][3] = 0.f;
    m_run[0] = m_run[1] = -INFINITY;
    l_run[0] = l_run[1] = 0.f;
  };
  reset_acc();

  for (int pi = 0; pi < npages; ++pi) {
    const uint32_t slot = (uint32_t)pi % kStages;
    const uint32_t parity = ((uint32_t)pi / kStages) & 1u;
    const int b = cc.b, nblk = cc.nblk, blk = cc.blk;
    const int h = cc.gh / p.HG, hg = cc.gh - h * p.HG;
    const int kv_len = __ldg(p.kv_lens + b);
    mbar_wait(&my_bars[slot], parity);

    const uint8_t* kpage = my_stages + (size_t)slot * T::kStageBytes;
    const uint8_t* vpage = kpage + T::kPageBytes;
    const uint32_t kaddr = smem_u32(kpage);
    const uint32_t vaddr = smem_u32(vpage);
    const int page_tok0 = blk * BS;
    const bool last_partial = (page_tok0 + BS > kv_len);

    if (last_partial) {
      // zero V columns of tokens >= kv_len so 0 * garbage cannot produce NaN
      // (attention_kernels.cuh:455-466 does the same per element)
      const int valid = kv_len - page_tok0;
      __nv_bfloat16* vp = reinterpret_cast<__nv_bfloat16*>(const_cast<uint8_t*>(vpage));
      for (int i = lane; i < D * BS; i += 32)
        if ((i % BS) >= valid) vp[i] = __float2bfloat16(0.f);
      fence_proxy_async();  // generic-proxy writes before a later TMA refill of this slot
      __syncwarp();
    }

#pragma unroll
    for (int tt = 0; tt < TT; ++tt) {
      // ---------------- S^T = K_tile x Q^T
      float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
      {
        // ldmatrix.x4 row address: matrix mi = lane/8 -> (token half = mi&1, chunk = 2j + mi>>1)
        const int mi = lane >> 3, r = lane & 7;
        const uint32_t base = kaddr + (uint32_t)(((mi >> 1) * BS + tt * 16 + (mi & 1) * 8 + r) * 16);
#pragma unroll
        for (int j = 0; j < KSTEPS; ++j) {
          uint32_t a[4];
          ldmatrix_x4(a, base + (uint32_t)(2 * j * BS * 16));
          if (j & 1) mma_bf16_16816(s1, a, qf[j][0], qf[j][1]);
          else       mma_bf16_16816(s0, a, qf[j][0], qf[j][1]);
        }
      }
      // s[0]=(tok g, head 2t) s[1]=(tok g, head 2t+1) s[2]=(tok g+8, head 2t) s[3]=(tok g+8, 2t+1)
      float x0 = (s0[0] + s1[0]) * p.scale_log2, x1 = (s0[1] + s1[1]) * p.scale_log2;
      float x2 = (s0[2] + s1[2]) * p.scale_log2, x3 = (s0[3] + s1[3]) * p.scale_log2;
      if (last_partial) {
        const int tok = page_tok0 + tt * 16 + g;
        if (tok >= kv_len) { x0 = -INFINITY; x1 = -INFINITY; }
        if (tok + 8 >= kv_len) { x2 = -INFINITY; x3 = -INFINITY; }
      }
      // ---------------- online softmax (per head = per column; rows live in lanes with equal t)
      float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
#pragma unroll
      for (int sh = 4; sh <= 16; sh <<= 1) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(kFull, mx0, sh));
        mx1 = fmaxf(mx1, __shfl_xor_sync(kFull, mx1, sh));
      }
      const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
      const float al0 = fast_exp2(m_run[0] - mn0), al1 = fast_exp2(m_run[1] - mn1);
      m_run[0] = mn0; m_run[1] = mn1;
      const float p0 = fast_exp2(x0 - mn0), p1 = fast_exp2(x1 - mn1);
      const float p2 = fast_exp2(x2 - mn0), p3 = fast_exp2(x3 - mn1);
      l_run[0] = l_run[0] * al0 + (p0 + p2);
      l_run[1] = l_run[1] * al1 + (p1 + p3);
      if (__any_sync(kFull, (al0 != 1.f) || (al1 != 1.f))) {
#pragma unroll
        for (int m = 0; m < MT; ++m) { o[m][0] *= al0; o[m][1] *= al1; o[m][2] *= al0; o[m][3] *= al1; }
      }
      // ---------------- P^T as the B operand: transpose the two 8x8 (tok x head) tiles
      const uint32_t pb0 = movmatrix_trans(pack_bf16x2(p0, p1));  // tokens 0..7
      const uint32_t pb1 = movmatrix_trans(pack_bf16x2(p2, p3));  // tokens 8..15
      // ---------------- O^T += V^T_tile x P^T
      {
        // matrix mi -> (dim half = mi&1, token half = mi>>1); row r = dim within the half
        const int mi = lane >> 3, r = lane & 7;
        const uint32_t base = vaddr + (uint32_t)((((mi & 1) * 8 + r) * BS + tt * 16 + (mi >> 1) * 8) * 2);
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          uint32_t a[4];
          ldmatrix_x4(a, base + (uint32_t)(m * 16 * BS * 2));
          mma_bf16_16816(o[m], a, pb0, pb1);
        }
      }
    }

    // ---------------- fused value_l2 score of the row's last (just filled) block
    if (p.seq_lens != nullptr && blk == nblk - 1 && hg == 0 && (kv_len % BS) == 0 &&
        (__ldg(p.seq_lens + b) % BS) == 0) {
      const __nv_bfloat16* vp = reinterpret_cast<const __nv_bfloat16*>(vpage);
      const float sc = warp_value_l2_head_score<D, BS>(
          [&](int tok, int d) { return __bfloat162float(vp[d * BS + tok]); });
      if (lane == 0)
        p.score_table[((size_t)__ldg(p.score_slots + b) * p.score_max_blocks + (nblk - 1)) * p.Hkv + h] = sc;
    }

    // ---------------- refill this slot with page pi + kStages
    __syncwarp();
    if (issued < npages) {
      if (lane == 0) issue((int)slot, ci, phys_next);
      cursor_next(ci, prefix, p.B, p.GH);
      if (++issued < npages) phys_next = table_entry(ci);
    }

    if (p.trace && pi == npages - 1 && lane == 0) p.trace[w * 4 + 2] = gtimer();
    // ---------------- segment end: the (row, group) is done, or this warp's range is
    const bool group_end = (blk == nblk - 1);
    if (!group_end && pi != npages - 1) {
      ++cc.blk;
      continue;
    }
    const int head0 = h * p.G + hg * 8;       // first q head of this group
    const int nheads = min(8, p.G - hg * 8);  // valid q heads in the group
    // l over the 8 row-lanes of each column
#pragma unroll
    for (int sh = 4; sh <= 16; sh <<= 1) {
      l_run[0] += __shfl_xor_sync(kFull, l_run[0], sh);
      l_run[1] += __shfl_xor_sync(kFull, l_run[1], sh);
    }
    // o[m][0]=(dim 16m+g, head 2t) [1]=(dim 16m+g, head 2t+1) [2]=(dim 16m+g+8, 2t) [3]=(.., 2t+1)
    if (seg_begin == 0 && group_end) {  // the whole group was ours: final result
      const float inv0 = 1.f / (l_run[0] + 1e-6f), inv1 = 1.f / (l_run[1] + 1e-6f);
      __nv_bfloat16* ob = p.out + (size_t)b * p.out_stride + (size_t)head0 * D;
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        if (2 * t < nheads) {
          ob[(size_t)(2 * t) * D + 16 * m + g] = __float2bfloat16_rn(o[m][0] * inv0);
          ob[(size_t)(2 * t) * D + 16 * m + g + 8] = __float2bfloat16_rn(o[m][2] * inv0);
        }
        if (2 * t + 1 < nheads) {
          ob[(size_t)(2 * t + 1) * D + 16 * m + g] = __float2bfloat16_rn(o[m][1] * inv1);
          ob[(size_t)(2 * t + 1) * D + 16 * m + g + 8] = __float2bfloat16_rn(o[m][3] * inv1);
        }
      }
    } else {
      // ---- stream-K fix-up: publish the partial, last arriver combines
      const int my_slot = 2 * (int)w + (first_seg ? 0 : 1);
      {
        float* po = p.part_o + (size_t)my_slot * 8 * D;
        float* pml = p.part_ml + (size_t)my_slot * 16;
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          po[(size_t)(2 * t) * D + 16 * m + g] = o[m][0];
          po[(size_t)(2 * t) * D + 16 * m + g + 8] = o[m][2];
          po[(size_t)(2 * t + 1) * D + 16 * m + g] = o[m][1];
          po[(size_t)(2 * t + 1) * D + 16 * m + g + 8] = o[m][3];
        }
        if (g == 0) {
          pml[(2 * t) * 2 + 0] = m_run[0];
          pml[(2 * t) * 2 + 1] = l_run[0];
          pml[(2 * t + 1) * 2 + 0] = m_run[1];
          pml[(2 * t + 1) * 2 + 1] = l_run[1];
        }
      }
      const long long gs = (long long)p.GH * prefix[b] + (long long)cc.gh * nblk;  // group's first page
      const int w_first = worker_of(gs, P, W);
      const int w_last = worker_of(gs + nblk - 1, P, W);
      const int n_contrib = w_last - w_first + 1;
      __threadfence();
      __syncwarp();
      int* tk = p.tickets + (size_t)b * p.GH + cc.gh;
      int ticket = 0;
      if (lane == 0) ticket = atomicAdd(tk, 1);
      ticket = __shfl_sync(kFull, ticket, 0);
      if (ticket == n_contrib - 1) {
        if (lane == 0) *tk = 0;  // leave the workspace clean for the next launch
        __threadfence();
        const int hh = lane >> 2;             // head of the group this lane combines
        const int d0 = (lane & 3) * (D / 4);  // its D/4 contiguous dims
        float M = -INFINITY;
        for (int s = 0; s < n_contrib; ++s) {
          const long long ws = w_first + s;
          const int sl = 2 * (int)ws + (range_lo(ws, P, W) >= gs ? 0 : 1);
          M = fmaxf(M, __ldcg(p.part_ml + (size_t)sl * 16 + hh * 2));
        }
        float L = 0.f;
        float acc[D / 4];
#pragma unroll
        for (int i = 0; i < D / 4; ++i) acc[i] = 0.f;
#pragma unroll 4
        for (int s = 0; s < n_contrib; ++s) {
          const long long ws = w_first + s;
          const int sl = 2 * (int)ws + (range_lo(ws, P, W) >= gs ? 0 : 1);
          const float ms = __ldcg(p.part_ml + (size_t)sl * 16 + hh * 2);
          const float ls = __ldcg(p.part_ml + (size_t)sl * 16 + hh * 2 + 1);
          const float wgt = fast_exp2(ms - M);
          L += wgt * ls;
          const float4* src = reinterpret_cast<const float4*>(p.part_o + ((size_t)sl * 8 + hh) * D + d0);
#pragma unroll
          for (int i = 0; i < D / 16; ++i) {
            const float4 v = __ldcg(src + i);
            acc[4 * i + 0] += wgt * v.x;
            acc[4 * i + 1] += wgt * v.y;
            acc[4 * i + 2] += wgt * v.z;
            acc[4 * i + 3] += wgt * v.w;
          }
        }
        if (hh < nheads) {
          const float inv = 1.f / (L + 1e-6f);
          uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)b * p.out_stride + (size_t)(head0 + hh) * D + d0);
#pragma unroll
          for (int i = 0; i < D / 32; ++i) {
            uint4 v;
            v.x = pack_bf16x2(acc[8 * i + 0] * inv, acc[8 * i + 1] * inv);
            v.y = pack_bf16x2(acc[8 * i + 2] * inv, acc[8 * i + 3] * inv);
            v.z = pack_bf16x2(acc[8 * i + 4] * inv, acc[8 * i + 5] * inv);
            v.w = pack_bf16x2(acc[8 * i + 6] * inv, acc[8 * i + 7] * inv);
            dst[i] = v;
          }
        }
      }
    }
    // ---- next segment
    if (pi != npages - 1) {
      cursor_next(cc, prefix, p.B, p.GH);
      seg_begin = 0;
      first_seg = false;
      load_q(qf, cc.b, cc.gh);
      reset_acc();
    }
  }
  if (p.trace && lane == 0) p.trace[w * 4 + 3] = gtimer();
}

// ---------------------------------------------------------------------------- host
// Workspace layout (fixed offsets so the ticket region stays zero between launches
// whatever the batch): [part_ml: 2*kMaxWorkers*16 f32][part_o: 2*kMaxWorkers*8*128 f32][tickets: B*GH i32]
constexpr int kMaxWarps = 8;
static unsigned long long* g_decode_trace = nullptr;
void set_decode_trace(void* buf) { g_decode_trace = static_cast<unsigned long long*>(buf); }
static size_t ws_workers(int num_sms) { return (size_t)num_sms * kMaxWarps; }
static size_t ws_off_part_o(int num_sms) { return (2 * ws_workers(num_sms) * 16 * sizeof(float) + 255) / 256 * 256; }
static size_t ws_off_tickets(int num_sms) {
  return ws_off_part_o(num_sms) + (2 * ws_workers(num_sms) * 8 * 128 * sizeof(float) + 255) / 256 * 256;
}

size_t decode_workspace_bytes(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms) {
  (void)D; (void)BS; (void)max_kv_len;
  const int G = Hkv > 0 ? Hq / Hkv : 1;
  const int HG = (G + 7) / 8;
  return ws_off_tickets(num_sms) + ((size_t)B * Hkv * HG * sizeof(int) + 255) / 256 * 256;
}

template <int D, int BS, int kWarps, int kStages>
static int launch_decode(const DecodeParams& p, int num_sms, int max_blocks, cudaStream_t stream) {
  using T = DecodeTraits<D, BS, kWarps, kStages>;
  static_assert(kWarps <= kMaxWarps, "workspace is sized for kMaxWarps warps per SM");
  auto kern = paged_decode_attn_score_kernel<D, BS, kWarps, kStages>;
  const size_t smem = T::smem_bytes(p.B);
  if (smem > 227 * 1024) return set_error(PC_ERR_INVALID, "decode: batch too large for the in-kernel schedule");
  static size_t smem_set[64] = {0};  // per device: largest dynamic smem opted in so far
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || smem_set[dev] < smem) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(decode)", e);
    if (dev >= 0 && dev < 64) smem_set[dev] = smem;
  }
  const long long pages_max = (long long)p.B * p.GH * max_blocks;
  long long workers = (pages_max + p.min_quantum - 1) / p.min_quantum;
  if (workers < 1) workers = 1;
  const int grid = (int)(workers < num_sms ? workers : num_sms);  // spread over SMs first
  kern<<<grid, T::kThreads, smem, stream>>>(p);
  return check_launch("paged_decode_attn_score_kernel");
}

int paged_decode_attn_score(const DecodeArgs& a, cudaStream_t stream) {
  if (a.B <= 0) return 0;
  if (a.Hq % a.Hkv != 0) return set_error(PC_ERR_INVALID, "num_heads must be a multiple of num_kv_heads");
  if ((a.out_stride % 8) != 0 || (reinterpret_cast<uintptr_t>(a.out) & 15) != 0 ||
      (a.q_stride % 2) != 0 || (reinterpret_cast<uintptr_t>(a.q) & 3) != 0)
    return set_error(PC_ERR_INVALID, "decode: out rows must be 16-byte aligned, query rows 4-byte aligned");
  const int num_sms = device_sm_count();
  if (num_sms <= 0) return PC_ERR_CUDA;
  const size_t need = decode_workspace_bytes(a.B, a.Hq, a.Hkv, a.D, a.BS, a.max_kv_len, num_sms);
  if (a.workspace_bytes < need)
    return set_error(PC_ERR_INVALID, "decode workspace too small (see pc_paged_decode_workspace_bytes)");
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  DecodeParams p;
  p.q = static_cast<const __nv_bfloat16*>(a.q);
  p.k_cache = static_cast<const __nv_bfloat16*>(a.k_cache);
  p.v_cache = static_cast<const __nv_bfloat16*>(a.v_cache);
  p.block_tables = a.block_tables;
  p.kv_lens = a.kv_lens;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.part_ml = reinterpret_cast<float*>(ws);
  p.part_o = reinterpret_cast<float*>(ws + ws_off_part_o(num_sms));
  p.tickets = reinterpret_cast<int*>(ws + ws_off_tickets(num_sms));
  p.seq_lens = a.score_table ? a.seq_lens : nullptr;
  p.score_slots = a.score_slots;
  p.score_table = a.score_table;
  p.score_max_blocks = a.score_max_blocks;
  p.B = a.B; p.Hq = a.Hq; p.Hkv = a.Hkv; p.G = a.Hq / a.Hkv; p.HG = (p.G + 7) / 8; p.GH = a.Hkv * p.HG;
  p.q_stride = a.q_stride; p.out_stride = a.out_stride;
  p.table_stride = a.table_stride; p.kv_block_stride = a.kv_block_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  p.min_quantum = 4;
  p.trace = g_decode_trace;
  const int max_blocks = a.max_kv_len > 0 ? (a.max_kv_len + a.BS - 1) / a.BS : 1;

  if (a.D == 128 && a.BS == 16) return launch_decode<128, 16, 8, 3>(p, num_sms, max_blocks, stream);
  if (a.D == 128 && a.BS == 32) return launch_decode<128, 32, 4, 3>(p, num_sms, max_blocks, stream);
  if (a.D == 64 && a.BS == 16) return launch_decode<64, 16, 8, 6>(p, num_sms, max_blocks, stream);
  if (a.D == 64 && a.BS == 32) return launch_decode<64, 32, 8, 3>(p, num_sms, max_blocks, stream);
  // mirrors TORCH_CHECK(false, "Unsupported head size / block size") in
  // csrc/attention/paged_attention_v1.cu:108-110,160-162
  return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size / block size (supported: head 64|128, block 16|32)");
}

}  // namespace pc
