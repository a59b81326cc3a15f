// Paged decode attention over (compacted) block tables with value_l2 block scoring
// fused into the same KV read.  sm_100a.
//
// Replaces csrc/attention/attention_kernels.cuh:86-668 (paged_attention_v1/v2 + reduce)
// and the extra V read of torch_cuda_sdpa.py:1518-1526 (last-block scoring).
//
// Design (HBM-bound; see DESIGN.md "decode kernel"):
//   * persistent grid: one CTA per SM, kWarps independent warps per CTA; a warp is a
//     worker that pulls (seq, kv-head, head-group, kv-split) items from a global
//     atomic counter.  No __syncthreads in the steady state.
//   * each warp owns a kStages-deep ring of page buffers.  A page = one (block,
//     kv-head) = 2 contiguous chunks in the reference cache layout
//     (K [D/8][BS][8], V [D][BS], vllm/attention/ops/paged_attn.py:49-62), fetched
//     with two TMA 1-D bulk copies (cp.async.bulk -> UBLKCP) that complete on an
//     mbarrier.  >= (kStages-1) * kWarps pages are in flight per SM.
//   * the whole GQA group (<= 8 q heads) is served from ONE read of the page:
//     S^T[16 tok x 8 heads] = K_tile(A, ldmatrix straight from the cache layout)
//     x Q^T(B, registers) with mma.sync m16n8k16 bf16/fp32; online softmax with warp
//     shuffles in the exp2 domain; P^T is transposed in registers (movmatrix) and
//     O^T[D x 8 heads] += V^T_tile(A, ldmatrix straight from the cache layout) x P^T.
//     Tensor cores are used because 2*G FMAs per KV byte exceed the SIMT fp32 budget
//     at HBM speed for G >= 4; the legacy HMMA path is far from saturated (HBM-bound).
//   * numerics follow attention_kernels.cuh: fp32 logits, probabilities rounded to
//     bf16 before P.V, fp32 accumulation, final 1/(sum + 1e-6).
//   * split-KV partials (m, l, O fp32) go to a workspace; the LAST warp to finish a
//     (seq, head, group) combines them in split order (deterministic), so the op is a
//     single launch.
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
  float* part_o;    // [items][8][D]
  float* part_ml;   // [items][8][2]
  int* counters;    // [0] work counter, [1 + group] arrival tickets
  const int* __restrict__ seq_lens;     // nullptr: no scoring
  const int* __restrict__ score_slots;  // [B]
  float* score_table;                   // [slots][score_max_blocks][Hkv]
  int score_max_blocks;
  int B, Hq, Hkv, G, HG;  // G = Hq/Hkv, HG = ceil(G/8)
  long long q_stride;         // elements between sequences in q
  long long out_stride;       // elements between sequences in out
  long long table_stride;     // ints between rows
  long long kv_block_stride;  // elements between blocks in k/v cache
  float scale_log2;
  int chunk_blocks, max_splits, total_items;
};

template <int D, int BS, int kWarps, int kStages>
struct DecodeTraits {
  static constexpr int kPageBytes = BS * D * 2;        // K or V page of one (block, head)
  static constexpr int kStageBytes = 2 * kPageBytes;   // K + V
  static constexpr int kSmemBytes = kWarps * kStages * kStageBytes + kWarps * kStages * 8;
  static constexpr int kThreads = kWarps * 32;
};

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

  uint8_t* my_stages = smem + (size_t)warp * kStages * T::kStageBytes;
  uint64_t* my_bars =
      reinterpret_cast<uint64_t*>(smem + (size_t)kWarps * kStages * T::kStageBytes) + warp * kStages;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&my_bars[s], 1);
  }
  mbar_fence_init();
  __syncthreads();

  uint32_t it = 0;  // pages consumed so far by this warp (ring position / phase)

  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(&p.counters[0], 1);
    item = __shfl_sync(kFull, item, 0);
    if (item >= p.total_items) break;

    // item -> (b, h, hg, split); splits of one group are adjacent so its workers run together
    const int split = item % p.max_splits;
    int rest = item / p.max_splits;
    const int hg = rest % p.HG;
    rest /= p.HG;
    const int h = rest % p.Hkv;
    const int b = rest / p.Hkv;

    const int kv_len = p.kv_lens[b];
    const int nblk = (kv_len + BS - 1) / BS;
    const int blk_begin = split * p.chunk_blocks;
    const int blk_end = min(nblk, blk_begin + p.chunk_blocks);
    const int head0 = h * p.G + hg * 8;              // first q head of this group
    const int nheads = min(8, p.G - hg * 8);         // valid q heads in the group

    if (blk_begin >= blk_end) {
      if (kv_len == 0 && split == 0) {  // empty sequence: defined output (zeros)
        for (int i = lane; i < nheads * D; i += 32)
          p.out[(size_t)b * p.out_stride + (size_t)(head0 + i / D) * D + (i % D)] =
              __float2bfloat16(0.f);
      }
      continue;
    }
    const int nsplit = (nblk + p.chunk_blocks - 1) / p.chunk_blocks;
    const int npages = blk_end - blk_begin;
    const int* table = p.block_tables + (size_t)b * p.table_stride;

    // ---- Q^T as the mma B operand: b0 = Q[head g][16j + 2t..], b1 = Q[head g][16j + 8 + 2t..]
    uint32_t qf[KSTEPS][2];
    {
      const bool valid = g < nheads;
      const __nv_bfloat16* qrow = p.q + (size_t)b * p.q_stride + (size_t)(head0 + (valid ? g : 0)) * D;
#pragma unroll
      for (int j = 0; j < KSTEPS; ++j) {
        qf[j][0] = valid ? *reinterpret_cast<const uint32_t*>(qrow + 16 * j + 2 * t) : 0u;
        qf[j][1] = valid ? *reinterpret_cast<const uint32_t*>(qrow + 16 * j + 8 + 2 * t) : 0u;
      }
    }

    // ---- block-table prefetch registers: entries [grp*32 + lane]
    int tbl_cur = (blk_begin + lane < blk_end) ? table[blk_begin + lane] : 0;
    int tbl_next = (blk_begin + 32 + lane < blk_end) ? table[blk_begin + 32 + lane] : 0;

    auto issue_page = [&](int pi /*page index within chunk*/, int phys_block) {
      // called by all lanes (phys_block already broadcast); lane 0 issues
      const uint32_t slot = (it + (uint32_t)pi) % kStages;  // see ring note below
      if (lane == 0) {
        uint8_t* dst = my_stages + (size_t)slot * T::kStageBytes;
        const size_t off = ((size_t)phys_block * p.kv_block_stride) + (size_t)h * (D * BS);
        mbar_expect_tx(&my_bars[slot], T::kStageBytes);
        tma_load_1d(dst, p.k_cache + off, T::kPageBytes, &my_bars[slot]);
        tma_load_1d(dst + T::kPageBytes, p.v_cache + off, T::kPageBytes, &my_bars[slot]);
      }
    };
    // ring note: `it` is the global count of pages this warp has consumed before this
    // item; page pi of the item lives in slot (it+pi) % kStages with parity
    // ((it+pi) / kStages) & 1.  A slot is refilled only after the page that occupied it
    // was consumed (program order + __syncwarp), so no "empty" barrier is needed.

    // prologue: fill the ring
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      if (s < npages) {
        const int pb = __shfl_sync(kFull, tbl_cur, s);  // kStages <= 32
        issue_page(s, pb);
      }
    }

    float o[MT][4];
#pragma unroll
    for (int m = 0; m < MT; ++m) o[m][0] = o[m][1] = o[m][2] = o[m][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};

    const bool do_score = (p.seq_lens != nullptr) && (hg == 0) && (blk_end == nblk) &&
                          (p.seq_lens[b] % BS == 0) && (kv_len % BS == 0);

    for (int pi = 0; pi < npages; ++pi) {
      const uint32_t pos = it + (uint32_t)pi;
      const uint32_t slot = pos % kStages;
      const uint32_t parity = (pos / kStages) & 1u;
      mbar_wait(&my_bars[slot], parity);

      const uint8_t* kpage = my_stages + (size_t)slot * T::kStageBytes;
      const uint8_t* vpage = kpage + T::kPageBytes;
      const uint32_t kaddr = smem_u32(kpage);
      const uint32_t vaddr = smem_u32(vpage);
      const int page_tok0 = (blk_begin + pi) * BS;
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
      if (do_score && (blk_begin + pi == nblk - 1)) {
        const __nv_bfloat16* vp = reinterpret_cast<const __nv_bfloat16*>(vpage);
        const float sc = warp_value_l2_head_score<D, BS>(
            [&](int tok, int d) { return __bfloat162float(vp[d * BS + tok]); });
        if (lane == 0)
          p.score_table[((size_t)p.score_slots[b] * p.score_max_blocks + (nblk - 1)) * p.Hkv + h] = sc;
      }

      // ---------------- refill this slot with page pi + kStages
      __syncwarp();
      const int nxt = pi + kStages;
      if (nxt < npages) {
        // table entry for chunk page `nxt`: group nxt/32 is either the current or next register
        if ((nxt & 31) == 0) {  // crossing into a new group of 32 entries
          tbl_cur = tbl_next;
          const int idx = blk_begin + nxt + 32 + lane;
          tbl_next = (idx < blk_end) ? table[idx] : 0;
        }
        const int pb = __shfl_sync(kFull, tbl_cur, nxt & 31);
        // slot of page nxt == slot of page pi
        if (lane == 0) {
          uint8_t* dst = my_stages + (size_t)slot * T::kStageBytes;
          const size_t off = ((size_t)pb * p.kv_block_stride) + (size_t)h * (D * BS);
          mbar_expect_tx(&my_bars[slot], T::kStageBytes);
          tma_load_1d(dst, p.k_cache + off, T::kPageBytes, &my_bars[slot]);
          tma_load_1d(dst + T::kPageBytes, p.v_cache + off, T::kPageBytes, &my_bars[slot]);
        }
      }
    }
    it += (uint32_t)npages;

    // ---------------- finish: l over the 8 row-lanes of each column
#pragma unroll
    for (int sh = 4; sh <= 16; sh <<= 1) {
      l_run[0] += __shfl_xor_sync(kFull, l_run[0], sh);
      l_run[1] += __shfl_xor_sync(kFull, l_run[1], sh);
    }
    // o[m][0]=(dim 16m+g, head 2t) [1]=(dim 16m+g, head 2t+1) [2]=(dim 16m+g+8, 2t) [3]=(.., 2t+1)
    if (nsplit == 1) {
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
      continue;
    }

    // ---------------- split-KV: publish the partial, last arriver combines
    {
      float* po = p.part_o + (size_t)item * 8 * D;
      float* pml = p.part_ml + (size_t)item * 16;
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
    __threadfence();
    __syncwarp();
    const int group = (b * p.Hkv + h) * p.HG + hg;
    int ticket = 0;
    if (lane == 0) ticket = atomicAdd(&p.counters[1 + group], 1);
    ticket = __shfl_sync(kFull, ticket, 0);
    if (ticket != nsplit - 1) continue;
    __threadfence();

    const int item0 = item - split;  // split 0 of this group
    for (int hh = 0; hh < nheads; ++hh) {
      float M = -INFINITY;
      for (int s = 0; s < nsplit; ++s)
        M = fmaxf(M, __ldcg(p.part_ml + (size_t)(item0 + s) * 16 + hh * 2));
      float L = 0.f;
      float acc[D / 32];
#pragma unroll
      for (int i = 0; i < D / 32; ++i) acc[i] = 0.f;
      for (int s = 0; s < nsplit; ++s) {
        const float ms = __ldcg(p.part_ml + (size_t)(item0 + s) * 16 + hh * 2);
        const float ls = __ldcg(p.part_ml + (size_t)(item0 + s) * 16 + hh * 2 + 1);
        const float w = fast_exp2(ms - M);
        L += w * ls;
        const float* src = p.part_o + ((size_t)(item0 + s) * 8 + hh) * D;
#pragma unroll
        for (int i = 0; i < D / 32; ++i) acc[i] += w * __ldcg(src + i * 32 + lane);
      }
      const float inv = 1.f / (L + 1e-6f);
      __nv_bfloat16* ob = p.out + (size_t)b * p.out_stride + (size_t)(head0 + hh) * D;
#pragma unroll
      for (int i = 0; i < D / 32; ++i) ob[i * 32 + lane] = __float2bfloat16_rn(acc[i] * inv);
    }
  }
}

// ---------------------------------------------------------------------------- host
template <int D, int BS, int kWarps, int kStages>
static int launch_decode(const DecodeParams& p, int num_sms, cudaStream_t stream) {
  using T = DecodeTraits<D, BS, kWarps, kStages>;
  auto kern = paged_decode_attn_score_kernel<D, BS, kWarps, kStages>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, T::kSmemBytes);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(decode)", e);
  const int workers_needed = (p.total_items + kWarps - 1) / kWarps;
  const int grid = max(1, min(num_sms, workers_needed));
  kern<<<grid, T::kThreads, T::kSmemBytes, stream>>>(p);
  return check_launch("paged_decode_attn_score_kernel");
}

struct DecodePlan {
  int chunk_blocks, max_splits, HG, total_items, groups;
  size_t off_part_o, off_part_ml, off_counters, bytes;
};

static DecodePlan plan_decode(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms,
                              int warps) {
  DecodePlan pl;
  const int G = Hq / Hkv;
  pl.HG = (G + 7) / 8;
  const int max_blocks = max(1, (max_kv_len + BS - 1) / BS);
  const long long groups = (long long)B * Hkv * pl.HG;
  const long long total_pages = groups * max_blocks;
  const long long workers = (long long)num_sms * warps;
  // aim at ~4 items per worker for balance; never below 4 pages per item
  long long chunk = (total_pages + workers * 4 - 1) / (workers * 4);
  if (chunk < 4) chunk = 4;
  if (chunk > max_blocks) chunk = max_blocks;
  pl.chunk_blocks = (int)chunk;
  pl.max_splits = (max_blocks + pl.chunk_blocks - 1) / pl.chunk_blocks;
  pl.groups = (int)groups;
  pl.total_items = (int)(groups * pl.max_splits);
  size_t off = 0;
  pl.off_counters = off;
  off += ((size_t)(1 + groups) * sizeof(int) + 255) / 256 * 256;
  pl.off_part_ml = off;
  off += ((size_t)pl.total_items * 16 * sizeof(float) + 255) / 256 * 256;
  pl.off_part_o = off;
  off += (size_t)pl.total_items * 8 * D * sizeof(float);
  pl.bytes = off;
  return pl;
}

// warps per CTA for each (D, BS): the ring (warps x stages x (K+V page)) must fit 227 KB
static int decode_warps(int D, int BS) { return (D == 128 && BS == 32) ? 4 : 8; }

size_t decode_workspace_bytes(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms) {
  return plan_decode(B, Hq, Hkv, D, BS, max_kv_len, num_sms, decode_warps(D, BS)).bytes;
}

int paged_decode_attn_score(const DecodeArgs& a, cudaStream_t stream) {
  if (a.B <= 0) return 0;
  if (a.Hq % a.Hkv != 0) return set_error(PC_ERR_INVALID, "num_heads must be a multiple of num_kv_heads");
  const int num_sms = device_sm_count();
  if (num_sms <= 0) return PC_ERR_CUDA;
  const DecodePlan pl = plan_decode(a.B, a.Hq, a.Hkv, a.D, a.BS, a.max_kv_len, num_sms, decode_warps(a.D, a.BS));
  if (a.workspace_bytes < pl.bytes)
    return set_error(PC_ERR_INVALID, "decode workspace too small (see pc_paged_decode_workspace_bytes)");
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  DecodeParams p;
  p.q = static_cast<const __nv_bfloat16*>(a.q);
  p.k_cache = static_cast<const __nv_bfloat16*>(a.k_cache);
  p.v_cache = static_cast<const __nv_bfloat16*>(a.v_cache);
  p.block_tables = a.block_tables;
  p.kv_lens = a.kv_lens;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.part_o = reinterpret_cast<float*>(ws + pl.off_part_o);
  p.part_ml = reinterpret_cast<float*>(ws + pl.off_part_ml);
  p.counters = reinterpret_cast<int*>(ws + pl.off_counters);
  p.seq_lens = a.score_table ? a.seq_lens : nullptr;
  p.score_slots = a.score_slots;
  p.score_table = a.score_table;
  p.score_max_blocks = a.score_max_blocks;
  p.B = a.B; p.Hq = a.Hq; p.Hkv = a.Hkv; p.G = a.Hq / a.Hkv; p.HG = pl.HG;
  p.q_stride = a.q_stride; p.out_stride = a.out_stride;
  p.table_stride = a.table_stride; p.kv_block_stride = a.kv_block_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  p.chunk_blocks = pl.chunk_blocks; p.max_splits = pl.max_splits; p.total_items = pl.total_items;

  cudaError_t e = cudaMemsetAsync(p.counters, 0, (size_t)(1 + pl.groups) * sizeof(int), stream);
  if (e != cudaSuccess) return set_cuda_error("cudaMemsetAsync(decode counters)", e);

  if (a.D == 128 && a.BS == 16) return launch_decode<128, 16, 8, 3>(p, num_sms, stream);
  if (a.D == 128 && a.BS == 32) return launch_decode<128, 32, 4, 3>(p, num_sms, stream);
  if (a.D == 64 && a.BS == 16) return launch_decode<64, 16, 8, 6>(p, num_sms, stream);
  if (a.D == 64 && a.BS == 32) return launch_decode<64, 32, 8, 3>(p, num_sms, stream);
  // mirrors TORCH_CHECK(false, "Unsupported head size / block size") in
  // csrc/attention/paged_attention_v1.cu:108-110,160-162
  return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size / block size (supported: head 64|128, block 16|32)");
}

}  // namespace pc
