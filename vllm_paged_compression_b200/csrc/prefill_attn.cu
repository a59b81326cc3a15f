// Prefill attention over the (streamingLLM-)pruned prompt with torch's TOP-LEFT aligned
// causal mask: query i of a request sees kept-kv j <= i   (Lq = prompt length,
// Lkv = kept length <= Lq).
//
// Replaces _run_sdpa_forward (vllm/attention/backends/torch_cuda_sdpa.py:1069-1127):
// per-request F.scaled_dot_product_attention(q[1,Hq,Lq,D], k[1,Hq,Lkv,D], v, is_causal=True)
// after repeat_interleave of K/V to Hq heads (:1078-1080).  Here GQA is index
// arithmetic (kv head = q head / G) and the pruned K/V is never materialised: rows are
// gathered on the fly with src = j < first ? j : j + pruned (kvcache_prunner.py:203-215).
//
// The product kernel is the tcgen05/TMEM one in prefill_attn_tc.cu; this file keeps the
// argument checks, the dispatcher, and the first implementation as the measured
// "recompiled mma.sync" baseline (pc_debug_set_prefill_variant(1)):
// flash-attention forward on mma.sync m16n8k16 (bf16 in, fp32
// accumulate, P rounded to bf16 before P.V like torch's flash/mem-efficient kernels),
// 64-query x 64-key tiles, 4 warps, cp.async double-buffered K/V tiles in XOR-swizzled
// shared memory.
#include "pc_common.cuh"
#include "pc_api_internal.h"

namespace pc {

struct PrefillParams {
  const __nv_bfloat16* __restrict__ q;
  const __nv_bfloat16* __restrict__ k;
  const __nv_bfloat16* __restrict__ v;
  __nv_bfloat16* __restrict__ out;
  const int* __restrict__ req_meta;  // [R][6] {q_start, q_len, kv_src_start, kv_len, first, pruned}
  int Hq, Hkv;
  long long q_stride, out_stride, key_stride, value_stride;
  float scale_log2;
};

constexpr int kBM = 64;  // query rows per CTA
constexpr int kBN = 64;  // kv rows per tile

// smem tile [rows][D] bf16, 16-byte chunks XOR-swizzled by (row & 7)
template <int D>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
  return (uint32_t)((row * (D / 8) + (chunk ^ (row & 7))) * 16);
}

template <int D>
__global__ void __launch_bounds__(128) prefill_attn_kernel(const PrefillParams p) {
  constexpr int CH = D / 8;      // 16-byte chunks per row
  constexpr int KS = D / 16;     // k-steps over head dim
  constexpr int NT = kBN / 8;    // n-tiles of S per warp row-block
  constexpr int OT = D / 8;      // n-tiles of O
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sQ = smem;                       // kBM * D * 2
  uint8_t* sK = sQ + kBM * D * 2;           // 2 stages
  uint8_t* sV = sK + 2 * kBN * D * 2;       // 2 stages

  const int r = blockIdx.z;
  const int hq = blockIdx.y;
  const int* m = p.req_meta + r * 6;
  const int q_start = m[0], q_len = m[1], kv_src = m[2], kv_len = m[3], first = m[4], pruned = m[5];
  const int i0 = blockIdx.x * kBM;
  if (i0 >= q_len) return;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;

  // number of kv tiles this q tile can see: j <= i, i < i0 + kBM
  const int kv_hi = min(kv_len, i0 + kBM);
  const int ntiles = (kv_hi + kBN - 1) / kBN;

  auto load_kv_tile = [&](int tile, int stage) {
    const int j0 = tile * kBN;
    uint8_t* dK = sK + stage * kBN * D * 2;
    uint8_t* dV = sV + stage * kBN * D * 2;
    for (int i = tid; i < kBN * CH; i += 128) {
      const int row = i / CH, c = i % CH;
      const int j = j0 + row;
      const bool ok = j < kv_len;
      const long long src = (long long)kv_src + (ok ? (j < first ? j : j + pruned) : 0);
      cp_async_16_zfill(dK + swz<D>(row, c), p.k + (size_t)src * p.key_stride + (size_t)hkv * D + c * 8, ok);
      cp_async_16_zfill(dV + swz<D>(row, c), p.v + (size_t)src * p.value_stride + (size_t)hkv * D + c * 8, ok);
    }
  };

  // ---- Q tile -> smem (zero rows past q_len), first K/V tile
  for (int i = tid; i < kBM * CH; i += 128) {
    const int row = i / CH, c = i % CH;
    const bool ok = (i0 + row) < q_len;
    const size_t off = (size_t)(q_start + (ok ? i0 + row : 0)) * p.q_stride + (size_t)hq * D + c * 8;
    cp_async_16_zfill(sQ + swz<D>(row, c), p.q + off, ok);
  }
  load_kv_tile(0, 0);
  cp_async_commit();

  float o[OT][4];
#pragma unroll
  for (int n = 0; n < OT; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  uint32_t qf[KS][4];

  for (int tile = 0; tile < ntiles; ++tile) {
    const int stage = tile & 1;
    if (tile + 1 < ntiles) load_kv_tile(tile + 1, stage ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();

    if (tile == 0) {  // Q fragments (A operand), once
      const int mi = lane >> 3, rr = lane & 7;
      const int row = warp * 16 + (mi & 1) * 8 + rr;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
        ldmatrix_x4(qf[ks], smem_u32(sQ) + swz<D>(row, 2 * ks + (mi >> 1)));
    }

    const uint32_t kbase = smem_u32(sK + stage * kBN * D * 2);
    const uint32_t vbase = smem_u32(sV + stage * kBN * D * 2);

    // ---- S = Q K^T  (16 x 64 per warp)
    float s[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int np = 0; np < NT / 2; ++np) {
        // matrices: (tokens 16np..+7, chunk 2ks), (same tokens, chunk 2ks+1),
        //           (tokens 16np+8..+15, chunk 2ks), (.., chunk 2ks+1)
        const int mi = lane >> 3, rr = lane & 7;
        const int row = np * 16 + (mi >> 1) * 8 + rr;
        uint32_t bfr[4];
        ldmatrix_x4(bfr, kbase + swz<D>(row, 2 * ks + (mi & 1)));
        mma_bf16_16816(s[2 * np], qf[ks], bfr[0], bfr[1]);
        mma_bf16_16816(s[2 * np + 1], qf[ks], bfr[2], bfr[3]);
      }
    }

    // ---- scale, mask (top-left causal + kv tail), online softmax
    const int j0 = tile * kBN;
    const int row0 = i0 + warp * 16 + g;  // query index of c0/c1; c2/c3 are row0 + 8
    const bool need_mask = (j0 + kBN - 1 > i0 + warp * 16) || (j0 + kBN > kv_len);
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
#pragma unroll
      for (int e = 0; e < 4; ++e) s[n][e] *= p.scale_log2;
      if (need_mask) {
        const int j = j0 + n * 8 + 2 * t;
        if (j > row0 || j >= kv_len) s[n][0] = -INFINITY;
        if (j + 1 > row0 || j + 1 >= kv_len) s[n][1] = -INFINITY;
        if (j > row0 + 8 || j >= kv_len) s[n][2] = -INFINITY;
        if (j + 1 > row0 + 8 || j + 1 >= kv_len) s[n][3] = -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(s[n][0], s[n][1]));
      mx1 = fmaxf(mx1, fmaxf(s[n][2], s[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(kFull, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(kFull, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(kFull, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(kFull, mx1, 2));
    const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
    // a fully masked row-tile keeps mn == -inf only before any visible key; guard the NaN
    const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
    const float al0 = fast_exp2(m_run[0] - ms0), al1 = fast_exp2(m_run[1] - ms1);
    m_run[0] = mn0; m_run[1] = mn1;
    float sum0 = 0.f, sum1 = 0.f;
    uint32_t pf[NT][2];
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      const float p0 = fast_exp2(s[n][0] - ms0), p1 = fast_exp2(s[n][1] - ms0);
      const float p2 = fast_exp2(s[n][2] - ms1), p3 = fast_exp2(s[n][3] - ms1);
      sum0 += p0 + p1; sum1 += p2 + p3;
      pf[n][0] = pack_bf16x2(p0, p1);
      pf[n][1] = pack_bf16x2(p2, p3);
    }
    l_run[0] = l_run[0] * al0 + sum0;
    l_run[1] = l_run[1] * al1 + sum1;
#pragma unroll
    for (int n = 0; n < OT; ++n) { o[n][0] *= al0; o[n][1] *= al0; o[n][2] *= al1; o[n][3] *= al1; }

    // ---- O += P V   (A = P from registers, B = V via ldmatrix.trans)
#pragma unroll
    for (int kk = 0; kk < kBN / 16; ++kk) {
      const uint32_t a[4] = {pf[2 * kk][0], pf[2 * kk][1], pf[2 * kk + 1][0], pf[2 * kk + 1][1]};
#pragma unroll
      for (int np = 0; np < OT / 2; ++np) {
        // matrices: (tok 16kk..+7, dims 16np..+7), (tok +8.., same dims),
        //           (tok 16kk..+7, dims 16np+8..), (tok +8.., dims 16np+8..)
        const int mi = lane >> 3, rr = lane & 7;
        const int row = kk * 16 + (mi & 1) * 8 + rr;
        uint32_t bfr[4];
        ldmatrix_x4_trans(bfr, vbase + swz<D>(row, 2 * np + (mi >> 1)));
        mma_bf16_16816(o[2 * np], a, bfr[0], bfr[1]);
        mma_bf16_16816(o[2 * np + 1], a, bfr[2], bfr[3]);
      }
    }
    __syncthreads();  // everyone done with this stage before it is refilled
  }
  cp_async_wait<0>();

  // ---- epilogue: rows g / g+8 of this warp, quad-reduced l
  l_run[0] += __shfl_xor_sync(kFull, l_run[0], 1);
  l_run[0] += __shfl_xor_sync(kFull, l_run[0], 2);
  l_run[1] += __shfl_xor_sync(kFull, l_run[1], 1);
  l_run[1] += __shfl_xor_sync(kFull, l_run[1], 2);
  const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
  const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
  const int ra = i0 + warp * 16 + g, rb = ra + 8;
#pragma unroll
  for (int n = 0; n < OT; ++n) {
    if (ra < q_len) {
      __nv_bfloat16* dst = p.out + (size_t)(q_start + ra) * p.out_stride + (size_t)hq * D + n * 8 + 2 * t;
      *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(o[n][0] * inv0, o[n][1] * inv0);
    }
    if (rb < q_len) {
      __nv_bfloat16* dst = p.out + (size_t)(q_start + rb) * p.out_stride + (size_t)hq * D + n * 8 + 2 * t;
      *reinterpret_cast<uint32_t*>(dst) = pack_bf16x2(o[n][2] * inv1, o[n][3] * inv1);
    }
  }
}

template <int D>
static int launch_prefill(const PrefillParams& p, const PrefillArgs& a, cudaStream_t stream) {
  const int smem = (kBM * D + 4 * kBN * D) * 2;
  auto kern = prefill_attn_kernel<D>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(prefill)", e);
  dim3 grid((a.max_q_len + kBM - 1) / kBM, a.Hq, a.num_reqs);
  kern<<<grid, 128, smem, stream>>>(p);
  return check_launch("prefill_attn_kernel");
}

static int g_prefill_variant = 0;
void set_prefill_variant(int v) { g_prefill_variant = v; }

int prefill_attn(const PrefillArgs& a, cudaStream_t stream) {
  if (a.num_reqs <= 0 || a.max_q_len <= 0) return 0;
  if (a.Hq % a.Hkv) return set_error(PC_ERR_INVALID, "num_heads must be a multiple of num_kv_heads");
  if ((a.q_stride | a.out_stride | a.key_stride | a.value_stride) % 8)
    return set_error(PC_ERR_INVALID, "q/out/key/value row strides must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) |
       reinterpret_cast<uintptr_t>(a.v) | reinterpret_cast<uintptr_t>(a.out)) & 15)
    return set_error(PC_ERR_INVALID, "q/k/v/out pointers must be 16-byte aligned");
  if (g_prefill_variant == 0) return prefill_attn_tc(a, stream);
  PrefillParams p;
  p.q = static_cast<const __nv_bfloat16*>(a.q);
  p.k = static_cast<const __nv_bfloat16*>(a.k);
  p.v = static_cast<const __nv_bfloat16*>(a.v);
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.req_meta = a.req_meta;
  p.Hq = a.Hq; p.Hkv = a.Hkv;
  p.q_stride = a.q_stride; p.out_stride = a.out_stride;
  p.key_stride = a.key_stride; p.value_stride = a.value_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  if (a.D == 128) return launch_prefill<128>(p, a, stream);
  if (a.D == 64) return launch_prefill<64>(p, a, stream);
  return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size (supported: 64|128)");
}

}  // namespace pc
