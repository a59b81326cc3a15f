// Token-level prompt pruning of HEAD's `local` / `global` / `inverse_key_l2` methods.
//
// Replaces the second branch of KVCachePruner.prune_prompt_inplace
// (vllm/attention/kvcache_prunner.py:220-279) driven by _prune_prompts_keys_and_values_new
// (vllm/attention/backends/torch_cuda_sdpa.py:1222-1296):
//   first slice  = tokens [0, bs)                           kept
//   middle slice = tokens [bs, q_len - bs - q_len % bs)     per kv HEAD the k lowest-scoring tokens
//                  are dropped (scores [T_mid, H]; torch.topk(largest=False, dim=0)), where
//                  k = T_mid - (budget - 2 bs)
//   last slice   = the rest                                 kept
// score = ||V[t,h,:]||_2 (local / global) or 1 / (||K[t,h,:]||_2 + 1e-8) (inverse_key_l2).
// The reference then re-packs the surviving (token, head) vectors with `middle[mask].view(-1, H, D)`:
// boolean-mask order is token-major, head-minor, so output row j holds the survivors number
// j*H .. j*H + H-1 of that order -- vectors of DIFFERENT source heads end up in one head column
// (SURVEY.md 0.5).  repack = 1 reproduces exactly that (the drop-in default); repack = 0 keeps
// every head's survivors in its own column (row j of head h = its j-th surviving token).
//
// Declared deviation (as for the block scores): scores are fp32 with the fixed
// reduction order of csrc/score_common.cuh (the reference reduces in bf16 with torch's order), and
// ties between equal scores go to the lowest token index (torch.topk leaves them unspecified).
//
// Byte work: K and V of the prompt read once for the scores (one of them) and once for the copy,
// written once.  Four small launches; nothing here is on the decode path.
#include "pc_common.cuh"
#include "pc_api_internal.h"

namespace pc {

struct TokenPruneParams {
  const __nv_bfloat16* __restrict__ key;
  const __nv_bfloat16* __restrict__ value;
  __nv_bfloat16* out_key;
  __nv_bfloat16* out_value;
  const int* __restrict__ req_meta;   // [R][6] {src_start, q_len, first, mid, k_drop, dst_start}
  const int* __restrict__ mid_offsets;  // [R+1] prefix sum of mid (rows of the workspace arrays)
  const int* __restrict__ tok_offsets;  // [R+1] prefix sum of q_len
  int num_reqs, Hkv, D, score_kind, repack;
  long long key_stride, value_stride, out_key_stride, out_value_stride;
  float* scores;            // [sum_mid][Hkv]
  unsigned char* keep;      // [sum_mid][Hkv]
  int* base;                // [sum_mid][Hkv or 1] exclusive prefix (see repack)
};

// ---- K1: scores.  One warp = 4 tokens of one head (8 lanes per token, score_common.cuh order).
template <int D>
__global__ void __launch_bounds__(256) token_score_kernel(const TokenPruneParams p, int total_mid) {
  const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, j = lane & 7, tq = lane >> 3;
  const int groups = (total_mid + 3) / 4;
  if (warp_g >= groups * p.Hkv) return;
  const int h = warp_g % p.Hkv;
  const int row = (warp_g / p.Hkv) * 4 + tq;     // row in the concatenated middle slices
  float acc = 0.f;
  const bool live = row < total_mid;
  if (live) {
    int lo = 0, hi = p.num_reqs;                 // request of this row
    while (hi - lo > 1) {
      const int m = (lo + hi) >> 1;
      if (p.mid_offsets[m] <= row) lo = m; else hi = m;
    }
    const int* m = p.req_meta + lo * 6;
    const size_t src = (size_t)m[0] + m[2] + (row - p.mid_offsets[lo]);
    const __nv_bfloat16* x = p.score_kind == 0 ? p.value + src * p.value_stride : p.key + src * p.key_stride;
    x += (size_t)h * D;
#pragma unroll 4
    for (int i = 0; i < D / 8; ++i) {
      const float v = __bfloat162float(x[8 * i + j]);
      acc = __fmaf_rn(v, v, acc);
    }
  }
  acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 1));
  acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 2));
  acc = __fadd_rn(acc, __shfl_xor_sync(kFull, acc, 4));
  float sc = __fsqrt_rn(acc);
  if (p.score_kind == 1) sc = __fdiv_rn(1.f, __fadd_rn(sc, 1e-8f));
  if (live && j == 0) p.scores[(size_t)row * p.Hkv + h] = sc;
}

// ---- K2: per (request, head) the k_drop smallest (score, token index) pairs are dropped.
// Scores are non-negative floats, so their bit patterns order like the values: 4-pass radix select
// of the k-th smallest key, then one ordered pass that also breaks ties by token index.
constexpr int kSelThreads = 256;
__global__ void __launch_bounds__(kSelThreads) token_select_kernel(const TokenPruneParams p) {
  __shared__ int hist[256];
  __shared__ int sh_digit, sh_remaining;
  __shared__ int wsum[kSelThreads / 32];
  const int r = blockIdx.x, h = blockIdx.y, tid = threadIdx.x;
  const int* m = p.req_meta + r * 6;
  const int mid = m[3], k_drop = m[4];
  if (mid <= 0) return;
  const size_t row0 = p.mid_offsets[r];
  const unsigned int* keys = reinterpret_cast<const unsigned int*>(p.scores);
  unsigned char* keep = p.keep;
  if (k_drop <= 0) {
    for (int t = tid; t < mid; t += kSelThreads) keep[(row0 + t) * p.Hkv + h] = 1;
    return;
  }
  unsigned int prefix = 0u, mask = 0u;
  int remaining = k_drop;   // 1-based rank of the wanted key among those matching the prefix
  for (int shift = 24; shift >= 0; shift -= 8) {
    hist[tid] = 0;
    __syncthreads();
    for (int t = tid; t < mid; t += kSelThreads) {
      const unsigned int key = keys[(row0 + t) * p.Hkv + h];
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1);
    }
    __syncthreads();
    if (tid == 0) {
      int cum = 0, d = 0;
      for (; d < 256; ++d) {
        if (cum + hist[d] >= remaining) break;
        cum += hist[d];
      }
      sh_digit = d;
      sh_remaining = remaining - cum;
    }
    __syncthreads();
    prefix |= (unsigned int)sh_digit << shift;
    mask |= 255u << shift;
    remaining = sh_remaining;
    __syncthreads();
  }
  // prefix = key of the k-th smallest; `remaining` of the tokens holding exactly that key go
  // (lowest token index first), every token with a smaller key goes
  int carry = 0;
  for (int base = 0; base < mid; base += kSelThreads) {
    const int t = base + tid;
    unsigned int key = 0xffffffffu;
    if (t < mid) key = keys[(row0 + t) * p.Hkv + h];
    const int eq = (t < mid && key == prefix) ? 1 : 0;
    int incl = eq;
#pragma unroll
    for (int sh = 1; sh < 32; sh <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, sh);
      if ((tid & 31) >= sh) incl += v;
    }
    if ((tid & 31) == 31) wsum[tid >> 5] = incl;
    __syncthreads();
    int off = carry, tot = 0;
#pragma unroll
    for (int i = 0; i < kSelThreads / 32; ++i) {
      const int v = wsum[i];
      off += (i < (tid >> 5)) ? v : 0;
      tot += v;
    }
    if (t < mid) {
      const int rank_eq = off + incl - eq;   // equal keys at lower token indices
      const bool drop = key < prefix || (eq && rank_eq < remaining);
      keep[(row0 + t) * p.Hkv + h] = drop ? 0 : 1;
    }
    carry += tot;
    __syncthreads();
  }
}

// ---- K3: exclusive prefix sums over the middle tokens of one request.
//   repack = 1: base[t] = survivors (token, head) pairs before token t (all heads)
//   repack = 0: base[t][h] = survivors of head h before token t
__global__ void __launch_bounds__(kSelThreads) token_scan_kernel(const TokenPruneParams p) {
  __shared__ int wsum[kSelThreads / 32];
  const int r = blockIdx.x, tid = threadIdx.x;
  const int hsel = p.repack ? -1 : (int)blockIdx.y;
  const int* m = p.req_meta + r * 6;
  const int mid = m[3];
  const size_t row0 = p.mid_offsets[r];
  int carry = 0;
  for (int base = 0; base < mid; base += kSelThreads) {
    const int t = base + tid;
    int cnt = 0;
    if (t < mid) {
      const unsigned char* kp = p.keep + (row0 + t) * p.Hkv;
      if (hsel >= 0) cnt = kp[hsel];
      else for (int h = 0; h < p.Hkv; ++h) cnt += kp[h];
    }
    int incl = cnt;
#pragma unroll
    for (int sh = 1; sh < 32; sh <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, sh);
      if ((tid & 31) >= sh) incl += v;
    }
    if ((tid & 31) == 31) wsum[tid >> 5] = incl;
    __syncthreads();
    int off = carry, tot = 0;
#pragma unroll
    for (int i = 0; i < kSelThreads / 32; ++i) {
      const int v = wsum[i];
      off += (i < (tid >> 5)) ? v : 0;
      tot += v;
    }
    if (t < mid) {
      if (hsel >= 0) p.base[(row0 + t) * p.Hkv + hsel] = off + incl - cnt;
      else p.base[row0 + t] = off + incl - cnt;
    }
    carry += tot;
    __syncthreads();
  }
}

// ---- K4: the copy.  One CTA per source token, one warp per (token, head) vector (D bf16).
template <int D>
__global__ void __launch_bounds__(256) token_gather_kernel(const TokenPruneParams p, int total_tokens) {
  const int tok_g = blockIdx.x;
  if (tok_g >= total_tokens) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int lo = 0, hi = p.num_reqs;
  while (hi - lo > 1) {
    const int mm = (lo + hi) >> 1;
    if (p.tok_offsets[mm] <= tok_g) lo = mm; else hi = mm;
  }
  const int* m = p.req_meta + lo * 6;
  const int src_start = m[0], q_len = m[1], first = m[2], mid = m[3], k_drop = m[4], dst_start = m[5];
  const int t = tok_g - p.tok_offsets[lo];
  if (t >= q_len) return;
  const size_t src = (size_t)src_start + t;
  const bool in_mid = t >= first && t < first + mid;
  const size_t mrow = (size_t)p.mid_offsets[lo] + (t - first);
  constexpr int VPL = D / 32;   // elements per lane
  for (int h = warp; h < p.Hkv; h += 8) {
    long long drow;
    int dh = h;
    if (!in_mid) {
      drow = (long long)dst_start + (t < first ? t : t - k_drop);
    } else {
      const unsigned char* kp = p.keep + mrow * p.Hkv;
      if (!kp[h]) continue;
      if (p.repack) {
        int flat = p.base[mrow];
        for (int hh = 0; hh < h; ++hh) flat += kp[hh];
        drow = (long long)dst_start + first + flat / p.Hkv;
        dh = flat % p.Hkv;
      } else {
        drow = (long long)dst_start + first + p.base[mrow * p.Hkv + h];
      }
    }
    const __nv_bfloat16* ks = p.key + src * p.key_stride + (size_t)h * D + lane * VPL;
    const __nv_bfloat16* vs = p.value + src * p.value_stride + (size_t)h * D + lane * VPL;
    __nv_bfloat16* kd = p.out_key + (size_t)drow * p.out_key_stride + (size_t)dh * D + lane * VPL;
    __nv_bfloat16* vd = p.out_value + (size_t)drow * p.out_value_stride + (size_t)dh * D + lane * VPL;
    if constexpr (VPL == 4) {
      *reinterpret_cast<uint2*>(kd) = __ldg(reinterpret_cast<const uint2*>(ks));
      *reinterpret_cast<uint2*>(vd) = __ldg(reinterpret_cast<const uint2*>(vs));
    } else {
      *reinterpret_cast<unsigned int*>(kd) = __ldg(reinterpret_cast<const unsigned int*>(ks));
      *reinterpret_cast<unsigned int*>(vd) = __ldg(reinterpret_cast<const unsigned int*>(vs));
    }
  }
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

size_t token_prune_workspace_bytes(long long total_mid, int Hkv) {
  if (total_mid < 0 || Hkv <= 0) return 0;
  const size_t n = (size_t)total_mid * Hkv;
  return align_up(n * 4, 256) + align_up(n, 256) + align_up(n * 4, 256) + 256;
}

int prompt_token_prune(const TokenPruneArgs& a, cudaStream_t stream) {
  if (a.num_reqs <= 0 || a.total_tokens <= 0) return 0;
  if (a.D != 64 && a.D != 128)
    return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size (supported: 64|128)");
  if ((a.key_stride | a.value_stride | a.out_key_stride | a.out_value_stride) % 4)
    return set_error(PC_ERR_INVALID, "prompt_token_prune: row strides must be multiples of 4 elements");
  if ((reinterpret_cast<uintptr_t>(a.key) | reinterpret_cast<uintptr_t>(a.value) |
       reinterpret_cast<uintptr_t>(a.out_key) | reinterpret_cast<uintptr_t>(a.out_value)) & 7)
    return set_error(PC_ERR_INVALID, "prompt_token_prune: tensors must be 8-byte aligned");
  if (a.workspace_bytes < token_prune_workspace_bytes(a.total_mid, a.Hkv) ||
      (reinterpret_cast<uintptr_t>(a.workspace) & 255))
    return set_error(PC_ERR_INVALID, "prompt_token_prune: workspace too small or not 256-byte aligned");
  const size_t n = (size_t)a.total_mid * a.Hkv;
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  TokenPruneParams p;
  p.key = static_cast<const __nv_bfloat16*>(a.key); p.value = static_cast<const __nv_bfloat16*>(a.value);
  p.out_key = static_cast<__nv_bfloat16*>(a.out_key); p.out_value = static_cast<__nv_bfloat16*>(a.out_value);
  p.req_meta = a.req_meta; p.mid_offsets = a.mid_offsets; p.tok_offsets = a.tok_offsets;
  p.num_reqs = a.num_reqs; p.Hkv = a.Hkv; p.D = a.D; p.score_kind = a.score_kind; p.repack = a.repack ? 1 : 0;
  p.key_stride = a.key_stride; p.value_stride = a.value_stride;
  p.out_key_stride = a.out_key_stride; p.out_value_stride = a.out_value_stride;
  p.scores = reinterpret_cast<float*>(ws);
  p.keep = ws + align_up(n * 4, 256);
  p.base = reinterpret_cast<int*>(ws + align_up(n * 4, 256) + align_up(n, 256));
  if (a.total_mid > 0) {
    const long long warps = (a.total_mid + 3) / 4 * (long long)a.Hkv;
    const unsigned blocks = (unsigned)((warps + 7) / 8);
    if (a.D == 128) token_score_kernel<128><<<blocks, 256, 0, stream>>>(p, (int)a.total_mid);
    else token_score_kernel<64><<<blocks, 256, 0, stream>>>(p, (int)a.total_mid);
    if (int rc = check_launch("token_score_kernel")) return rc;
    token_select_kernel<<<dim3(a.num_reqs, a.Hkv), kSelThreads, 0, stream>>>(p);
    if (int rc = check_launch("token_select_kernel")) return rc;
    token_scan_kernel<<<dim3(a.num_reqs, p.repack ? 1 : a.Hkv), kSelThreads, 0, stream>>>(p);
    if (int rc = check_launch("token_scan_kernel")) return rc;
  }
  if (a.D == 128) token_gather_kernel<128><<<(unsigned)a.total_tokens, 256, 0, stream>>>(p, (int)a.total_tokens);
  else token_gather_kernel<64><<<(unsigned)a.total_tokens, 256, 0, stream>>>(p, (int)a.total_tokens);
  return check_launch("token_gather_kernel");
}

}  // namespace pc
