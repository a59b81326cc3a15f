// reshape_and_cache + streamingLLM prompt pruning + prompt block scoring in ONE pass.
//
// Replaces csrc/cache_kernels.cu:152-205 (reshape_and_cache_kernel: grid = tokens,
// scalar 2-byte stores, V stores strided by block_size) and, for prefill, the python
// slice/cat/copy of KVCachePruner.prune_prompt_inplace
// (vllm/attention/kvcache_prunner.py:183-218) driven by
// _prune_prompts_keys_and_values_new (vllm/attention/backends/torch_cuda_sdpa.py:1222-1296)
// plus the per-block get_block_score loop (torch_cuda_sdpa.py:1278-1288).
//
// Unit of work = one PAGE = (tile of up to BS consecutive kept tokens of one request,
// kv head): 128 threads gather the tile's K/V rows (16-byte coalesced reads, the prune
// gather is index arithmetic: src = j < first ? j : j + pruned), transpose through
// shared memory and emit the cache layout K [D/8][BS][8] / V [D][BS] with 16-byte
// stores that are contiguous across lanes whenever the tile maps onto one block
// (always true for prefill).  HBM-bound: N*Hkv*D*2 B read + the same written.
#include "pc_common.cuh"
#include "score_common.cuh"
#include "pc_api_internal.h"

namespace pc {

struct CacheWriteParams {
  const __nv_bfloat16* __restrict__ key;
  const __nv_bfloat16* __restrict__ value;
  __nv_bfloat16* __restrict__ k_cache;
  __nv_bfloat16* __restrict__ v_cache;
  const long long* __restrict__ slot_mapping;
  const int* __restrict__ req_meta;      // [R][5] {src_start, kept, first, pruned, dst_start}
  const int* __restrict__ tile_offsets;  // [R+1]
  int num_reqs, num_tokens, Hkv;
  long long key_stride, value_stride, kv_block_stride;
  ScorePool score;
};

template <int D, int BS>
__global__ void __launch_bounds__(128) cache_write_kernel(const CacheWriteParams p) {
  constexpr int VEC = D / 8;          // 16-byte vectors per row per head
  constexpr int KROW = VEC + 1;       // padded K tile row (uint4 units)
  constexpr int VROW = D + 8;         // padded V tile row (bf16 units), keeps 16-B alignment
  __shared__ uint4 ks[BS * KROW];
  __shared__ __align__(16) __nv_bfloat16 vs[BS * VROW];
  __shared__ long long slot_s[BS];
  __shared__ long long src_s[BS];
  __shared__ int flags_s[2];  // [0] = tile maps onto one whole block, in order

  const int tile = blockIdx.x;
  const int h = blockIdx.y;
  const int tid = threadIdx.x;

  // ---- which request / which block of it
  int r = 0, src_start = 0, kept = p.num_tokens, first = p.num_tokens, pruned = 0, dst_start = 0,
      tile0 = 0;
  if (p.req_meta != nullptr) {
    int lo = 0, hi = p.num_reqs;  // last r with tile_offsets[r] <= tile
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (p.tile_offsets[mid] <= tile) lo = mid; else hi = mid;
    }
    r = lo;
    tile0 = p.tile_offsets[r];
    const int* m = p.req_meta + r * 5;
    src_start = m[0]; kept = m[1]; first = m[2]; pruned = m[3]; dst_start = m[4];
  }
  const int j = tile - tile0;  // block index inside the request's table
  const int ntok = min(BS, kept - j * BS);
  if (ntok <= 0) return;

  if (tid < BS) {
    long long slot = -1, src = 0;
    if (tid < ntok) {
      const int jj = j * BS + tid;
      slot = p.slot_mapping[dst_start + jj];
      src = (long long)src_start + (jj < first ? jj : jj + pruned);
    }
    slot_s[tid] = slot;
    src_s[tid] = src;
  }
  __syncthreads();
  if (tid == 0) {
    bool whole = (ntok == BS) && (slot_s[0] >= 0) && (slot_s[0] % BS == 0);
    for (int i = 1; whole && i < BS; ++i) whole = (slot_s[i] == slot_s[0] + i);
    flags_s[0] = whole ? 1 : 0;
  }

  // ---- gather K/V rows of head h (c fastest: 16 lanes read one 256-byte row segment)
  for (int i = tid; i < BS * VEC; i += 128) {
    const int tl = i / VEC, c = i % VEC;
    uint4 kv4 = make_uint4(0, 0, 0, 0), vv4 = make_uint4(0, 0, 0, 0);
    if (tl < ntok) {
      const size_t ko = (size_t)src_s[tl] * p.key_stride + (size_t)h * D + c * 8;
      const size_t vo = (size_t)src_s[tl] * p.value_stride + (size_t)h * D + c * 8;
      kv4 = __ldg(reinterpret_cast<const uint4*>(p.key + ko));
      vv4 = __ldg(reinterpret_cast<const uint4*>(p.value + vo));
    }
    ks[tl * KROW + c] = kv4;
    *reinterpret_cast<uint4*>(&vs[tl * VROW + c * 8]) = vv4;
  }
  __syncthreads();
  const bool whole = flags_s[0] != 0;

  // ---- K: [c][tok] 16-byte units, tok fastest -> 256 contiguous bytes per chunk
  for (int i = tid; i < BS * VEC; i += 128) {
    const int tl = i % BS, c = i / BS;
    const long long slot = slot_s[tl];
    if (slot < 0) continue;
    const long long blk = slot / BS;
    const int off = (int)(slot % BS);
    __nv_bfloat16* dst = p.k_cache + (size_t)blk * p.kv_block_stride + (size_t)h * (D * BS) +
                         (size_t)(c * BS + off) * 8;
    *reinterpret_cast<uint4*>(dst) = ks[tl * KROW + c];
  }

  // ---- V: [d][tok]
  if (whole) {
    const long long blk = slot_s[0] / BS;
    __nv_bfloat16* base = p.v_cache + (size_t)blk * p.kv_block_stride + (size_t)h * (D * BS);
    for (int i = tid; i < D * (BS / 8); i += 128) {
      const int hv = i % (BS / 8), d = i / (BS / 8);
      __align__(16) __nv_bfloat16 tmp[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) tmp[k] = vs[(hv * 8 + k) * VROW + d];
      *reinterpret_cast<uint4*>(base + (size_t)d * BS + hv * 8) = *reinterpret_cast<uint4*>(tmp);
    }
  } else {
    for (int i = tid; i < D * BS; i += 128) {
      const int tl = i % BS, d = i / BS;
      const long long slot = slot_s[tl];
      if (slot < 0) continue;
      const long long blk = slot / BS;
      const int off = (int)(slot % BS);
      p.v_cache[(size_t)blk * p.kv_block_stride + (size_t)h * (D * BS) + (size_t)d * BS + off] =
          vs[tl * VROW + d];
    }
  }

  // ---- prompt block score (full kept blocks only), per kv head, fixed fp32 order
  if (p.score.n_pools > 0 && whole && tid < 32) {
    const float sc = warp_value_l2_head_score<D, BS>(
        [&](int tok, int d) { return __bfloat162float(vs[tok * VROW + d]); });
    if (tid == 0) score_store(p.score, slot_s[0] / BS, h, sc);
  }
}

template <int D, int BS>
static int launch_cache_write(const CacheWriteParams& p, int tiles, cudaStream_t stream) {
  dim3 grid(tiles, p.Hkv);
  cache_write_kernel<D, BS><<<grid, 128, 0, stream>>>(p);
  return check_launch("cache_write_kernel");
}

int cache_write(const CacheWriteArgs& a, cudaStream_t stream) {
  if (a.key_stride % 8 || a.value_stride % 8)
    return set_error(PC_ERR_INVALID, "key/value row strides must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(a.key) | reinterpret_cast<uintptr_t>(a.value) |
       reinterpret_cast<uintptr_t>(a.k_cache) | reinterpret_cast<uintptr_t>(a.v_cache)) & 15)
    return set_error(PC_ERR_INVALID, "key/value/cache pointers must be 16-byte aligned");
  CacheWriteParams p;
  p.key = static_cast<const __nv_bfloat16*>(a.key);
  p.value = static_cast<const __nv_bfloat16*>(a.value);
  p.k_cache = static_cast<__nv_bfloat16*>(a.k_cache);
  p.v_cache = static_cast<__nv_bfloat16*>(a.v_cache);
  p.slot_mapping = a.slot_mapping;
  p.req_meta = a.req_meta;
  p.tile_offsets = a.tile_offsets;
  p.num_reqs = a.num_reqs;
  p.num_tokens = a.num_tokens;
  p.Hkv = a.Hkv;
  p.key_stride = a.key_stride;
  p.value_stride = a.value_stride;
  p.kv_block_stride = a.kv_block_stride;
  p.score = a.score;
  const int tiles = a.req_meta ? a.total_tiles : (a.num_tokens + a.BS - 1) / a.BS;
  if (tiles <= 0) return 0;
  if (a.D == 128 && a.BS == 16) return launch_cache_write<128, 16>(p, tiles, stream);
  if (a.D == 128 && a.BS == 32) return launch_cache_write<128, 32>(p, tiles, stream);
  if (a.D == 64 && a.BS == 16) return launch_cache_write<64, 16>(p, tiles, stream);
  if (a.D == 64 && a.BS == 32) return launch_cache_write<64, 32>(p, tiles, stream);
  return set_error(PC_ERR_UNSUPPORTED,
                   "Unsupported head size / block size (supported: head 64|128, block 16|32)");
}

}  // namespace pc
