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
#include <atomic>

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

// ------------------------------------------------------------------------------------------
// Prompt-time fast path: one CTA = one cache BLOCK (BS kept tokens) x HC kv heads.
//
// In the cache a block is one contiguous run per pool: K [Hkv][D/8][BS][8], V [Hkv][D][BS].
//   K needs no element transposition, only a permutation of 16-byte units ([tok][h][c] ->
//     [h][c][tok]): a warp loads 8 tokens x 4 consecutive units (64 contiguous bytes per
//     token row) and stores 4 x 128 contiguous bytes -- registers only, full sectors both ways.
//   V is a real [tok][h*D] -> [h*D][tok] transposition: the BS token rows arrive by TMA 1-D bulk
//     copies (one per gathered row: the streamingLLM prune is the row index) into a padded
//     staging tile, ldmatrix.trans + stmatrix move 8x8 tiles into the block image, and ONE
//     bulk store (cp.async.bulk shared -> global) writes the HC*D*BS*2 contiguous bytes.
//   The prompt block score (S3) reads the staged V rows: zero extra HBM bytes.
// Tiles that do not map onto one whole block in slot order (the last, partial tile of a prompt;
// callers with arbitrary slot mappings) take an element-wise path in the same kernel.
__device__ __forceinline__ void ldmatrix_x4_trans_(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void stmatrix_x4(uint32_t addr, const uint32_t (&r)[4]) {
  asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3])
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

constexpr int kBlockThreads = 256;

template <int D, int BS, int HC>
struct BlockWriteSmem {
  static constexpr int kRowBytes = HC * D * 2;            // one token's rows of the CTA's heads
  static constexpr int kPitch = kRowBytes + 16;           // staging pitch: ldmatrix rows hit distinct banks
  static constexpr int kStage = BS * kPitch;
  static constexpr int kImage = HC * D * BS * 2;          // the V block image of these heads
  static constexpr int kBytes = kStage + kImage + 64;
};

template <int D, int BS, int HC>
__global__ void __launch_bounds__(kBlockThreads) cache_write_block_kernel(const CacheWriteParams p) {
  using L = BlockWriteSmem<D, BS, HC>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* vstage = smem;
  uint8_t* vimage = smem + L::kStage;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::kStage + L::kImage);
  constexpr int NW = kBlockThreads / 32;

  const int tile = blockIdx.x;
  const int h0 = blockIdx.y * HC;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- which request: r = #{1 <= i < R : tile_offsets[i] <= tile}; every load independent, so
  // the CTA pays one memory latency here instead of a dependent binary search
  int r = 0;
  for (int base = 1; base < p.num_reqs; base += kBlockThreads) {
    const int i = base + tid;
    r += __syncthreads_count(i < p.num_reqs && __ldg(p.tile_offsets + i) <= tile);
  }
  const int* m = p.req_meta + r * 5;
  const int src_start = __ldg(m), kept = __ldg(m + 1), first = __ldg(m + 2), pruned = __ldg(m + 3),
            dst_start = __ldg(m + 4);
  const int j = tile - __ldg(p.tile_offsets + r);
  const int ntok = min(BS, kept - j * BS);
  if (ntok <= 0) return;
  // source row of the tile's token tk: the streamingLLM prune is index arithmetic
  auto src_row = [&](int tk) -> size_t {
    const int jj = j * BS + tk;
    return (size_t)src_start + (size_t)(jj < first ? jj : jj + pruned);
  };

  // ---- V rows leave for shared memory as soon as the request is known (no slot needed)
  if (warp == 0) {
    if (lane == 0) {
      mbar_init(bar, 1);
      mbar_fence_init();
      mbar_expect_tx(bar, ntok * L::kRowBytes);
    }
    __syncwarp();
    if (lane < ntok)
      tma_load_1d(vstage + lane * L::kPitch, p.value + src_row(lane) * p.value_stride + (size_t)h0 * D,
                  L::kRowBytes, bar);
  }

  // ---- K: unit (tok, u) -> (u, tok), u = 16-byte unit inside the CTA's heads' row.  One warp
  // instruction = 8 tokens x 4 units: 64 contiguous bytes per token row in, 4 x 128 contiguous
  // bytes out.  The first batch of loads is in flight before the slots are known.
  constexpr int U = L::kRowBytes / 16;          // units per token row
  constexpr int TH = BS / 8;                    // 8-token groups
  constexpr int GROUPS = TH * (U / 4);
  constexpr int KB = 8;                         // loads in flight per thread
  const int tok_lo = lane & 7, q4 = lane >> 3;
  uint4 kreg[KB];
  auto k_load = [&](int g0) {
#pragma unroll
    for (int i = 0; i < KB; ++i) {
      const int gi = g0 + i * NW + warp;
      if (gi < GROUPS) {
        const int tok = (gi % TH) * 8 + tok_lo, u = (gi / TH) * 4 + q4;
        if (tok < ntok)
          kreg[i] = __ldg(reinterpret_cast<const uint4*>(p.key + src_row(tok) * p.key_stride + (size_t)h0 * D) + u);
      }
    }
  };
  k_load(0);

  // ---- slots: the tile must be one whole block, in order
  const long long slot0 = p.slot_mapping[dst_start + j * BS];
  bool ok = (ntok == BS) && slot0 >= 0 && (slot0 % BS == 0);
  if (tid < ntok) ok = ok && (p.slot_mapping[dst_start + j * BS + tid] == slot0 + tid);
  const bool whole = __syncthreads_and(ok) != 0;   // also publishes the mbarrier init

  if (!whole) {
    // element-wise path (csrc/cache_kernels.cu:152-205 semantics; slot < 0 = skip)
    mbar_wait(bar, 0);                             // the staged rows must land before the CTA exits
    for (int i = tid; i < ntok * HC * D; i += kBlockThreads) {
      const int tl = i / (HC * D), e = i - tl * (HC * D);
      const int h = e / D, d = e - h * D;
      const long long slot = p.slot_mapping[dst_start + j * BS + tl];
      if (slot < 0) continue;
      const long long blk = slot / BS;
      const int off = (int)(slot - blk * BS);
      const size_t base = (size_t)blk * p.kv_block_stride + (size_t)(h0 + h) * (D * BS);
      p.k_cache[base + (size_t)((d >> 3) * BS + off) * 8 + (d & 7)] =
          p.key[src_row(tl) * p.key_stride + (size_t)(h0 + h) * D + d];
      p.v_cache[base + (size_t)d * BS + off] = p.value[src_row(tl) * p.value_stride + (size_t)(h0 + h) * D + d];
    }
    return;
  }

  const long long blk = slot0 / BS;
  const size_t gbase = (size_t)blk * p.kv_block_stride + (size_t)h0 * (D * BS);
  {
    uint4* kdst = reinterpret_cast<uint4*>(p.k_cache + gbase);
    for (int g0 = 0; g0 < GROUPS; g0 += KB * NW) {
      if (g0 > 0) k_load(g0);
#pragma unroll
      for (int i = 0; i < KB; ++i) {
        const int gi = g0 + i * NW + warp;
        if (gi < GROUPS) {
          const int tok = (gi % TH) * 8 + tok_lo, u = (gi / TH) * 4 + q4;
          kdst[u * BS + tok] = kreg[i];
        }
      }
    }
  }

  // ---- V: staged rows -> transposed block image
  mbar_wait(bar, 0);
  {
    constexpr int TQ = BS / 8;                    // 8-token groups per block
    constexpr int EG = 4 / TQ;                    // 8-element groups per ldmatrix.x4
    constexpr int OPS = HC * D / (8 * EG);
    const int mj = lane >> 3, rr = lane & 7;      // this lane addresses row rr of matrix mj
    const int tg = mj % TQ, eg = mj / TQ;
    const uint32_t s_stage = smem_u32(vstage), s_image = smem_u32(vimage);
    for (int op = warp; op < OPS; op += NW) {
      const int e0 = op * 8 * EG;
      uint32_t frag[4];
      // source tile mj: tokens 8 tg .. +7, elements e0 + 8 eg .. +7 (row = token)
      ldmatrix_x4_trans_(frag, s_stage + (uint32_t)((tg * 8 + rr) * L::kPitch + (e0 + 8 * eg) * 2));
      // transposed tile mj: rows = elements e0 + 8 eg + rr, 8 tokens (16 bytes) at token group tg
      stmatrix_x4(s_image + (uint32_t)((e0 + 8 * eg + rr) * (BS * 2) + tg * 16), frag);
    }
  }
  fence_proxy_async();   // the image was written through the generic proxy
  __syncthreads();
  if (tid == 0) bulk_store(p.v_cache + gbase, vimage, L::kImage);
  // ---- prompt block score per kv head, fixed fp32 order (one warp per head), while the store drains
  if (p.score.n_pools > 0) {
    for (int h = warp; h < HC; h += NW) {
      const float sc = warp_value_l2_head_score<D, BS>([&](int tok, int d) {
        return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(vstage + tok * L::kPitch + (h * D + d) * 2));
      });
      if (lane == 0) score_store(p.score, blk, h0 + h, sc);
    }
  }
  if (tid == 0) bulk_commit_wait_read();
}

template <int D, int BS, int HC>
static int launch_block_write(const CacheWriteParams& p, int tiles, cudaStream_t stream) {
  using L = BlockWriteSmem<D, BS, HC>;
  auto kern = cache_write_block_kernel<D, BS, HC>;
  static std::atomic<int> opted[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !opted[dev].load(std::memory_order_acquire)) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes);
    if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(cache_write_block)", e);
    if (dev >= 0 && dev < 64) opted[dev].store(1, std::memory_order_release);
  }
  dim3 grid(tiles, p.Hkv / HC);
  kern<<<grid, kBlockThreads, L::kBytes, stream>>>(p);
  return check_launch("cache_write_block_kernel");
}

// ------------------------------------------------------------------------------------------
// Prompt-time streaming path: PERSISTENT CTAs, one per SM, each walking a contiguous range of
// work items (item = one cache block x HC kv heads) through a multi-stage shared-memory pipeline.
//
// The block kernel above is one short-lived CTA per item: request lookup -> slot check -> loads
// -> stores are a dependent chain that every CTA pays again, and only ~3 CTAs per SM overlap
// (ncu, profiles/r02_cache_write_block_ncu_summary.txt: 36 % warps active, 4.6 TB/s at 64
// prompts).  Here one PRODUCER warp runs ahead of eight CONSUMER warps:
//   producer : item -> (request, block) with the request carried in registers (items of a CTA are
//              consecutive blocks, a request boundary is crossed every ~34 items), the block's slots
//              prefetched one item ahead, then BS + BS TMA 1-D row copies (K rows and V rows of the
//              item's HC heads, gathered: the streamingLLM prune is the row index) into stage s,
//              completing on full[s]; NST stages (128 KiB of loads in flight per SM at D 128)
//   consumers: K units [tok][u] -> [u][tok] from shared memory to the pool with 16-byte stores that
//              are contiguous across the warp; V rows -> transposed block image (ldmatrix.trans +
//              stmatrix) -> ONE bulk store per item (two images, the store of item i drains while
//              item i+1 is transposed); prompt block scores from the staged V rows; then empty[s].
struct StreamDesc {      // written by the producer, read by the consumers after full[s]
  long long blk;         // physical block (whole items)
  int ntok, whole, h0, pad;
  long long pad1;        // 32 bytes: the per-token slots follow
};
static_assert(sizeof(StreamDesc) == 32, "StreamDesc must be 32 bytes");

constexpr int kStreamConsumerWarps = 8;
constexpr int kStreamThreads = (kStreamConsumerWarps + 1) * 32;

template <int D, int BS, int HC, int NST>
struct StreamSmem {
  static constexpr int kRow = HC * D * 2;               // one token's rows of the item's heads
  static constexpr int kPitch = kRow + 16;              // staging pitch (bank spread for ldmatrix / LDS.128)
  static constexpr int kTensor = BS * kPitch;           // K (or V) rows of one item
  static constexpr int kStageIn = 2 * kTensor;
  static constexpr int kImage = HC * D * BS * 2;        // V block image of the item's heads
  static constexpr int kDesc = 32 + BS * 8;             // StreamDesc + per-token slots
  static constexpr int kOffImage = NST * kStageIn;
  static constexpr int kOffDesc = kOffImage + 2 * kImage;
  static constexpr int kOffBar = kOffDesc + NST * kDesc;
  static constexpr int kBytes = kOffBar + 2 * NST * 8;
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

template <int D, int BS, int HC, int NST>
__global__ void __launch_bounds__(kStreamThreads, 1) cache_write_stream_kernel(const CacheWriteParams p, int total_tiles) {
  using L = StreamSmem<D, BS, HC, NST>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::kOffBar);
  uint64_t* empty = full + NST;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int HG = p.Hkv / HC;
  const long long n_items = (long long)total_tiles * HG;
  // contiguous item range of this CTA
  const long long it_begin = n_items * blockIdx.x / gridDim.x;
  const long long it_end = n_items * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kStreamConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kStreamConsumerWarps) {
    // =============================================================== producer
    if (it_begin >= it_end) return;
    int r = 0, t_lo = 0, t_hi = 0;
    int src_start = 0, kept = 0, first = 0, pruned = 0, dst_start = 0;
    auto locate = [&](int tile, bool from_scratch) {   // all lanes, uniform result
      if (from_scratch) {
        int cnt = 0;   // r = #{1 <= i < R : tile_offsets[i] <= tile}: independent loads
        for (int i = 1 + lane; i < p.num_reqs; i += 32) cnt += (__ldg(p.tile_offsets + i) <= tile) ? 1 : 0;
#pragma unroll
        for (int sh = 16; sh >= 1; sh >>= 1) cnt += __shfl_xor_sync(kFull, cnt, sh);
        r = cnt;
      } else {
        int adv;
        do {       // request ends are monotone: the lanes whose request ends at or before `tile` form a prefix
          const int i = r + 1 + lane;
          const int e = (i < p.num_reqs) ? __ldg(p.tile_offsets + i) : 0x7fffffff;
          adv = __popc(__ballot_sync(kFull, e <= tile));
          r += adv;
        } while (adv == 32);
      }
      int v = 0;
      if (lane < 5) v = __ldg(p.req_meta + r * 5 + lane);
      else if (lane < 7) v = __ldg(p.tile_offsets + r + (lane - 5));
      src_start = __shfl_sync(kFull, v, 0); kept = __shfl_sync(kFull, v, 1); first = __shfl_sync(kFull, v, 2);
      pruned = __shfl_sync(kFull, v, 3); dst_start = __shfl_sync(kFull, v, 4);
      t_lo = __shfl_sync(kFull, v, 5); t_hi = __shfl_sync(kFull, v, 6);
    };
    auto load_slot = [&](int tile) -> long long {     // lane's slot of the item's block (after locate)
      const int j = tile - t_lo;
      const int ntok = min(BS, kept - j * BS);
      return (lane < ntok) ? __ldg(p.slot_mapping + dst_start + j * BS + lane) : -1;
    };
    int tile = (int)(it_begin / HG);
    const int last_tile = (int)((it_end - 1) / HG);
    int stage = 0;
    uint32_t phase = 0;
    long long it = it_begin;
    bool first_locate = true;
    constexpr int PF = 4;          // blocks whose slots are fetched together (one memory latency per group)
    while (it < it_end) {
      if (first_locate || tile >= t_hi) locate(tile, first_locate);
      first_locate = false;
      const int n = min(PF, min(t_hi - tile, last_tile - tile + 1));
      long long sl[PF];
#pragma unroll
      for (int k = 0; k < PF; ++k) sl[k] = (k < n) ? load_slot(tile + k) : -1;
#pragma unroll
      for (int k = 0; k < PF; ++k) {
        if (k >= n) break;
        const long long slot = sl[k];
        const int j = tile - t_lo;
        const int ntok = min(BS, kept - j * BS);
        // whole block, in slot order?
        const long long slot0 = __shfl_sync(kFull, slot, 0);
        const bool ok = (ntok == BS) && slot0 >= 0 && (slot0 % BS == 0);
        const bool mine = (lane >= ntok) || (slot == slot0 + lane);
        const bool whole = ok && (__ballot_sync(kFull, mine) == kFull);
        // source row of this lane's token (the streamingLLM prune is index arithmetic)
        const int jj = j * BS + lane;
        const size_t srow = (size_t)src_start + (size_t)(jj < first ? jj : jj + pruned);
        for (int hg = (int)(it - (long long)tile * HG); hg < HG && it < it_end; ++hg, ++it) {
          mbar_wait(&empty[stage], phase ^ 1u);
          uint8_t* st = smem + (size_t)stage * L::kStageIn;
          StreamDesc* d = reinterpret_cast<StreamDesc*>(smem + L::kOffDesc + stage * L::kDesc);
          if (lane < BS) reinterpret_cast<long long*>(d + 1)[lane] = slot;
          if (lane == 0) {
            d->blk = slot0 / BS; d->ntok = ntok; d->whole = whole ? 1 : 0; d->h0 = hg * HC; d->pad = 0;
          }
          __syncwarp();
          if (ntok > 0) {
            if (lane == 0) mbar_expect_tx(&full[stage], 2u * (uint32_t)ntok * L::kRow);
            __syncwarp();
            if (lane < ntok) {
              tma_load_1d(st + lane * L::kPitch, p.key + srow * p.key_stride + (size_t)hg * HC * D, L::kRow, &full[stage]);
              tma_load_1d(st + L::kTensor + lane * L::kPitch, p.value + srow * p.value_stride + (size_t)hg * HC * D,
                          L::kRow, &full[stage]);
            }
          } else if (lane == 0) {
            mbar_arrive(&full[stage]);
          }
          if (++stage == NST) { stage = 0; phase ^= 1u; }
        }
        ++tile;
      }
    }
    return;
  }

  // ================================================================= consumers
  const int ctid = threadIdx.x;                 // 0 .. 255
  constexpr int CT = kStreamConsumerWarps * 32;
  int stage = 0;
  uint32_t phase = 0;
  int ib = 0;                                    // V image buffer of this item
  for (long long it = it_begin; it < it_end; ++it) {
    // the bulk store that last read image `ib` (two items ago) must be done reading it
    if (ctid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    mbar_wait(&full[stage], phase);
    named_bar_sync(1, CT);
    const uint8_t* kst = smem + (size_t)stage * L::kStageIn;
    const uint8_t* vst = kst + L::kTensor;
    uint8_t* vimage = smem + L::kOffImage + ib * L::kImage;
    const StreamDesc d = *reinterpret_cast<const StreamDesc*>(smem + L::kOffDesc + stage * L::kDesc);
    const long long* slots = reinterpret_cast<const long long*>(smem + L::kOffDesc + stage * L::kDesc + 32);
    if (d.whole) {
      const size_t gbase = (size_t)d.blk * p.kv_block_stride + (size_t)d.h0 * (D * BS);
      // ---- K: unit g = u * BS + tok of the destination (contiguous across the warp)
      {
        uint4* kdst = reinterpret_cast<uint4*>(p.k_cache + gbase);
        constexpr int UNITS = BS * L::kRow / 16;
#pragma unroll 4
        for (int g = ctid; g < UNITS; g += CT) {
          const int u = g / BS, tok = g % BS;
          kdst[g] = *reinterpret_cast<const uint4*>(kst + tok * L::kPitch + u * 16);
        }
      }
      // ---- V: staged rows -> transposed block image
      {
        constexpr int TQ = BS / 8;                    // 8-token groups per block
        constexpr int EG = 4 / TQ;                    // 8-element groups per ldmatrix.x4
        constexpr int OPS = HC * D / (8 * EG);
        const int mj = lane >> 3, rr = lane & 7;
        const int tg = mj % TQ, eg = mj / TQ;
        const uint32_t s_stage = smem_u32(vst), s_image = smem_u32(vimage);
        for (int op = warp; op < OPS; op += kStreamConsumerWarps) {
          const int e0 = op * 8 * EG;
          uint32_t frag[4];
          ldmatrix_x4_trans_(frag, s_stage + (uint32_t)((tg * 8 + rr) * L::kPitch + (e0 + 8 * eg) * 2));
          stmatrix_x4(s_image + (uint32_t)((e0 + 8 * eg + rr) * (BS * 2) + tg * 16), frag);
        }
      }
      fence_proxy_async();   // the image was written through the generic proxy
      named_bar_sync(2, CT);
      if (ctid == 0) {
        bulk_store(p.v_cache + gbase, vimage, L::kImage);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      // ---- prompt block score per kv head, fixed fp32 order (one warp per head)
      if (p.score.n_pools > 0) {
        for (int h = warp; h < HC; h += kStreamConsumerWarps) {
          const float sc = warp_value_l2_head_score<D, BS>([&](int tok, int dd) {
            return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(vst + tok * L::kPitch + (h * D + dd) * 2));
          });
          if (lane == 0) score_store(p.score, d.blk, d.h0 + h, sc);
        }
      }
      ib ^= 1;
    } else if (d.ntok > 0) {
      // element-wise path (csrc/cache_kernels.cu:152-205 semantics; slot < 0 = skip), from the staged rows
      for (int i = ctid; i < d.ntok * HC * D; i += CT) {
        const int tl = i / (HC * D), e = i - tl * (HC * D);
        const int h = e / D, dd = e - h * D;
        const long long slot = slots[tl];
        if (slot < 0) continue;
        const long long blk = slot / BS;
        const int off = (int)(slot - blk * BS);
        const size_t base = (size_t)blk * p.kv_block_stride + (size_t)(d.h0 + h) * (D * BS);
        p.k_cache[base + (size_t)((dd >> 3) * BS + off) * 8 + (dd & 7)] =
            *reinterpret_cast<const __nv_bfloat16*>(kst + tl * L::kPitch + e * 2);
        p.v_cache[base + (size_t)dd * BS + off] = *reinterpret_cast<const __nv_bfloat16*>(vst + tl * L::kPitch + e * 2);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[stage]);
    if (++stage == NST) { stage = 0; phase ^= 1u; }
  }
  if (ctid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int D, int BS, int HC, int NST>
static int launch_stream_write(const CacheWriteParams& p, int tiles, cudaStream_t stream) {
  using L = StreamSmem<D, BS, HC, NST>;
  static_assert(L::kBytes <= 227 * 1024, "stream pipeline does not fit shared memory");
  auto kern = cache_write_stream_kernel<D, BS, HC, NST>;
  static std::atomic<int> opted[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !opted[dev].load(std::memory_order_acquire)) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes);
    if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(cache_write_stream)", e);
    if (dev >= 0 && dev < 64) opted[dev].store(1, std::memory_order_release);
  }
  const int sms = device_sm_count();
  if (sms <= 0) return PC_ERR_CUDA;
  const long long items = (long long)tiles * (p.Hkv / HC);
  const int grid = (int)(items < sms ? items : sms);
  kern<<<grid, kStreamThreads, L::kBytes, stream>>>(p, tiles);
  return check_launch("cache_write_stream_kernel");
}

// stages: as many as fit next to the two V images, at most 4
template <int D, int BS, int HC>
constexpr int stream_stages() {
  int n = 4;
  while (n > 2 && StreamSmem<D, BS, HC, 4>::kStageIn * n + 2 * StreamSmem<D, BS, HC, 4>::kImage + 4096 > 220 * 1024) --n;
  return n;
}
// heads per item: as many as keep a stage at <= 64 KiB of rows (K + V)
template <int D, int BS>
static int dispatch_stream_write(const CacheWriteParams& p, int tiles, cudaStream_t stream) {
  constexpr int kMaxHC = (D * BS <= 2048) ? 8 : (D * BS <= 4096 ? 4 : 2);   // 2 * BS * HC * D * 2 bytes per stage
  constexpr int H8 = kMaxHC >= 8 ? 8 : kMaxHC, H4 = kMaxHC >= 4 ? 4 : kMaxHC;
  if (p.Hkv % 8 == 0 && kMaxHC >= 8) return launch_stream_write<D, BS, H8, stream_stages<D, BS, H8>()>(p, tiles, stream);
  if (p.Hkv % 4 == 0 && kMaxHC >= 4) return launch_stream_write<D, BS, H4, stream_stages<D, BS, H4>()>(p, tiles, stream);
  if (p.Hkv % 2 == 0) return launch_stream_write<D, BS, 2, stream_stages<D, BS, 2>()>(p, tiles, stream);
  return launch_stream_write<D, BS, 1, stream_stages<D, BS, 1>()>(p, tiles, stream);
}

#ifdef PC_DEBUG
static int g_force_block_hc = 0;   // heads per CTA override (sweeps)
void set_cache_write_block_heads(int hc) { g_force_block_hc = hc; }
#endif

// heads per CTA of the block kernel: as many as keep staging + image under ~75 KB (3 CTAs / SM)
template <int D, int BS>
static int dispatch_block_write(const CacheWriteParams& p, int tiles, cudaStream_t stream) {
  constexpr int kMaxHC = (D * BS <= 2048) ? 8 : (D * BS <= 4096 ? 4 : 2);   // D*BS*2*2 bytes per head
#ifdef PC_DEBUG
  if (g_force_block_hc > 0 && p.Hkv % g_force_block_hc == 0) {
    switch (g_force_block_hc) {
      case 1: return launch_block_write<D, BS, 1>(p, tiles, stream);
      case 2: return launch_block_write<D, BS, 2>(p, tiles, stream);
      case 4: if (kMaxHC >= 4) return launch_block_write<D, BS, (kMaxHC >= 4 ? 4 : 1)>(p, tiles, stream); break;
      case 8: if (kMaxHC >= 8) return launch_block_write<D, BS, (kMaxHC >= 8 ? 8 : 1)>(p, tiles, stream); break;
    }
  }
#endif
  if (kMaxHC >= 8 && p.Hkv % 8 == 0) return launch_block_write<D, BS, (kMaxHC >= 8 ? 8 : kMaxHC)>(p, tiles, stream);
  if (kMaxHC >= 4 && p.Hkv % 4 == 0) return launch_block_write<D, BS, (kMaxHC >= 4 ? 4 : kMaxHC)>(p, tiles, stream);
  if (p.Hkv % 2 == 0) return launch_block_write<D, BS, 2>(p, tiles, stream);
  return launch_block_write<D, BS, 1>(p, tiles, stream);
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
  if (a.req_meta != nullptr) {   // prompt-time: whole blocks
    // The streaming kernel hands ONE item at a time to all consumer warps: it wins when an item
    // (block x heads per CTA) is 64 KiB of rows (us per launch, stream / block kernel, 285 MB:
    // Hkv 8 53.7 / 62.2, block 32 67.2 / 77.8); with 1-4 kv heads per rank (TP 2 / 4 / 8) items are
    // 8-32 KiB and many small concurrent CTAs (the block kernel) do better (Hkv 4 64.2 / 61.0,
    // Hkv 2 101 / 62, Hkv 1 179 / 71) -- profiles/r02_cache_write_stream.txt.
    const int hc = a.Hkv % 8 == 0 ? 8 : (a.Hkv % 4 == 0 ? 4 : (a.Hkv % 2 == 0 ? 2 : 1));
    const int hc_cap = (a.D * a.BS <= 2048) ? 8 : (a.D * a.BS <= 4096 ? 4 : 2);
    const long long item_bytes = 2LL * a.BS * (hc < hc_cap ? hc : hc_cap) * a.D * 2;
    bool streaming = item_bytes >= 64 * 1024 || (item_bytes >= 32 * 1024 && a.Hkv >= 8);
#ifdef PC_DEBUG
    if (g_force_block_hc > 0) streaming = false;   // A/B runs: the block kernel with that many heads per CTA
    if (g_force_block_hc < 0) streaming = true;
#endif
    if (streaming) {
      if (a.D == 128 && a.BS == 16) return dispatch_stream_write<128, 16>(p, tiles, stream);
      if (a.D == 128 && a.BS == 32) return dispatch_stream_write<128, 32>(p, tiles, stream);
      if (a.D == 64 && a.BS == 16) return dispatch_stream_write<64, 16>(p, tiles, stream);
      if (a.D == 64 && a.BS == 32) return dispatch_stream_write<64, 32>(p, tiles, stream);
      if (a.D == 128 && a.BS == 8) return dispatch_stream_write<128, 8>(p, tiles, stream);
      if (a.D == 64 && a.BS == 8) return dispatch_stream_write<64, 8>(p, tiles, stream);
    }
    if (a.D == 128 && a.BS == 16) return dispatch_block_write<128, 16>(p, tiles, stream);
    if (a.D == 128 && a.BS == 32) return dispatch_block_write<128, 32>(p, tiles, stream);
    if (a.D == 64 && a.BS == 16) return dispatch_block_write<64, 16>(p, tiles, stream);
    if (a.D == 64 && a.BS == 32) return dispatch_block_write<64, 32>(p, tiles, stream);
    if (a.D == 128 && a.BS == 8) return dispatch_block_write<128, 8>(p, tiles, stream);
    if (a.D == 64 && a.BS == 8) return dispatch_block_write<64, 8>(p, tiles, stream);
    return set_error(PC_ERR_UNSUPPORTED,
                     "Unsupported head size / block size (supported: head 64|128, block 8|16|32)");
  }
  if (a.D == 128 && a.BS == 16) return launch_cache_write<128, 16>(p, tiles, stream);
  if (a.D == 128 && a.BS == 32) return launch_cache_write<128, 32>(p, tiles, stream);
  if (a.D == 64 && a.BS == 16) return launch_cache_write<64, 16>(p, tiles, stream);
  if (a.D == 64 && a.BS == 32) return launch_cache_write<64, 32>(p, tiles, stream);
  if (a.D == 128 && a.BS == 8) return launch_cache_write<128, 8>(p, tiles, stream);
  if (a.D == 64 && a.BS == 8) return launch_cache_write<64, 8>(p, tiles, stream);
  return set_error(PC_ERR_UNSUPPORTED,
                   "Unsupported head size / block size (supported: head 64|128, block 8|16|32)");
}

}  // namespace pc
