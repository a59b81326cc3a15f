// Block-granular maintenance ops required by the backend surface
// (TorchCUDASDPABackend.swap_blocks / copy_blocks,
// vllm/attention/backends/torch_cuda_sdpa.py:75-88 -> PagedAttention.swap_blocks /
// copy_blocks, vllm/attention/ops/paged_attn.py:233-254).
// copy_blocks replaces csrc/cache_kernels.cu:69-150 (one layer per call): a block is a
// contiguous block_bytes run in both K and V pools, copied with 16-byte accesses.
#include "pc_common.cuh"
#include "pc_api_internal.h"

namespace pc {

__global__ void __launch_bounds__(256) copy_blocks_kernel(uint4* __restrict__ k_cache,
                                                          uint4* __restrict__ v_cache,
                                                          const long long* __restrict__ pairs,
                                                          long long vec_per_block) {
  const long long src = pairs[2 * blockIdx.x], dst = pairs[2 * blockIdx.x + 1];
  uint4* pool = blockIdx.y == 0 ? k_cache : v_cache;
  const uint4* s = pool + src * vec_per_block;
  uint4* d = pool + dst * vec_per_block;
  for (long long i = threadIdx.x; i < vec_per_block; i += blockDim.x) d[i] = s[i];
}

int copy_blocks(void* k_cache, void* v_cache, const long long* pairs, int num_pairs,
                long long block_bytes, cudaStream_t stream) {
  if (num_pairs <= 0) return 0;
  if (block_bytes % 16) return set_error(PC_ERR_INVALID, "block_bytes must be a multiple of 16");
  dim3 grid(num_pairs, 2);
  copy_blocks_kernel<<<grid, 256, 0, stream>>>(static_cast<uint4*>(k_cache),
                                               static_cast<uint4*>(v_cache), pairs,
                                               block_bytes / 16);
  return check_launch("copy_blocks_kernel");
}

// W2: the slot-zeroing write of `streamingLLM-1` / `inverse_key_l2`.  The reference appends
// all-zero K/V rows to the step's reshape_and_cache call (the TmpCacheSingleton rows,
// vllm/attention/backends/torch_cuda_sdpa.py:919-933,962-975, slots from :1396-1401 and
// :1444-1454); here it is one launch for EVERY layer of the step (grid slots x layers): the
// slots are chosen on the host before the step, so the zeros can land before the first layer's
// attention and the layers' attention kernels keep their dependent-launch chain.
// One CTA zeroes the K row ([D/8][.][8] -> D/8 runs of 16 bytes per head) and the V row (D x Hkv
// scalars at stride BS) of one slot for all kv heads.
struct ZeroSlotsParams {
  __nv_bfloat16* const* k_caches;   // DEVICE [layers] or nullptr (then k0 / v0 = the one layer)
  __nv_bfloat16* const* v_caches;
  __nv_bfloat16* k0;
  __nv_bfloat16* v0;
  const long long* __restrict__ slots;   // [layers][slot_stride], < 0 = skip
  long long slot_stride;
  int Hkv, D, BS;
  long long kv_block_stride;
};

__global__ void __launch_bounds__(128) zero_cache_slots_kernel(const ZeroSlotsParams p) {
  const int layer = blockIdx.y;
  const long long slot = p.slots[(size_t)layer * p.slot_stride + blockIdx.x];
  if (slot < 0) return;
  __nv_bfloat16* kc = p.k_caches ? p.k_caches[layer] : p.k0;
  __nv_bfloat16* vc = p.v_caches ? p.v_caches[layer] : p.v0;
  const long long blk = slot / p.BS;
  const int off = (int)(slot - blk * p.BS);
  const size_t base = (size_t)blk * p.kv_block_stride;
  const int HD = p.Hkv * p.D;
  // K: element (h, d) lives at h*D*BS + (d/8)*BS*8 + off*8 + d%8 -> one 16-byte run per (h, d/8)
  for (int i = threadIdx.x; i < HD / 8; i += blockDim.x) {
    const int h = i / (p.D / 8), c = i - h * (p.D / 8);
    *reinterpret_cast<uint4*>(kc + base + (size_t)h * p.D * p.BS + ((size_t)c * p.BS + off) * 8) = make_uint4(0u, 0u, 0u, 0u);
  }
  // V: element (h, d) at h*D*BS + d*BS + off
  for (int i = threadIdx.x; i < HD; i += blockDim.x)
    vc[base + (size_t)i * p.BS + off] = __float2bfloat16(0.f);
}

int zero_cache_slots(const ZeroSlotsArgs& a, cudaStream_t stream) {
  if (a.num_slots <= 0 || a.num_layers <= 0) return 0;
  if (a.D % 8 || a.Hkv <= 0 || a.BS <= 0) return set_error(PC_ERR_INVALID, "zero_cache_slots: bad shape");
  if (a.num_layers > 65535) return set_error(PC_ERR_INVALID, "zero_cache_slots: too many layers for one launch");
  ZeroSlotsParams p;
  p.k_caches = reinterpret_cast<__nv_bfloat16* const*>(a.k_caches);
  p.v_caches = reinterpret_cast<__nv_bfloat16* const*>(a.v_caches);
  p.k0 = static_cast<__nv_bfloat16*>(a.k0); p.v0 = static_cast<__nv_bfloat16*>(a.v0);
  p.slots = a.slots; p.slot_stride = a.slot_stride;
  p.Hkv = a.Hkv; p.D = a.D; p.BS = a.BS; p.kv_block_stride = a.kv_block_stride;
  zero_cache_slots_kernel<<<dim3(a.num_slots, a.num_layers), 128, 0, stream>>>(p);
  return check_launch("zero_cache_slots_kernel");
}

}  // namespace pc
