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

}  // namespace pc
