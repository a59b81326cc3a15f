// Thin torch C++ extension over the C ABI of libpagedevict_b200.so: registers the path's ops with
// the torch dispatcher so that `torch.ops.<ns>.<op>` callers work without ctypes.
//
//   namespace pagedevict_b200 : every op of include/pagedevict_b200.h a torch caller needs
//   aliases (optional)        : pagedevict_b200::alias_vllm_ops() additionally defines
//       _C::paged_attention_v1 / _C::paged_attention_v2 and _C_cache_ops::reshape_and_cache
//     with the fork's schemas (csrc/torch_bindings.cpp:27-51,444-450), so the unpatched
//     vllm/_custom_ops.py:70-130,984-996 wrappers land in this library.  It must run INSTEAD of
//     importing the fork's own vllm._C (two definitions of one schema are an error).
//
// No kernels here: tensors are unpacked into pointers / strides / the current CUDA stream and
// handed to the extern "C" entry points; failures become TORCH_CHECK errors with pc_last_error().
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include <map>
#include <memory>
#include <mutex>
#include <string>

#include "../../include/pagedevict_b200.h"

namespace {

void check(int rc) { TORCH_CHECK(rc == 0, pc_last_error()); }

void* stream_of(const at::Tensor& t) {
  return static_cast<void*>(c10::cuda::getCurrentCUDAStream(t.device().index()).stream());
}

void require_bf16_cuda(const at::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor: this library has no CPU path");
  TORCH_CHECK(t.scalar_type() == at::kBFloat16, name, ": only bf16 KV cache / activations are built");
}

// key_cache [nb, Hkv, D/8, bs, 8], value_cache [nb, Hkv, D, bs] (vllm/attention/ops/paged_attn.py:49-62)
struct CacheDims { int Hkv, D, bs; int64_t block_stride; };
CacheDims cache_dims(const at::Tensor& key_cache, const at::Tensor& value_cache) {
  TORCH_CHECK(value_cache.dim() == 4 && key_cache.dim() == 5, "key_cache / value_cache must be 5-D / 4-D");
  CacheDims d;
  d.Hkv = (int)value_cache.size(1); d.D = (int)value_cache.size(2); d.bs = (int)value_cache.size(3);
  d.block_stride = value_cache.stride(0);
  TORCH_CHECK(key_cache.stride(0) == d.block_stride, "key / value cache need equal block strides");
  return d;
}

void reshape_and_cache(at::Tensor& key, at::Tensor& value, at::Tensor& key_cache, at::Tensor& value_cache,
                       at::Tensor& slot_mapping, const std::string& kv_cache_dtype, double k_scale,
                       double v_scale) {
  TORCH_CHECK(kv_cache_dtype == "auto" && k_scale == 1.0 && v_scale == 1.0,
              "only kv_cache_dtype='auto' with k_scale == v_scale == 1 is supported");
  require_bf16_cuda(key, "key");
  require_bf16_cuda(key_cache, "key_cache");
  TORCH_CHECK(slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous(),
              "slot_mapping must be a contiguous int64 tensor");
  const c10::cuda::CUDAGuard guard(key.device());
  const CacheDims d = cache_dims(key_cache, value_cache);
  check(pc_reshape_and_cache(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                             slot_mapping.data_ptr<int64_t>(), (int32_t)slot_mapping.numel(), d.Hkv, d.D, d.bs,
                             key.stride(0), value.stride(0), d.block_stride, PC_BF16, stream_of(key)));
}

// one zero-filled split-KV workspace per (device, stream), grown on demand (the kernel leaves it
// clean after every launch, include/pagedevict_b200.h)
at::Tensor& workspace(const at::Tensor& like, size_t need) {
  static std::mutex mu;
  static std::map<std::pair<int, void*>, at::Tensor> ws;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair((int)like.device().index(), stream_of(like));
  at::Tensor& t = ws[key];
  if (!t.defined() || (size_t)t.numel() < need)
    t = at::zeros({(int64_t)std::max<size_t>(need, 1 << 20)}, like.options().dtype(at::kByte));
  return t;
}

void paged_attention_v1(at::Tensor& out, at::Tensor& query, at::Tensor& key_cache, at::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, at::Tensor& block_tables, at::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<at::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double k_scale, double v_scale, int64_t tp_rank,
                        int64_t blocksparse_local_blocks, int64_t blocksparse_vert_stride,
                        int64_t blocksparse_block_size, int64_t blocksparse_head_sliding_step) {
  TORCH_CHECK(!alibi_slopes.has_value(), "alibi_slopes are not supported by the paged-eviction backend");
  TORCH_CHECK(kv_cache_dtype == "auto" && k_scale == 1.0 && v_scale == 1.0,
              "only kv_cache_dtype='auto' with k_scale == v_scale == 1 is supported");
  TORCH_CHECK(blocksparse_vert_stride <= 1, "blocksparse attention is not supported");
  (void)tp_rank; (void)blocksparse_local_blocks; (void)blocksparse_block_size; (void)blocksparse_head_sliding_step;
  require_bf16_cuda(query, "query");
  require_bf16_cuda(key_cache, "key_cache");
  TORCH_CHECK(query.dim() == 3 && out.sizes() == query.sizes(), "query / out must be [B, Hq, D]");
  TORCH_CHECK(block_tables.scalar_type() == at::kInt && seq_lens.scalar_type() == at::kInt &&
                  block_tables.stride(-1) == 1,
              "block_tables / seq_lens must be int32, table rows contiguous");
  const c10::cuda::CUDAGuard guard(query.device());
  const CacheDims d = cache_dims(key_cache, value_cache);
  TORCH_CHECK(d.Hkv == num_kv_heads && d.bs == block_size, "cache shape does not match num_kv_heads / block_size");
  const int B = (int)query.size(0), Hq = (int)query.size(1);
  const size_t need = pc_paged_decode_workspace_bytes(B, Hq, d.Hkv, d.D, d.bs, (int32_t)max_seq_len);
  TORCH_CHECK(need > 0, "Unsupported head size / block size (supported: head 64|128, block 16|32)");
  at::Tensor& ws = workspace(query, need);
  check(pc_paged_decode_attn_score(out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                   block_tables.data_ptr<int32_t>(), block_tables.stride(0),
                                   seq_lens.data_ptr<int32_t>(), B, Hq, d.Hkv, d.D, d.bs, (int32_t)max_seq_len,
                                   (float)scale, query.stride(0), out.stride(0), d.block_stride, nullptr, nullptr,
                                   ws.data_ptr(), (size_t)ws.numel(), PC_BF16, stream_of(query)));
}

// one split-KV kernel: the v2 temporaries are accepted and left untouched
void paged_attention_v2(at::Tensor& out, at::Tensor& exp_sums, at::Tensor& max_logits, at::Tensor& tmp_out,
                        at::Tensor& query, at::Tensor& key_cache, at::Tensor& value_cache, int64_t num_kv_heads,
                        double scale, at::Tensor& block_tables, at::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<at::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double k_scale, double v_scale, int64_t tp_rank,
                        int64_t blocksparse_local_blocks, int64_t blocksparse_vert_stride,
                        int64_t blocksparse_block_size, int64_t blocksparse_head_sliding_step) {
  (void)exp_sums; (void)max_logits; (void)tmp_out;
  paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                     max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank, blocksparse_local_blocks,
                     blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step);
}

void zero_cache_slots(at::Tensor& key_cache, at::Tensor& value_cache, at::Tensor& slots) {
  require_bf16_cuda(key_cache, "key_cache");
  TORCH_CHECK(slots.is_cuda() && slots.scalar_type() == at::kLong && slots.is_contiguous(),
              "slots must be a contiguous int64 CUDA tensor");
  const c10::cuda::CUDAGuard guard(key_cache.device());
  const CacheDims d = cache_dims(key_cache, value_cache);
  check(pc_zero_cache_slots(key_cache.data_ptr(), value_cache.data_ptr(), slots.data_ptr<int64_t>(),
                            (int32_t)slots.numel(), d.Hkv, d.D, d.bs, d.block_stride, stream_of(key_cache)));
}

void prefill_attn_topleft_causal(at::Tensor& out, at::Tensor& query, at::Tensor& key, at::Tensor& value,
                                 at::Tensor& req_meta, int64_t max_q_len, double scale) {
  require_bf16_cuda(query, "query");
  TORCH_CHECK(query.dim() == 3 && key.dim() == 3 && value.dim() == 3, "q/k/v must be [T, H, D]");
  TORCH_CHECK(req_meta.is_cuda() && req_meta.scalar_type() == at::kInt && req_meta.is_contiguous() &&
                  req_meta.dim() == 2 && req_meta.size(1) == 6,
              "req_meta must be a contiguous int32 CUDA tensor [R, 6]");
  const c10::cuda::CUDAGuard guard(query.device());
  const int64_t rows = std::min(query.size(0), std::min(key.size(0), value.size(0)));
  check(pc_prefill_attn_topleft_causal(out.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(),
                                       req_meta.data_ptr<int32_t>(), (int32_t)req_meta.size(0), (int32_t)max_q_len,
                                       rows, (int32_t)query.size(1), (int32_t)key.size(1), (int32_t)query.size(2),
                                       (float)scale, query.stride(0), out.stride(0), key.stride(0), value.stride(0),
                                       nullptr, PC_BF16, stream_of(query)));
}

const char* kSchemaV1 =
    "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, "
    "float scale, Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, Tensor? alibi_slopes, "
    "str kv_cache_dtype, float k_scale, float v_scale, int tp_rank, int blocksparse_local_blocks, "
    "int blocksparse_vert_stride, int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()";
const char* kSchemaV2 =
    "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out, Tensor query, "
    "Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, "
    "int block_size, int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, float k_scale, float v_scale, "
    "int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride, int blocksparse_block_size, "
    "int blocksparse_head_sliding_step) -> ()";
const char* kSchemaRC =
    "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping, "
    "str kv_cache_dtype, float k_scale, float v_scale) -> ()";

std::unique_ptr<torch::Library> g_alias_c, g_alias_cache;

// Define the fork's three op names in ITS namespaces, implemented by this library.
void alias_vllm_ops() {
  static std::once_flag once;
  std::call_once(once, [] {
    g_alias_c = std::make_unique<torch::Library>(torch::Library::FRAGMENT, "_C", std::nullopt, __FILE__, __LINE__);
    g_alias_c->def(kSchemaV1);
    g_alias_c->impl("paged_attention_v1", c10::DispatchKey::CUDA, TORCH_FN(paged_attention_v1));
    g_alias_c->def(kSchemaV2);
    g_alias_c->impl("paged_attention_v2", c10::DispatchKey::CUDA, TORCH_FN(paged_attention_v2));
    g_alias_cache =
        std::make_unique<torch::Library>(torch::Library::FRAGMENT, "_C_cache_ops", std::nullopt, __FILE__, __LINE__);
    g_alias_cache->def(kSchemaRC);
    g_alias_cache->impl("reshape_and_cache", c10::DispatchKey::CUDA, TORCH_FN(reshape_and_cache));
  });
}

int64_t abi_version() { return pc_abi_version(); }

}  // namespace

TORCH_LIBRARY(pagedevict_b200, m) {
  m.def(kSchemaV1);
  m.def(kSchemaV2);
  m.def(kSchemaRC);
  m.def("zero_cache_slots(Tensor! key_cache, Tensor! value_cache, Tensor slots) -> ()");
  m.def("prefill_attn_topleft_causal(Tensor! out, Tensor query, Tensor key, Tensor value, Tensor req_meta, "
        "int max_q_len, float scale) -> ()");
  m.def("alias_vllm_ops() -> ()", &alias_vllm_ops);
  m.def("abi_version() -> int", &abi_version);
}

TORCH_LIBRARY_IMPL(pagedevict_b200, CUDA, m) {
  m.impl("paged_attention_v1", TORCH_FN(paged_attention_v1));
  m.impl("paged_attention_v2", TORCH_FN(paged_attention_v2));
  m.impl("reshape_and_cache", TORCH_FN(reshape_and_cache));
  m.impl("zero_cache_slots", TORCH_FN(zero_cache_slots));
  m.impl("prefill_attn_topleft_causal", TORCH_FN(prefill_attn_topleft_causal));
}
