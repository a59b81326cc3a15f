// TEST INFRASTRUCTURE ONLY (oracle/): binding shim for the reference's own CPU
// kernels. Nothing under vllm_paged_compression_b200/ may load the resulting
// library; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs do.
//
// The reference sources are compiled where they lie
// (/root/reference/csrc/cpu/{attention,cache}.cpp, never copied); this file only
// declares their entry points (csrc/cpu/torch_bindings.cpp:8-25 declares the same
// prototypes) and registers them under a private op namespace, because the
// reference's own torch_bindings.cpp also registers ~20 unrelated ops whose sources
// we do not build.
#include <torch/all.h>
#include <torch/library.h>

#include <string>

// csrc/cpu/attention.cpp:458-476
void paged_attention_v1(
    torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
    int64_t max_seq_len, const c10::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, double k_scale, double v_scale,
    const int64_t tp_rank, const int64_t blocksparse_local_blocks,
    const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
    const int64_t blocksparse_head_sliding_step);

// csrc/cpu/attention.cpp:773-797
void paged_attention_v2(
    torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
    torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int64_t num_kv_heads, double scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
    int64_t max_seq_len, const c10::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, double k_scale, double v_scale,
    const int64_t tp_rank, const int64_t blocksparse_local_blocks,
    const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
    const int64_t blocksparse_head_sliding_step);

// csrc/cpu/cache.cpp:108-136
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value,
                       torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, double k_scale,
                       double v_scale);

TORCH_LIBRARY(pc_ref_cpu, m) {
  m.def(
      "paged_attention_v1(Tensor! out, Tensor query, Tensor key_cache, Tensor "
      "value_cache, int num_kv_heads, float scale, Tensor block_tables, Tensor "
      "seq_lens, int block_size, int max_seq_len, Tensor? alibi_slopes, str "
      "kv_cache_dtype, float k_scale, float v_scale, int tp_rank, int "
      "blocksparse_local_blocks, int blocksparse_vert_stride, int "
      "blocksparse_block_size, int blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v1", torch::kCPU, &paged_attention_v1);
  m.def(
      "paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, "
      "Tensor! tmp_out, Tensor query, Tensor key_cache, Tensor value_cache, int "
      "num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int "
      "block_size, int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, "
      "float k_scale, float v_scale, int tp_rank, int blocksparse_local_blocks, "
      "int blocksparse_vert_stride, int blocksparse_block_size, int "
      "blocksparse_head_sliding_step) -> ()");
  m.impl("paged_attention_v2", torch::kCPU, &paged_attention_v2);
  m.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! "
      "value_cache, Tensor slot_mapping, str kv_cache_dtype, float k_scale, float "
      "v_scale) -> ()");
  m.impl("reshape_and_cache", torch::kCPU, &reshape_and_cache);
}
