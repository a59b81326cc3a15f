// Internal (C++) side of the C ABI: argument structs + error plumbing shared by the
// kernel translation units and c_api.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/pagedevict_b200.h"

namespace pc {

// Device-side view of pc_score_pool (passed by value in kernel params).
struct ScorePool {
  float* pools[PC_MAX_PEERS];
  int n_pools;      // 0 = scoring off
  int local;
  int heads_total;
  int head_offset;
};
inline ScorePool make_score_pool(const pc_score_pool* s) {
  ScorePool sp;
  for (int i = 0; i < PC_MAX_PEERS; ++i) sp.pools[i] = nullptr;
  sp.n_pools = 0; sp.local = 0; sp.heads_total = 0; sp.head_offset = 0;
  if (s != nullptr) {
    for (int i = 0; i < s->n_pools && i < PC_MAX_PEERS; ++i) sp.pools[i] = s->pools[i];
    sp.n_pools = s->n_pools; sp.local = s->local; sp.heads_total = s->heads_total;
    sp.head_offset = s->head_offset;
  }
  return sp;
}
int check_score_pool(const pc_score_pool* s, int num_kv_heads, const char* who);

int set_error(int code, const char* msg);
int set_cuda_error(const char* what, cudaError_t e);
int check_launch(const char* kernel_name);
int device_sm_count();

struct DecodeArgs {
  void* out;
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  long long table_stride;
  const int* kv_lens;
  int B, Hq, Hkv, D, BS, max_kv_len;
  float scale;
  long long q_stride, out_stride, kv_block_stride;
  const int* seq_lens;
  ScorePool score;
  void* workspace;
  size_t workspace_bytes;
  // fused reshape_and_cache of the step's token (nullptr = off)
  const void* new_k;
  const void* new_v;
  const long long* slot_mapping;
  long long new_k_stride, new_v_stride;
  const int* plan;  // device {chunk_pages, team} or nullptr (host plans from max_kv_len)
  int pdl;          // programmatic dependent launch (see pc_decode_step.pdl)
};
size_t decode_workspace_bytes(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms);
void decode_plan(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms, int* out2);  // {chunk_pages, team}
int paged_decode_attn_score(const DecodeArgs& a, cudaStream_t stream);
// static-scratch variant (pc_paged_decode_attn_score_static): device address + size of the file-scope
// workspace of the current device, size check for this launch, counters / tickets cleared on `stream`
int decode_static_scratch(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, cudaStream_t stream,
                          void** out_ptr, size_t* out_bytes);
#ifdef PC_DEBUG
void set_decode_trace(void* buf);  // per-warp globaltimer stamps, nullptr = off
void set_decode_plan_override(int chunk_pages, int team);  // 0 = planner's choice
void set_decode_two_cta(int v);    // (128, 16): two 4-warp CTAs per SM instead of one 8-warp CTA
#endif

struct CacheWriteArgs {
  const void* key;
  const void* value;
  void* k_cache;
  void* v_cache;
  const long long* slot_mapping;
  const int* req_meta;      // [R][5] or nullptr (plain reshape_and_cache)
  const int* tile_offsets;  // [R+1] or nullptr
  int num_reqs, total_tiles, num_tokens;
  int Hkv, D, BS;
  long long key_stride, value_stride, kv_block_stride;
  ScorePool score;
};
int cache_write(const CacheWriteArgs& a, cudaStream_t stream);

struct EvictSelectArgs {
  const int* block_tables;
  long long table_stride;
  const int* n_blocks;
  const int* seq_lens;
  const int* kv_lens;
  int B;
  pc_evict_policy pol;
  ScorePool score;
  int* out_tables;
  long long out_stride;
  int* out_kv_lens;
  int* out_victim;
  int pre_adjusted;
};
int evict_select_compact(const EvictSelectArgs& a, cudaStream_t stream);

struct EvictLayersArgs {
  const int* block_tables;
  long long table_stride;
  const int* n_blocks;
  const int* seq_lens;
  int B, L;
  pc_evict_policy pol;
  const float* const* pools;   // DEVICE array [L] of this rank's score pools, or nullptr (streamingLLM)
  int heads_total;
  int* out_tables;
  long long out_stride, out_layer_stride;
  int* out_victims;             // decode_method 2: IN (the host's choice per triggered row) and OUT
};
int evict_select_compact_layers(const EvictLayersArgs& a, cudaStream_t stream);

struct PrefillArgs {
  void* out;
  const void* q;
  const void* k;
  const void* v;
  const int* req_meta;  // [R][6]
  int num_reqs, max_q_len, Hq, Hkv, D;
  float scale;
  long long q_stride, out_stride, key_stride, value_stride;
  long long num_tokens;  // rows of q / k / v (bounds of the TMA tensor maps)
  const int* req_meta_host;  // HOST copy of req_meta for argument validation, or nullptr
};
int prefill_attn(const PrefillArgs& a, cudaStream_t stream);     // tcgen05/TMEM kernels (prefill_attn_tc.cu: checks + dispatch)
bool prefill_attn_dq_supported(const PrefillArgs& a);                        // dual-query-tile kernel (prefill_attn_dq.cu):
int prefill_attn_dq(const PrefillArgs& a, bool box8, cudaStream_t stream);   // prefill_attn() sends it D 128, prompts <= kDqMaxQLen
constexpr int kDqMaxQLen = 6144;
#ifdef PC_DEBUG
bool prefill_attn_pp_supported(const PrefillArgs& a);            // two-query-tile ping-pong kernel (prefill_attn_pp.cu),
int prefill_attn_pp(const PrefillArgs& a, cudaStream_t stream);  // measured, not adopted: PC_DEBUG library only
#endif
#ifdef PC_DEBUG
void set_prefill_tc_persistent(int v);  // 1 = persistent CTAs, 0 = one CTA per query tile, -1 = rule
void set_prefill_tc_tile(int tn);       // kv tile width of the tcgen05 kernel (64 | 128)
int prefill_dq_trace(int on, void* host_out, size_t bytes);  // clock stamps of CTA 0 of the dual-query-tile kernel
void set_prefill_dq(int v);             // dual-query-tile kernel: 1 = always (when supported), 0 = never, -1 = rule
void set_prefill_pp(int v);             // ping-pong kernel: 1 = always (when supported), 0 = never, -1 = rule
void set_prefill_tc_xitem(int v);       // 1 = the next item's first S MMA is issued before the current item's last P.V (product), 0 = after
void set_cache_write_block_heads(int hc);  // heads per CTA of cache_write_block_kernel, 0 = rule
#endif

struct TokenPruneArgs {
  const void* key;
  const void* value;
  void* out_key;
  void* out_value;
  const int* req_meta;      // DEVICE [R][6] {src_start, q_len, first, mid, k_drop, dst_start}
  const int* mid_offsets;   // DEVICE [R+1]
  const int* tok_offsets;   // DEVICE [R+1]
  int num_reqs;
  long long total_tokens, total_mid;
  int Hkv, D, score_kind, repack;
  long long key_stride, value_stride, out_key_stride, out_value_stride;
  void* workspace;
  size_t workspace_bytes;
};
size_t token_prune_workspace_bytes(long long total_mid, int Hkv);
int prompt_token_prune(const TokenPruneArgs& a, cudaStream_t stream);

struct ZeroSlotsArgs {
  void* const* k_caches;   // DEVICE [num_layers] pool pointers, or nullptr: k0 / v0 is the one layer
  void* const* v_caches;
  void* k0;
  void* v0;
  const long long* slots;  // DEVICE [num_layers][slot_stride]
  long long slot_stride;
  int num_slots, num_layers, Hkv, D, BS;
  long long kv_block_stride;
};
int zero_cache_slots(const ZeroSlotsArgs& a, cudaStream_t stream);

int copy_blocks(void* k_cache, void* v_cache, const long long* pairs, int num_pairs,
                long long block_bytes, cudaStream_t stream);

}  // namespace pc
