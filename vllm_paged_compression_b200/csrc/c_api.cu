// extern "C" entry points of libpagedevict_b200.so (see include/pagedevict_b200.h).
#include <stdio.h>
#include <string.h>

#include "pc_api_internal.h"

namespace pc {

static thread_local char g_err[512] = "";

int set_error(int code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int set_cuda_error(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return PC_ERR_CUDA;
}
int check_launch(const char* kernel_name) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error(kernel_name, e);
  return PC_OK;
}
int device_sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_cuda_error("cudaGetDevice", e); return -1; }
  if (dev != cached_dev) {
    e = cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) { set_cuda_error("cudaDeviceGetAttribute", e); return -1; }
    cached_dev = dev;
  }
  return cached;
}

int check_score_pool(const pc_score_pool* s, int num_kv_heads, const char* who) {
  if (s == nullptr) return PC_OK;
  char msg[160];
  if (s->n_pools < 1 || s->n_pools > PC_MAX_PEERS || s->local < 0 || s->local >= s->n_pools) {
    snprintf(msg, sizeof(msg), "%s: score pool needs 1 <= n_pools <= %d and 0 <= local < n_pools", who, PC_MAX_PEERS);
    return set_error(PC_ERR_INVALID, msg);
  }
  for (int i = 0; i < s->n_pools; ++i)
    if (s->pools[i] == nullptr) {
      snprintf(msg, sizeof(msg), "%s: score pool %d is a null pointer", who, i);
      return set_error(PC_ERR_INVALID, msg);
    }
  if (s->heads_total < 1 || s->head_offset < 0 || s->head_offset + num_kv_heads > s->heads_total) {
    snprintf(msg, sizeof(msg), "%s: score pool heads [%d, %d) do not fit heads_total %d", who, s->head_offset,
             s->head_offset + num_kv_heads, s->heads_total);
    return set_error(PC_ERR_INVALID, msg);
  }
  return PC_OK;
}

static int check_dtype(int dtype) {
  if (dtype != PC_BF16) return set_error(PC_ERR_UNSUPPORTED, "only bf16 KV cache / activations are built");
  return PC_OK;
}

}  // namespace pc

using namespace pc;

// extras pc_decode_layer adds to the attention launch: the fused cache write of the step's
// token and the device-side plan
struct FuseExtras {
  const void* key; const void* value; const long long* slot_mapping;
  long long key_stride, value_stride; const int* plan; int pdl;
};
static int decode_attn_impl(void* out, const void* query, const void* key_cache, const void* value_cache,
                            const int32_t* block_tables, int64_t table_stride, const int32_t* kv_lens,
                            int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                            int32_t block_size, int32_t max_kv_len, float scale, int64_t q_stride,
                            int64_t out_stride, int64_t kv_block_stride, const int32_t* seq_lens,
                            const pc_score_pool* score, void* workspace, size_t workspace_bytes,
                            int32_t dtype, void* stream, const FuseExtras& g_fuse);

// the library is compiled with -fvisibility=hidden: only the C ABI is exported
#pragma GCC visibility push(default)
extern "C" {

int pc_abi_version(void) { return PC_ABI_VERSION; }
size_t pc_sizeof_struct(int32_t which) {
  switch (which) {
    case 0: return sizeof(pc_evict_policy);
    case 1: return sizeof(pc_score_pool);
    case 2: return sizeof(pc_decode_step);
    default: return 0;
  }
}
const char* pc_last_error(void) { return g_err; }
int pc_device_sm_count(void) { return device_sm_count(); }

int pc_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                         const int64_t* slot_mapping, int32_t num_tokens, int32_t num_kv_heads,
                         int32_t head_size, int32_t block_size, int64_t key_stride,
                         int64_t value_stride, int64_t kv_block_stride, int32_t dtype,
                         void* stream) {
  if (int rc = check_dtype(dtype)) return rc;
  if (num_tokens == 0) return PC_OK;
  if (!key || !value || !key_cache || !value_cache || !slot_mapping)
    return set_error(PC_ERR_INVALID, "pc_reshape_and_cache: null pointer");
  CacheWriteArgs a{};
  a.key = key; a.value = value; a.k_cache = key_cache; a.v_cache = value_cache;
  a.slot_mapping = reinterpret_cast<const long long*>(slot_mapping);
  a.req_meta = nullptr; a.tile_offsets = nullptr; a.num_reqs = 1; a.total_tiles = 0;
  a.num_tokens = num_tokens; a.Hkv = num_kv_heads; a.D = head_size; a.BS = block_size;
  a.key_stride = key_stride; a.value_stride = value_stride; a.kv_block_stride = kv_block_stride;
  a.score = make_score_pool(nullptr);
  return cache_write(a, static_cast<cudaStream_t>(stream));
}

int pc_prune_prompt_and_cache(const void* key, const void* value, void* key_cache,
                              void* value_cache, const int64_t* slot_mapping,
                              const int32_t* req_meta, const int32_t* tile_offsets,
                              int32_t num_reqs, int32_t total_tiles, int32_t num_kv_heads,
                              int32_t head_size, int32_t block_size, int64_t key_stride,
                              int64_t value_stride, int64_t kv_block_stride,
                              const pc_score_pool* score, int32_t dtype, void* stream) {
  if (int rc = check_dtype(dtype)) return rc;
  if (num_reqs == 0 || total_tiles == 0) return PC_OK;
  if (!key || !value || !key_cache || !value_cache || !slot_mapping || !req_meta || !tile_offsets)
    return set_error(PC_ERR_INVALID, "pc_prune_prompt_and_cache: null pointer");
  if (int rc = check_score_pool(score, num_kv_heads, "pc_prune_prompt_and_cache")) return rc;
  CacheWriteArgs a{};
  a.key = key; a.value = value; a.k_cache = key_cache; a.v_cache = value_cache;
  a.slot_mapping = reinterpret_cast<const long long*>(slot_mapping);
  a.req_meta = req_meta; a.tile_offsets = tile_offsets; a.num_reqs = num_reqs;
  a.total_tiles = total_tiles; a.num_tokens = 0;
  a.Hkv = num_kv_heads; a.D = head_size; a.BS = block_size;
  a.key_stride = key_stride; a.value_stride = value_stride; a.kv_block_stride = kv_block_stride;
  a.score = make_score_pool(score);
  return cache_write(a, static_cast<cudaStream_t>(stream));
}

int pc_evict_select_compact(const int32_t* block_tables, int64_t table_stride,
                            const int32_t* n_blocks, const int32_t* seq_lens,
                            const int32_t* kv_lens, int32_t batch, const pc_evict_policy* policy,
                            const pc_score_pool* score, int32_t num_kv_heads, int32_t* out_tables,
                            int64_t out_stride, int32_t* out_kv_lens, int32_t* out_victim,
                            int32_t kv_lens_pre_adjusted, void* stream) {
  if (batch == 0) return PC_OK;
  if (!block_tables || !n_blocks || !seq_lens || !policy || !out_tables || !out_victim)
    return set_error(PC_ERR_INVALID, "pc_evict_select_compact: null pointer");
  if (out_kv_lens && !kv_lens)
    return set_error(PC_ERR_INVALID, "pc_evict_select_compact: out_kv_lens needs kv_lens");
  (void)num_kv_heads;
  if (int rc = check_score_pool(score, 0, "pc_evict_select_compact")) return rc;
  EvictSelectArgs a{};
  a.block_tables = block_tables; a.table_stride = table_stride; a.n_blocks = n_blocks;
  a.seq_lens = seq_lens; a.kv_lens = kv_lens; a.B = batch; a.pol = *policy;
  a.score = make_score_pool(score);
  a.out_tables = out_tables; a.out_stride = out_stride; a.out_kv_lens = out_kv_lens;
  a.out_victim = out_victim; a.pre_adjusted = kv_lens_pre_adjusted;
  return evict_select_compact(a, static_cast<cudaStream_t>(stream));
}

size_t pc_paged_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                       int32_t head_size, int32_t block_size,
                                       int32_t max_kv_len) {
  const int sms = device_sm_count();
  if (sms <= 0 || batch <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads) return 0;
  if ((head_size != 64 && head_size != 128) || (block_size != 8 && block_size != 16 && block_size != 32)) return 0;
  return decode_workspace_bytes(batch, num_heads, num_kv_heads, head_size, block_size, max_kv_len, sms);
}

int pc_paged_decode_plan(int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                         int32_t block_size, int32_t max_kv_len, int32_t num_sms, int32_t* out_plan) {
  if (!out_plan) return set_error(PC_ERR_INVALID, "pc_paged_decode_plan: null out_plan");
  out_plan[0] = out_plan[1] = 0;
  const int sms = num_sms > 0 ? num_sms : device_sm_count();
  if (sms <= 0) return PC_ERR_CUDA;
  if (batch <= 0 || num_kv_heads <= 0 || num_heads % num_kv_heads)
    return set_error(PC_ERR_INVALID, "pc_paged_decode_plan: bad shape");
  if ((head_size != 64 && head_size != 128) || (block_size != 8 && block_size != 16 && block_size != 32))
    return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size / block size (supported: head 64|128, block 8|16|32)");
  int pl[2];
  decode_plan(batch, num_heads, num_kv_heads, head_size, block_size, max_kv_len, sms, pl);
  out_plan[0] = pl[0];
  out_plan[1] = pl[1];
  return PC_OK;
}

int pc_paged_decode_attn_score(void* out, const void* query, const void* key_cache,
                               const void* value_cache, const int32_t* block_tables,
                               int64_t table_stride, const int32_t* kv_lens, int32_t batch,
                               int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                               int32_t block_size, int32_t max_kv_len, float scale,
                               int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                               const int32_t* seq_lens, const pc_score_pool* score,
                               void* workspace, size_t workspace_bytes, int32_t dtype,
                               void* stream) {
  const FuseExtras none = {nullptr, nullptr, nullptr, 0, 0, nullptr, 0};
  return decode_attn_impl(out, query, key_cache, value_cache, block_tables, table_stride, kv_lens, batch,
                          num_heads, num_kv_heads, head_size, block_size, max_kv_len, scale, q_stride,
                          out_stride, kv_block_stride, seq_lens, score, workspace, workspace_bytes, dtype,
                          stream, none);
}

int pc_paged_decode_attn_score_static(void* out, const void* query, const void* key_cache,
                                      const void* value_cache, const int32_t* block_tables,
                                      int64_t table_stride, const int32_t* kv_lens, int32_t batch,
                                      int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                                      int32_t block_size, int32_t max_kv_len, float scale,
                                      int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                                      const int32_t* seq_lens, const pc_score_pool* score,
                                      int32_t dtype, void* stream) {
  if (int rc = check_dtype(dtype)) return rc;
  if (batch == 0) return PC_OK;
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  if (int rc = decode_static_scratch(batch, num_heads, num_kv_heads, head_size, block_size, max_kv_len,
                                     static_cast<cudaStream_t>(stream), &scratch, &scratch_bytes))
    return rc;
  const FuseExtras none = {nullptr, nullptr, nullptr, 0, 0, nullptr, 0};
  return decode_attn_impl(out, query, key_cache, value_cache, block_tables, table_stride, kv_lens, batch,
                          num_heads, num_kv_heads, head_size, block_size, max_kv_len, scale, q_stride,
                          out_stride, kv_block_stride, seq_lens, score, scratch, scratch_bytes, dtype,
                          stream, none);
}
}  // extern "C"
#pragma GCC visibility pop

static int decode_attn_impl(void* out, const void* query, const void* key_cache, const void* value_cache,
                            const int32_t* block_tables, int64_t table_stride, const int32_t* kv_lens,
                            int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                            int32_t block_size, int32_t max_kv_len, float scale, int64_t q_stride,
                            int64_t out_stride, int64_t kv_block_stride, const int32_t* seq_lens,
                            const pc_score_pool* score, void* workspace, size_t workspace_bytes,
                            int32_t dtype, void* stream, const FuseExtras& g_fuse) {
  if (int rc = check_dtype(dtype)) return rc;
  if (batch == 0) return PC_OK;
  if (!out || !query || !key_cache || !value_cache || !block_tables || !kv_lens || !workspace)
    return set_error(PC_ERR_INVALID, "pc_paged_decode_attn_score: null pointer");
  if (score && !seq_lens)
    return set_error(PC_ERR_INVALID, "pc_paged_decode_attn_score: scoring needs seq_lens");
  if (int rc = check_score_pool(score, num_kv_heads, "pc_paged_decode_attn_score")) return rc;
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15)
    return set_error(PC_ERR_INVALID, "key/value cache must be 16-byte aligned (TMA bulk copy)");
  if ((kv_block_stride % 8) || (q_stride % 2) || (reinterpret_cast<uintptr_t>(query) & 3))
    return set_error(PC_ERR_INVALID, "kv_block_stride % 8 == 0, q_stride even, query 4-byte aligned required");
  if (reinterpret_cast<uintptr_t>(workspace) & 255)
    return set_error(PC_ERR_INVALID, "workspace must be 256-byte aligned");
  DecodeArgs a{};
  a.out = out; a.q = query; a.k_cache = key_cache; a.v_cache = value_cache;
  a.block_tables = block_tables; a.table_stride = table_stride; a.kv_lens = kv_lens;
  a.B = batch; a.Hq = num_heads; a.Hkv = num_kv_heads; a.D = head_size; a.BS = block_size;
  a.max_kv_len = max_kv_len; a.scale = scale;
  a.q_stride = q_stride; a.out_stride = out_stride; a.kv_block_stride = kv_block_stride;
  a.seq_lens = seq_lens; a.score = make_score_pool(score);
  a.workspace = workspace; a.workspace_bytes = workspace_bytes;
  a.new_k = g_fuse.key; a.new_v = g_fuse.value; a.slot_mapping = g_fuse.slot_mapping;
  a.new_k_stride = g_fuse.key_stride; a.new_v_stride = g_fuse.value_stride; a.plan = g_fuse.plan; a.pdl = g_fuse.pdl;
  return paged_decode_attn_score(a, static_cast<cudaStream_t>(stream));
}

#pragma GCC visibility push(default)
extern "C" {

int pc_evict_select_compact_layers(const int32_t* block_tables, int64_t table_stride,
                                   const int32_t* n_blocks, const int32_t* seq_lens, int32_t batch,
                                   int32_t num_layers, const pc_evict_policy* policy,
                                   const float* const* score_pools, int32_t heads_total,
                                   int32_t* out_tables, int64_t out_stride, int64_t out_layer_stride,
                                   int32_t* out_victims, void* stream) {
  if (batch == 0 || num_layers == 0) return PC_OK;
  if (!block_tables || !n_blocks || !seq_lens || !policy || !out_tables || !out_victims)
    return set_error(PC_ERR_INVALID, "pc_evict_select_compact_layers: null pointer");
  if (out_layer_stride < (int64_t)batch * out_stride)
    return set_error(PC_ERR_INVALID, "pc_evict_select_compact_layers: out_layer_stride < batch * out_stride");
  EvictLayersArgs a{};
  a.block_tables = block_tables; a.table_stride = table_stride; a.n_blocks = n_blocks; a.seq_lens = seq_lens;
  a.B = batch; a.L = num_layers; a.pol = *policy; a.pools = score_pools; a.heads_total = heads_total;
  a.out_tables = out_tables; a.out_stride = out_stride; a.out_layer_stride = out_layer_stride;
  a.out_victims = out_victims;
  return evict_select_compact_layers(a, static_cast<cudaStream_t>(stream));
}

int pc_decode_layer(const pc_decode_step* st, void* out, const void* query, const void* key,
                    const void* value, void* key_cache, void* value_cache,
                    const pc_score_pool* score, int32_t* out_victim, const int32_t* attend_tables,
                    int64_t attend_stride) {
  if (!st) return set_error(PC_ERR_INVALID, "pc_decode_layer: null step");
  if (st->batch == 0) return PC_OK;
  // W1: this step's K/V rows into the pool -- inside the attention kernel when fused
  const bool fuse = st->fuse_cache_write && !(st->any_trigger && st->attend_uncompacted);
  if (!fuse)
    if (int rc = pc_reshape_and_cache(key, value, key_cache, value_cache, st->slot_mapping, st->batch,
                                      st->num_kv_heads, st->head_size, st->block_size, st->key_stride,
                                      st->value_stride, st->kv_block_stride, st->dtype, st->stream))
      return rc;
  const int32_t* tables = st->block_tables;
  int64_t stride = st->table_stride;
  const bool preselected = attend_tables != nullptr;   // pc_evict_select_compact_layers ran for this step
  if (preselected) {
    if (!fuse) return set_error(PC_ERR_INVALID, "pc_decode_layer: attend_tables needs the fused cache write");
    tables = attend_tables;
    stride = attend_stride;
  } else if (st->any_trigger) {
    // P3 + S4 + C1: victims from the scores of blocks filled on EARLIER steps, compacted tables
    if (!out_victim || !st->compact_tables || !st->n_blocks)
      return set_error(PC_ERR_INVALID, "pc_decode_layer: trigger step needs out_victim, compact_tables, n_blocks");
    if (int rc = pc_evict_select_compact(st->block_tables, st->table_stride, st->n_blocks, st->seq_lens,
                                         nullptr, st->batch, &st->policy, score, st->num_kv_heads,
                                         st->compact_tables, st->compact_stride, nullptr, out_victim, 1,
                                         st->stream))
      return rc;
    if (!st->attend_uncompacted) { tables = st->compact_tables; stride = st->compact_stride; }
  }
  // A2 + S2: attention over the (compacted) tables, scoring the blocks that fill this step
  // PDL only when the preceding kernel is not this call's own evict_select / cache write
  FuseExtras g_fuse = {nullptr, nullptr, nullptr, 0, 0, st->plan,
                       (st->pdl && fuse && (preselected || !st->any_trigger)) ? 1 : 0};
  if (fuse) {
    if (!key || !value || !st->slot_mapping)
      return set_error(PC_ERR_INVALID, "pc_decode_layer: fused cache write needs key, value, slot_mapping");
    g_fuse.key = key; g_fuse.value = value;
    g_fuse.slot_mapping = reinterpret_cast<const long long*>(st->slot_mapping);
    g_fuse.key_stride = st->key_stride; g_fuse.value_stride = st->value_stride;
  }
  return decode_attn_impl(out, query, key_cache, value_cache, tables, stride, st->kv_lens,
                                            st->batch, st->num_heads, st->num_kv_heads, st->head_size,
                                            st->block_size, st->max_kv_len, st->scale, st->q_stride,
                                            st->out_stride, st->kv_block_stride,
                          score ? st->seq_lens : nullptr, score, st->workspace,
                          st->workspace_bytes, st->dtype, st->stream, g_fuse);
}

int pc_prefill_attn_topleft_causal(void* out, const void* query, const void* key,
                                   const void* value, const int32_t* req_meta,
                                   int32_t num_reqs, int32_t max_q_len, int64_t num_tokens,
                                   int32_t num_heads,
                                   int32_t num_kv_heads, int32_t head_size, float scale,
                                   int64_t q_stride, int64_t out_stride, int64_t key_stride,
                                   int64_t value_stride, const int32_t* req_meta_host, int32_t dtype,
                                   void* stream) {
  if (int rc = check_dtype(dtype)) return rc;
  if (num_reqs == 0) return PC_OK;
  if (!out || !query || !key || !value || !req_meta)
    return set_error(PC_ERR_INVALID, "pc_prefill_attn_topleft_causal: null pointer");
  PrefillArgs a{};
  a.out = out; a.q = query; a.k = key; a.v = value; a.req_meta = req_meta;
  a.num_reqs = num_reqs; a.max_q_len = max_q_len; a.num_tokens = num_tokens; a.Hq = num_heads; a.Hkv = num_kv_heads;
  a.D = head_size; a.scale = scale; a.q_stride = q_stride; a.out_stride = out_stride;
  a.key_stride = key_stride; a.value_stride = value_stride;
  a.req_meta_host = req_meta_host;
  return prefill_attn(a, static_cast<cudaStream_t>(stream));
}

int pc_copy_blocks(void* key_cache, void* value_cache, const int64_t* pairs, int32_t num_pairs,
                   int64_t block_bytes, void* stream) {
  if (num_pairs == 0) return PC_OK;
  if (!key_cache || !value_cache || !pairs)
    return set_error(PC_ERR_INVALID, "pc_copy_blocks: null pointer");
  return copy_blocks(key_cache, value_cache, reinterpret_cast<const long long*>(pairs), num_pairs,
                     block_bytes, static_cast<cudaStream_t>(stream));
}

size_t pc_prompt_token_prune_workspace_bytes(int64_t total_mid, int32_t num_kv_heads) {
  return token_prune_workspace_bytes(total_mid, num_kv_heads);
}

int pc_prompt_token_prune(const void* key, const void* value, void* out_key, void* out_value,
                          const int32_t* req_meta, const int32_t* mid_offsets, const int32_t* tok_offsets,
                          int32_t num_reqs, int64_t total_tokens, int64_t total_mid, int32_t num_kv_heads,
                          int32_t head_size, int32_t score_kind, int32_t head_mixing_repack,
                          int64_t key_stride, int64_t value_stride, int64_t out_key_stride,
                          int64_t out_value_stride, void* workspace, size_t workspace_bytes, int32_t dtype,
                          void* stream) {
  if (int rc = check_dtype(dtype)) return rc;
  if (num_reqs == 0 || total_tokens == 0) return PC_OK;
  if (!key || !value || !out_key || !out_value || !req_meta || !mid_offsets || !tok_offsets || !workspace)
    return set_error(PC_ERR_INVALID, "pc_prompt_token_prune: null pointer");
  if (score_kind != 0 && score_kind != 1)
    return set_error(PC_ERR_INVALID, "pc_prompt_token_prune: score_kind must be 0 (value_l2) or 1 (inverse_key_l2)");
  TokenPruneArgs a{};
  a.key = key; a.value = value; a.out_key = out_key; a.out_value = out_value;
  a.req_meta = req_meta; a.mid_offsets = mid_offsets; a.tok_offsets = tok_offsets;
  a.num_reqs = num_reqs; a.total_tokens = total_tokens; a.total_mid = total_mid;
  a.Hkv = num_kv_heads; a.D = head_size; a.score_kind = score_kind; a.repack = head_mixing_repack;
  a.key_stride = key_stride; a.value_stride = value_stride;
  a.out_key_stride = out_key_stride; a.out_value_stride = out_value_stride;
  a.workspace = workspace; a.workspace_bytes = workspace_bytes;
  return prompt_token_prune(a, static_cast<cudaStream_t>(stream));
}

int pc_zero_cache_slots(void* key_cache, void* value_cache, const int64_t* slots, int32_t num_slots,
                        int32_t num_kv_heads, int32_t head_size, int32_t block_size,
                        int64_t kv_block_stride, void* stream) {
  if (num_slots == 0) return PC_OK;
  if (!key_cache || !value_cache || !slots) return set_error(PC_ERR_INVALID, "pc_zero_cache_slots: null pointer");
  ZeroSlotsArgs a{};
  a.k_caches = nullptr; a.v_caches = nullptr; a.k0 = key_cache; a.v0 = value_cache;
  a.slots = reinterpret_cast<const long long*>(slots); a.slot_stride = num_slots;
  a.num_slots = num_slots; a.num_layers = 1; a.Hkv = num_kv_heads; a.D = head_size; a.BS = block_size;
  a.kv_block_stride = kv_block_stride;
  return zero_cache_slots(a, static_cast<cudaStream_t>(stream));
}

int pc_zero_cache_slots_layers(void* const* key_caches, void* const* value_caches, const int64_t* slots,
                               int64_t slot_stride, int32_t num_slots, int32_t num_layers,
                               int32_t num_kv_heads, int32_t head_size, int32_t block_size,
                               int64_t kv_block_stride, void* stream) {
  if (num_slots == 0 || num_layers == 0) return PC_OK;
  if (!key_caches || !value_caches || !slots)
    return set_error(PC_ERR_INVALID, "pc_zero_cache_slots_layers: null pointer");
  if (slot_stride < num_slots) return set_error(PC_ERR_INVALID, "pc_zero_cache_slots_layers: slot_stride < num_slots");
  ZeroSlotsArgs a{};
  a.k_caches = key_caches; a.v_caches = value_caches; a.k0 = nullptr; a.v0 = nullptr;
  a.slots = reinterpret_cast<const long long*>(slots); a.slot_stride = slot_stride;
  a.num_slots = num_slots; a.num_layers = num_layers; a.Hkv = num_kv_heads; a.D = head_size; a.BS = block_size;
  a.kv_block_stride = kv_block_stride;
  return zero_cache_slots(a, static_cast<cudaStream_t>(stream));
}

int pc_swap_blocks(const void* src, void* dst, const int64_t* pairs, int32_t num_pairs,
                   int64_t block_bytes, void* stream) {
  if (num_pairs == 0) return PC_OK;
  if (!src || !dst || !pairs) return set_error(PC_ERR_INVALID, "pc_swap_blocks: null pointer");
  // one cudaMemcpyAsync per RUN of pairs whose source and destination blocks are both consecutive
  // (csrc/cache_kernels.cu:24-66 issues one per pair: a 32 KiB block costs ~4 us per call = 8 GB/s;
  // swap-in / swap-out lists of a sequence are mostly runs).  Direction inferred by UVA.
  for (int i = 0; i < num_pairs;) {
    int n = 1;
    while (i + n < num_pairs && pairs[2 * (i + n)] == pairs[2 * i] + n && pairs[2 * (i + n) + 1] == pairs[2 * i + 1] + n) ++n;
    const char* s = static_cast<const char*>(src) + pairs[2 * i] * block_bytes;
    char* d = static_cast<char*>(dst) + pairs[2 * i + 1] * block_bytes;
    const cudaError_t e = cudaMemcpyAsync(d, s, (size_t)block_bytes * n, cudaMemcpyDefault,
                                          static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return set_cuda_error("pc_swap_blocks cudaMemcpyAsync", e);
    i += n;
  }
  return PC_OK;
}

#ifdef PC_DEBUG
/* Debug hooks, only in the PC_DEBUG build (libpagedevict_b200_debug.so, `make debug`); the
 * product library neither exports them nor reads any environment variable.
 * decode trace: per-warp globaltimer stamps, [warps][8]. */
void pc_debug_set_decode_trace(void* device_buf) { pc::set_decode_trace(device_buf); }
void pc_debug_set_decode_plan(int32_t chunk_pages, int32_t team) { pc::set_decode_plan_override(chunk_pages, team); }
void pc_debug_set_prefill_tc_tile(int32_t tn) { pc::set_prefill_tc_tile(tn); }
void pc_debug_set_prefill_tc_persistent(int32_t v) { pc::set_prefill_tc_persistent(v); }
void pc_debug_set_cache_write_block_heads(int32_t hc) { pc::set_cache_write_block_heads(hc); }
void pc_debug_set_prefill_pp(int32_t v) { pc::set_prefill_pp(v); }
void pc_debug_set_prefill_dq(int32_t v) { pc::set_prefill_dq(v); }
int32_t pc_debug_prefill_dq_trace(int32_t on, void* host_out, size_t bytes) { return pc::prefill_dq_trace(on, host_out, bytes); }
void pc_debug_set_prefill_tc_xitem(int32_t v) { pc::set_prefill_tc_xitem(v); }
void pc_debug_set_decode_two_cta(int32_t v) { pc::set_decode_two_cta(v); }
#endif

}  // extern "C"
#pragma GCC visibility pop
