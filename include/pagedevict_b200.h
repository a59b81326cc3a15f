/*
 * pagedevict_b200.h -- C ABI of libpagedevict_b200.so: the B200 (sm_100a) native
 * implementation of the PagedCompression / "paged eviction" hot path of
 * krishnateja95/vllm-Paged-Compression.
 *
 * This is the drop-in boundary.  The reference binds this path through a torch C++
 * extension (TORCH_LIBRARY ops called from vllm/_custom_ops.py); each entry point
 * below states the reference interface it replaces.  Plain pointers and sizes only:
 * every pointer is a CUDA DEVICE pointer unless marked HOST; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  All calls are
 * asynchronous on `stream`.  No call allocates or frees device memory, keeps a
 * reference to caller memory after returning, or synchronises the device.
 *
 * Error convention: every function returns 0 on success or a negative pc_status; the
 * message is available from pc_last_error() (thread-local).  This mirrors the
 * reference's TORCH_CHECK(false, "...") -> RuntimeError convention
 * (csrc/attention/paged_attention_v1.cu:108-110,160-162); the Python host raises
 * RuntimeError(pc_last_error()).
 *
 * Layouts (vllm/attention/ops/paged_attn.py:40-62), element type bf16:
 *   key_cache   [num_blocks][num_kv_heads][head_size/8][block_size][8]
 *   value_cache [num_blocks][num_kv_heads][head_size][block_size]
 * kv_block_stride = elements between consecutive blocks (num_kv_heads*head_size*block_size
 * for a dense pool).  Supported: head_size 64|128, block_size 8|16|32, dtype bf16.
 */
#ifndef PAGEDEVICT_B200_H_
#define PAGEDEVICT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PC_ABI_VERSION 6
#define PC_MAX_PEERS 8

typedef enum pc_status {
  PC_OK = 0,
  PC_ERR_INVALID = -1,     /* bad argument (shape, alignment, null pointer) */
  PC_ERR_UNSUPPORTED = -2, /* head size / block size / dtype not built */
  PC_ERR_CUDA = -3         /* CUDA runtime error at launch */
} pc_status;

typedef enum pc_dtype { PC_BF16 = 0 } pc_dtype;

/* Eviction policy = vllm/config.py:933-991 PagedEvictConfig (budget mode, HEAD) plus
 * the README-era percentage / evict_freq fields (benchmarks/llama3.1_8b_test.txt:1). */
typedef struct pc_evict_policy {
  int32_t prune_type;      /* 0 = budget, 1 = percentage (evict_freq) */
  int32_t block_size;
  int32_t cache_budget;    /* budget mode */
  int32_t evict_freq;      /* percentage mode: one block per evict_freq blocks */
  int32_t initial_blocks;  /* sink blocks, reference hard-codes 1 */
  int32_t num_block_merge; /* percentage mode lower bound, reference default 2 */
  int32_t decode_method;   /* 0 = streamingLLM (victim = block 1), 1 = value_l2 argmin,
                            * 2 = victim chosen by the HOST (HEAD as shipped: enable_random_evict = True,
                            *     torch_cuda_sdpa.py:802 -- a random.randint per triggered row, :1504-1513):
                            *     the caller pre-fills out_victim / out_victims, the kernel applies the
                            *     trigger test and compacts; an index outside [1, n_blocks) evicts nothing */
  int32_t topk_blocks;     /* <0: global window [1, n-1); >0: local window [1, 1+topk) */
} pc_evict_policy;

/* value_l2 block scores.  Replaces the per-request python lists of BlockL2NormManager
 * (vllm/attention/block_l2norm_manager.py:8-106).  One fp32 POOL per layer, indexed by
 * PHYSICAL block id like the KV pool itself: pool[block_id][kv_head] = sum over the
 * block's tokens of ||V[t, head, :]||_2 (fixed fp32 order, csrc/score_common.cuh).  A
 * block is scored once, when it fills; evicting a block needs no score bookkeeping (the
 * table entry goes, the pool row is simply rewritten when the block is reused and fills
 * again), so there is nothing to compact and nothing keyed by request or batch row.
 * Tensor parallelism over KV heads: every rank holds a pool with ALL heads
 * (heads_total columns); a rank computes its own heads [head_offset, head_offset +
 * num_kv_heads) and the scoring kernels store each partial straight into every rank's
 * pool through peer memory (pools[0..n_pools), NVLink P2P; n_pools == 1: no peers).
 * Victim selection then reads the local pool (pools[local]) and sums heads in ascending
 * global order, so every rank picks the same victim, bit-identical to TP = 1. */
typedef struct pc_score_pool {
  float* pools[PC_MAX_PEERS]; /* device pointers, [num_blocks][heads_total] each */
  int32_t n_pools;            /* pools to store into (>= 1) */
  int32_t local;              /* index of this rank's own pool in pools[] */
  int32_t heads_total;        /* kv heads over all ranks */
  int32_t head_offset;        /* first global kv head computed by this rank */
} pc_score_pool;

int pc_abi_version(void);
/* sizeof() of the structs above as this library was compiled (binding self-check). */
size_t pc_sizeof_struct(int32_t which /* 0 pc_evict_policy, 1 pc_score_pool, 2 pc_decode_step */);
const char* pc_last_error(void);
/* SM count of the current device (grid sizing is derived from it). */
int pc_device_sm_count(void);

/* ---- W1: reshape_and_cache --------------------------------------------------------
 * Replaces _C_cache_ops::reshape_and_cache (csrc/torch_bindings.cpp:444-450,
 * csrc/cache_kernels.cu:152-205,263-287; python vllm/_custom_ops.py:984-996).
 * key/value [num_tokens][num_kv_heads][head_size] with row strides key_stride /
 * value_stride (elements, multiples of 8, 16-byte aligned rows).  slot_mapping int64
 * [num_tokens]; slot < 0 = skip (PAD_SLOT_ID).  Bit-exact copy. */
int pc_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                         const int64_t* slot_mapping, int32_t num_tokens, int32_t num_kv_heads,
                         int32_t head_size, int32_t block_size, int64_t key_stride,
                         int64_t value_stride, int64_t kv_block_stride, int32_t dtype,
                         void* stream);

/* ---- R1 + W1 + S3: prompt prune, cache write and prompt block scoring in one pass ---
 * Replaces _prune_prompts_keys_and_values_new + prune_prompt_inplace + _update_kvcache
 * (vllm/attention/backends/torch_cuda_sdpa.py:1222-1296,1129-1171;
 * vllm/attention/kvcache_prunner.py:167-218) and the prompt block scoring loop
 * (torch_cuda_sdpa.py:1278-1288).  No intermediate pruned K/V tensor is materialised.
 * req_meta int32 [num_reqs][5] = {src_start, kept, first, pruned, dst_start}: request r
 * keeps source rows src_start + (j < first ? j : j + pruned), j in [0, kept), written
 * at slot_mapping[dst_start + j].  tile_offsets int32 [num_reqs+1] = prefix sum of
 * ceil(kept/block_size).  score (HOST struct) may be NULL (no scoring); otherwise every
 * FULL kept block gets pool[block_id][head_offset + h] (fp32 per-head value_l2 partial,
 * fixed reduction order, see csrc/score_common.cuh) in every pool of *score. */
int pc_prune_prompt_and_cache(const void* key, const void* value, void* key_cache,
                              void* value_cache, const int64_t* slot_mapping,
                              const int32_t* req_meta, const int32_t* tile_offsets,
                              int32_t num_reqs, int32_t total_tiles, int32_t num_kv_heads,
                              int32_t head_size, int32_t block_size, int64_t key_stride,
                              int64_t value_stride, int64_t kv_block_stride,
                              const pc_score_pool* score /* HOST */, int32_t dtype, void* stream);

/* ---- P3 + S4 + C1 + C2: decode trigger, victim argmin, table/score compaction -------
 * Replaces _prune_decode_reqs_streamingLLM / _prune_decode_reqs_topk /
 * _create_updated_block_tables (torch_cuda_sdpa.py:1375-1433,1478-1586) and
 * BlockL2NormManager.get_updated_block_tables (block_l2norm_manager.py:95-106).
 * Row r owns n_blocks[r] table entries.  Triggered rows (policy on seq_lens[r] and
 * n_blocks[r]*block_size) lose entry out_victim[r]: out_tables row = input row with that
 * column deleted, shifted left, 0-padded to out_stride.  Other rows are copied;
 * out_victim = -1.  out_kv_lens[r] = kv_lens[r] - (triggered && !kv_lens_pre_adjusted ?
 * block_size : 0).  value_l2: score of table entry i = (sum over all heads_total heads,
 * ascending, of pool[table[i]][h]) / heads_total, read from score->pools[score->local];
 * ties in the argmin go to the lowest index (strict <, torch_cuda_sdpa.py:1536-1541).
 * decode_method 2: out_victim is IN/OUT -- on entry the host's choice for every row.
 * num_kv_heads is unused when score is given (kept for call-site symmetry). */
int pc_evict_select_compact(const int32_t* block_tables, int64_t table_stride,
                            const int32_t* n_blocks, const int32_t* seq_lens,
                            const int32_t* kv_lens, int32_t batch,
                            const pc_evict_policy* policy /* HOST */,
                            const pc_score_pool* score /* HOST, NULL for streamingLLM */,
                            int32_t num_kv_heads, int32_t* out_tables, int64_t out_stride,
                            int32_t* out_kv_lens, int32_t* out_victim,
                            int32_t kv_lens_pre_adjusted, void* stream);

/* ---- A2 + S2: paged decode attention with fused last-block scoring ------------------
 * Replaces _C::paged_attention_v1 / paged_attention_v2 (csrc/torch_bindings.cpp:27-51,
 * csrc/attention/attention_kernels.cuh:86-668, python vllm/_custom_ops.py:70-130,
 * vllm/attention/ops/paged_attn.py:87-191) and the last-block scoring read of
 * torch_cuda_sdpa.py:1518-1526.  query/out [batch][num_heads][head_size] with row
 * strides q_stride / out_stride (elements).  kv_lens = physical KV lengths.  If
 * score != NULL and seq_lens != NULL, rows with seq_lens[b] % block_size == 0 (and a
 * full last block) get pool[table[b][ceil(kv_len/bs)-1]][head_offset + h] for every
 * local kv head h.
 * workspace: pc_paged_decode_workspace_bytes(max_kv_len) bytes -- enough for every plan the
 * planner produces for a longest row of up to max_kv_len tokens -- 256-byte aligned.  It must be
 * ZERO-FILLED by the caller before its first use; every call leaves it ready for the
 * next one (arrival tickets are reset in-kernel), so one workspace per stream is
 * allocated once and reused.  A single kernel launch, no memset. */
size_t pc_paged_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                       int32_t head_size, int32_t block_size,
                                       int32_t max_kv_len);
int pc_paged_decode_attn_score(void* out, const void* query, const void* key_cache,
                               const void* value_cache, const int32_t* block_tables,
                               int64_t table_stride, const int32_t* kv_lens, int32_t batch,
                               int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                               int32_t block_size, int32_t max_kv_len, float scale,
                               int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                               const int32_t* seq_lens, const pc_score_pool* score /* HOST */,
                               void* workspace, size_t workspace_bytes, int32_t dtype,
                               void* stream);

/* ---- pc_paged_decode_attn_score without a caller workspace ------------------------------
 * No counterpart in the reference: added for callers that
 * cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same arguments, semantics and results as pc_paged_decode_attn_score;
 * the split-KV workspace is a file-scope device array inside the library of
 * PC_DECODE_STATIC_SCRATCH_BYTES (16 MiB) that every GPU which loads the library holds until
 * exit.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine):
 * one cudaMemsetAsync of the counters + this launch's arrival tickets and one kernel launch on
 * `stream`.  A launch whose plan needs more scratch than the array holds (very large batch x
 * context products: > ~3000 split chunks at head size 128) is rejected with PC_ERR_INVALID and a
 * message naming both sizes -- use pc_paged_decode_attn_score with a caller workspace then.
 * EVERY call on a device shares the one array: calls that can overlap on the GPU (different
 * streams, or graph branches without a dependency) race on it and must be serialised by the
 * caller; calls on one stream are safe. */
#define PC_DECODE_STATIC_SCRATCH_BYTES (16u << 20)
int pc_paged_decode_attn_score_static(void* out, const void* query, const void* key_cache,
                                      const void* value_cache, const int32_t* block_tables,
                                      int64_t table_stride, const int32_t* kv_lens, int32_t batch,
                                      int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                                      int32_t block_size, int32_t max_kv_len, float scale,
                                      int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                                      const int32_t* seq_lens, const pc_score_pool* score /* HOST */,
                                      int32_t dtype, void* stream);

/* ---- one decode step of one layer in ONE call --------------------------------------
 * The decode branch of TorchCUDASDPABackendImpl.forward (torch_cuda_sdpa.py:896-1062):
 * reshape_and_cache of the step's K/V -> [on steps where a row triggers: victim
 * selection + table compaction into compact_tables] -> paged decode attention over the
 * (compacted) tables with the fused scoring of freshly filled blocks.  Everything that
 * is the same for all layers of an engine step lives in struct pc_decode_step, a HOST
 * struct built once per step; the per-layer call passes only that layer's tensors.
 * any_trigger: 0 = no row evicts this step (compaction skipped, attention reads
 * block_tables); != 0 = run selection (out_victim must be non-NULL, int32 [batch]) and
 * attend over compact_tables unless attend_uncompacted is set (HEAD's streamingLLM
 * behaviour).  kv_lens are the lengths the host passes (already
 * reduced by block_size on trigger rows, vllm/worker/model_runner.py:540-543). */
typedef struct pc_decode_step {
  const int64_t* slot_mapping;   /* [batch] */
  const int32_t* block_tables;   /* [batch][table_stride] */
  int64_t table_stride;
  const int32_t* kv_lens;        /* [batch] physical KV lengths handed to attention */
  const int32_t* seq_lens;       /* [batch] logical lengths (trigger / score test) */
  const int32_t* n_blocks;       /* [batch] table entries each row owns (needed if any_trigger) */
  int32_t* compact_tables;       /* [batch][compact_stride] scratch (needed if any_trigger) */
  int64_t compact_stride;
  int32_t batch, num_heads, num_kv_heads, head_size, block_size, max_kv_len;
  float scale;
  int32_t any_trigger, attend_uncompacted;
  int64_t q_stride, key_stride, value_stride, out_stride, kv_block_stride;
  pc_evict_policy policy;
  void* workspace;               /* see pc_paged_decode_workspace_bytes */
  size_t workspace_bytes;
  int32_t dtype;
  void* stream;
  /* ABI 3 */
  int32_t fuse_cache_write;      /* != 0: the attention kernel writes this step's K/V rows itself
                                  * (no separate reshape_and_cache launch).  The caller guarantees
                                  * what the reference's builder produces for a decode row
                                  * (torch_cuda_sdpa.py:565-598): slot_mapping[r] ==
                                  * last_table_entry(r) * block_size + (kv_lens[r] - 1) % block_size.
                                  * Ignored (separate write) when attend_uncompacted is set. */
  const int32_t* plan;           /* DEVICE int32[2] = {chunk_pages, team} from pc_paged_decode_plan, read
                                  * by the kernel at run time instead of being planned from max_kv_len at
                                  * launch time -- lets a captured (CUDA graph) launch be replayed while
                                  * kv lengths move; NULL = plan on the host.  The workspace
                                  * (pc_paged_decode_workspace_bytes(max_kv_len)) covers every plan the
                                  * planner produces for lengths up to max_kv_len. */
  int32_t pdl;                   /* != 0: launch the attention kernel with programmatic stream
                                  * serialization (PDL): its prologue and first ring of KV-page loads
                                  * overlap the tail of the PRECEDING kernel of the stream.  The caller
                                  * promises that that kernel does not write this layer's KV pool, the
                                  * block tables or the length arrays (true for the preceding layer's
                                  * call and for the model's projection kernels).  Q and the step's
                                  * K/V rows are read after the dependency wait.  Only honoured with
                                  * fuse_cache_write on non-trigger steps. */
} pc_decode_step;
/* The launch plan the host planner picks for this shape, to fill pc_decode_step.plan:
 * out_plan[0] = pages per chunk (split-KV granularity), out_plan[1] = warps per team (1|2|4|8:
 * warps of one CTA that share a chunk and combine through shared memory). */
int pc_paged_decode_plan(int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_size,
                         int32_t block_size, int32_t max_kv_len,
                         int32_t num_sms /* 0 = the current device's; a pure function of its arguments otherwise */,
                         int32_t* out_plan /* HOST int32[2] */);
/* attend_tables (optional): tables [batch][attend_stride] this layer attends over INSTEAD of the
 * step's -- the layer's rows of pc_evict_select_compact_layers' output.  No selection is run by
 * this call then (out_victim is not written), and the attention kernel keeps its PDL launch. */
int pc_decode_layer(const pc_decode_step* step /* HOST */, void* out, const void* query,
                    const void* key, const void* value, void* key_cache, void* value_cache,
                    const pc_score_pool* score /* HOST, NULL = no value_l2 */,
                    int32_t* out_victim /* [batch] or NULL */,
                    const int32_t* attend_tables /* DEVICE or NULL */, int64_t attend_stride);

/* P3 + S4 + C1 for EVERY layer of an engine step in one launch (grid batch x layers).  A
 * layer's victims depend on the block scores of EARLIER steps and on the step's tables only
 * (torch_cuda_sdpa.py:1478-1560 reads BlockL2NormManager state written by previous forwards),
 * so all layers' selections can precede the first layer's attention: one launch instead of one
 * per layer, and the attention kernels of a trigger step stay PDL-chained.
 * score_pools: DEVICE array [num_layers] of this rank's pool of each layer (pc_score_pool
 * .pools[local]), NULL for streamingLLM.  out_tables [num_layers][batch][out_stride] (layer
 * stride out_layer_stride ints), out_victims [num_layers][batch] (-1 = row did not trigger;
 * decode_method 2: IN/OUT, pre-filled with the host's choices in (layer, row) order). */
int pc_evict_select_compact_layers(const int32_t* block_tables, int64_t table_stride,
                                   const int32_t* n_blocks, const int32_t* seq_lens, int32_t batch,
                                   int32_t num_layers, const pc_evict_policy* policy /* HOST */,
                                   const float* const* score_pools /* DEVICE */, int32_t heads_total,
                                   int32_t* out_tables, int64_t out_stride, int64_t out_layer_stride,
                                   int32_t* out_victims, void* stream);

/* ---- A1: prefill attention, TOP-LEFT aligned causal mask over the pruned prompt -----
 * Replaces _run_sdpa_forward (torch_cuda_sdpa.py:1069-1127): per request
 * scaled_dot_product_attention(q[Lq], k[Lkv], v[Lkv], is_causal=True), GQA by head
 * grouping instead of repeat_interleave; query i sees kv j <= i.
 * K/V rows are read through the same gather as pc_prune_prompt_and_cache, so the
 * pruned K/V never exists in memory.  req_meta int32 [num_reqs][6] =
 * {q_start, q_len, kv_src_start, kv_len(kept), first, pruned}; first and pruned are
 * multiples of 8 (whole blocks of 8 / 16 / 32; 8-row TMA boxes are used when they are not multiples
 * of 16, which needs req_meta_host).  num_tokens = rows of query / key / value (the
 * bound of the TMA tensor maps the tcgen05 kernel loads through).
 * req_meta_host: HOST copy of req_meta (the caller built it on the host anyway) or NULL.
 * When given, every row is validated before the launch (ranges inside num_tokens, first /
 * pruned multiples of 8) and a violation returns PC_ERR_INVALID; with NULL a violating row
 * reaches the kernel's own guard, which traps (and poisons the CUDA context). */
int pc_prefill_attn_topleft_causal(void* out, const void* query, const void* key,
                                   const void* value, const int32_t* req_meta,
                                   int32_t num_reqs, int32_t max_q_len, int64_t num_tokens,
                                   int32_t num_heads,
                                   int32_t num_kv_heads, int32_t head_size, float scale,
                                   int64_t q_stride, int64_t out_stride, int64_t key_stride,
                                   int64_t value_stride, const int32_t* req_meta_host /* HOST or NULL */,
                                   int32_t dtype, void* stream);

/* ---- f3: token-level prompt prune of `local` / `global` / `inverse_key_l2` ---------------
 * Replaces the second branch of KVCachePruner.prune_prompt_inplace
 * (vllm/attention/kvcache_prunner.py:220-279) for a whole prefill batch.  Request r:
 * req_meta int32 [num_reqs][6] = {src_start, q_len, first, mid, k_drop, dst_start}: source rows
 * [src_start, src_start + q_len); tokens [0, first) and [first + mid, q_len) are copied; of the
 * middle tokens [first, first + mid) every kv HEAD drops its k_drop lowest-scoring ones
 * (score_kind 0: ||V[t,h,:]||_2, 1: 1 / (||K[t,h,:]||_2 + 1e-8); fp32, fixed order of
 * csrc/score_common.cuh; equal scores: lowest token index goes first).  The q_len - k_drop
 * output rows start at out row dst_start.  head_mixing_repack != 0 reproduces the reference's
 * `middle[mask].view(-1, H, D)` (survivors in token-major, head-minor order re-chunked H at a
 * time, so a column mixes source heads -- SURVEY.md 0.5); 0 keeps each head's survivors in its
 * own column.  mid_offsets / tok_offsets int32 [num_reqs + 1] = prefix sums of mid / q_len.
 * workspace: pc_prompt_token_prune_workspace_bytes(total_mid, heads) bytes, 256-byte aligned. */
size_t pc_prompt_token_prune_workspace_bytes(int64_t total_mid, int32_t num_kv_heads);
int pc_prompt_token_prune(const void* key, const void* value, void* out_key, void* out_value,
                          const int32_t* req_meta, const int32_t* mid_offsets, const int32_t* tok_offsets,
                          int32_t num_reqs, int64_t total_tokens, int64_t total_mid, int32_t num_kv_heads,
                          int32_t head_size, int32_t score_kind, int32_t head_mixing_repack,
                          int64_t key_stride, int64_t value_stride, int64_t out_key_stride,
                          int64_t out_value_stride, void* workspace, size_t workspace_bytes, int32_t dtype,
                          void* stream);

/* ---- W2: slot zeroing of `streamingLLM-1` / `inverse_key_l2` --------------------------
 * Replaces the all-zero K/V rows the reference appends to the decode step's reshape_and_cache
 * call (TmpCacheSingleton rows, torch_cuda_sdpa.py:919-933,962-975; slots chosen at :1396-1401 --
 * slot (seq_len % bs - 1) of the table's 2nd block -- and :1444-1454 -- a random slot of a random
 * middle block).  slots int64 [num_slots] (DEVICE), < 0 = skip: the K and V rows of every kv head
 * at each slot become zero.  The _layers form zeroes [num_layers][slot_stride] slots in the
 * pools key_caches[l] / value_caches[l] (DEVICE arrays of pool base pointers) in ONE launch: the
 * slots are host decisions known before the step, so the zeros land ahead of the first layer's
 * attention and the layers' attention kernels keep their PDL chain. */
int pc_zero_cache_slots(void* key_cache, void* value_cache, const int64_t* slots, int32_t num_slots,
                        int32_t num_kv_heads, int32_t head_size, int32_t block_size,
                        int64_t kv_block_stride, void* stream);
int pc_zero_cache_slots_layers(void* const* key_caches /* DEVICE */, void* const* value_caches /* DEVICE */,
                               const int64_t* slots, int64_t slot_stride, int32_t num_slots,
                               int32_t num_layers, int32_t num_kv_heads, int32_t head_size,
                               int32_t block_size, int64_t kv_block_stride, void* stream);

/* ---- block maintenance (backend surface, torch_cuda_sdpa.py:75-88) ------------------
 * pc_copy_blocks replaces _C_cache_ops::copy_blocks (csrc/cache_kernels.cu:69-150) for
 * one layer: pairs int64 [num_pairs][2] = {src_block, dst_block}, same pool.
 * pc_swap_blocks replaces swap_blocks (csrc/cache_kernels.cu:24-66): block-granular
 * copies between two pools (either may be pinned host memory); pairs is a HOST array.  Runs of
 * pairs whose source and destination blocks are both consecutive travel as one copy. */
int pc_copy_blocks(void* key_cache, void* value_cache, const int64_t* pairs, int32_t num_pairs,
                   int64_t block_bytes, void* stream);
int pc_swap_blocks(const void* src, void* dst, const int64_t* pairs /* HOST */,
                   int32_t num_pairs, int64_t block_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PAGEDEVICT_B200_H_ */
