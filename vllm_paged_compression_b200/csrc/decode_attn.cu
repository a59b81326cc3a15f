// Paged decode attention over (compacted) block tables with value_l2 block scoring
// fused into the same KV read.  sm_100a.
//
// Replaces csrc/attention/attention_kernels.cuh:86-668 (paged_attention_v1/v2 + reduce)
// and the extra V read of torch_cuda_sdpa.py:1518-1526 (last-block scoring).
//
// Design (HBM-bound; see DESIGN.md "decode kernel"):
//   * unit of work = one PAGE = (sequence b, kv head h [, 8-q-head group hg], block):
//     two contiguous chunks of the reference cache layout (K [D/8][BS][8], V [D][BS],
//     vllm/attention/ops/paged_attn.py:49-62), fetched with two TMA 1-D bulk copies
//     (cp.async.bulk -> UBLKCP) that complete on one mbarrier.
//   * schedule: every (row, group) is cut into CHUNKS of `chunk_pages` consecutive pages
//     (a fixed, deterministic partition); chunks are numbered in (b, h, hg, chunk) order.
//     A TEAM of T = 1, 2, 4 or 8 warps of one CTA is a worker (T from the host planner):
//     member j of the team streams pages begin + j, + T, ... of the chunk through its own
//     page ring, and at the end of the chunk the members hand their (m, l, O) to member 0
//     through SHARED memory (one mbarrier pair per member) -- no global partial, no fence,
//     no ticket.  With T = 1 a single warp needs ~0.7 us per page whatever its neighbours
//     do (dependent chain + a 3-deep ring), so a launch with fewer chunks than warps is
//     latency-bound (the 64 x ~590-token bench shape: 512 chunks on 1184 warps); teams put
//     every warp of an SM on the SM's chunks.  A team's first chunk is static (id = team x
//     grid + cta, so neighbouring chunks land on different SMs); with T = 1 further chunks
//     come from a global atomic counter, with T > 1 the planner keeps the launch to one
//     wave (chunks <= teams).  The hand-out must be dynamic: per-SM HBM bandwidth differs by
//     ~+-15 % across the chip (measured with the trace hook, profiles/), a static equal
//     split loses that much at the tail.  Everything on the hand-out path is prefetched
//     one chunk ahead -- the counter atomic, the chunk descriptor (binary search of a
//     per-row prefix sum that every CTA derives from the device-side kv_lens), the block
//     table entries and Q -- and the kStages-deep page ring keeps running ACROSS chunk
//     boundaries, so a warp never drains its loads in flight.  No __syncthreads after
//     the prologue.
//   * the whole GQA group (<= 8 q heads) is served from ONE read of the page:
//     S^T[16 tok x 8 heads] = K_tile(A, ldmatrix straight from the cache layout)
//     x Q^T(B, registers) with mma.sync m16n8k16 bf16/fp32; online softmax with warp
//     shuffles in the exp2 domain; P^T is transposed in registers (movmatrix) and
//     O^T[D x 8 heads] += V^T_tile(A, ldmatrix straight from the cache layout) x P^T.
//     Tensor cores are used because 2*G FMAs per KV byte exceed the SIMT fp32 budget
//     at HBM speed for G >= 4; the legacy HMMA path is far from saturated (HBM-bound).
//   * numerics follow attention_kernels.cuh: fp32 logits, probabilities rounded to
//     bf16 before P.V, fp32 accumulation, final 1/(sum + 1e-6).
//   * a group with more than one chunk is finished split-KV style: every chunk publishes
//     (m, l, O) fp32 to its own workspace slot and the LAST one to arrive (acq_rel
//     ticket) combines them in chunk order (deterministic), warp-parallel over heads x
//     dims, while the warp's ring already holds the next chunk's pages.  Tickets and the
//     chunk counter are reset in-kernel, so the op is a single launch with no memset.
//   * scoring: when seq_lens[b] % BS == 0 the row's last block just filled; the warp
//     that owns that page computes the per-head value_l2 partial from the V page it
//     already has in shared memory (zero extra HBM bytes) and stores it at
//     pool[physical block][head] of every rank's score pool (peer stores under TP).
#include "pc_common.cuh"
#include "score_common.cuh"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <atomic>
#include <type_traits>

#include "pc_api_internal.h"

namespace pc {

struct DecodeParams {
  const __nv_bfloat16* __restrict__ q;
  const __nv_bfloat16* __restrict__ k_cache;
  const __nv_bfloat16* __restrict__ v_cache;
  const int* __restrict__ block_tables;
  const int* __restrict__ kv_lens;
  __nv_bfloat16* __restrict__ out;
  float* part;      // split-KV partials, one slot per chunk id: [8][D] O then [8][2] (m, l); the slot
                    // stride depends on D only, so the layout is the same whatever chunk plan a
                    // (graph-replayed) launch reads from device memory
  int cap_chunks;   // slots the workspace holds
  int* counters;    // [0] next chunk id - launched warps, [1] warps that have exited; zero between launches
  int* tickets;     // [B * GH] arrival counters, zero between launches
  const int* __restrict__ seq_lens;     // nullptr: no scoring
  ScorePool score;                      // per-block value_l2 partials, indexed by physical block
  int B, Hq, Hkv, G, HG, GH;  // G = Hq/Hkv, HG = ceil(G/8), GH = Hkv*HG groups per row
  long long q_stride;         // elements between sequences in q
  long long out_stride;       // elements between sequences in out
  long long table_stride;     // ints between rows
  long long kv_block_stride;  // elements between blocks in k/v cache
  float scale_log2;
  int chunk_pages;            // pages per chunk (the last chunk of a group may be shorter)
  int team;                   // warps per team (1, 2, 4, 8; <= team_cap)
  int team_cap;               // largest team the launch has shared memory for
  const int* __restrict__ plan;  // device-side {chunk_pages, team} overriding the two fields above (CUDA-graph replay), or nullptr
  // fused reshape_and_cache of this step's token (nullptr = the caller wrote the cache):
  // the warp that owns a row's LAST page patches the new token's K/V into the page it has
  // in shared memory and stores it to the pool at slot_mapping[b]
  const __nv_bfloat16* __restrict__ new_k;
  const __nv_bfloat16* __restrict__ new_v;
  const long long* __restrict__ slot_mapping;
  __nv_bfloat16* k_cache_w;
  __nv_bfloat16* v_cache_w;
  long long new_k_stride, new_v_stride;
  int pdl;  // launched with programmatic stream serialization: page prefetch runs ahead of the wait
  unsigned long long* trace;  // debug: [warps][8]: 4 globaltimer stamps, chunks, combines, smid, membar ns (nullptr = off)
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// ticket arrival: releases this warp's partial (after __syncwarp) and acquires the others'
__device__ __forceinline__ int atom_add_acq_rel(int* p, int v) {
  int old;
  asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

constexpr int kFifo = 8;  // chunk descriptors in flight per warp (> kStages)

struct ChunkDesc {  // 32 bytes
  int id, b, gh, blk_begin, blk_end, nblk, kv_len, pad;
};

template <int D, int BS, int kWarps, int kStages>
struct DecodeTraits {
  static constexpr int kPageBytes = BS * D * 2;        // K or V page of one (block, head)
  static constexpr int kStageBytes = 2 * kPageBytes;   // K + V
  static constexpr int kRingBytes = kWarps * kStages * kStageBytes;
  static constexpr int kBarBytes = kWarps * kStages * 8;
  static constexpr int kFifoBytes = kWarps * kFifo * (int)sizeof(ChunkDesc);
  static constexpr int kTeamBarBytes = 2 * kWarps * 8;          // pfull[kWarps], pfree[kWarps]
  // a member's hand-over buffer: its O^T fragments [D/4 regs][32 lanes] fp32 + (m, l) of the 8 heads
  static constexpr int kPartBytes = 32 * (D / 4) * 4 + 64;
  static constexpr int kThreads = kWarps * 32;
  static constexpr int followers(int team) { return team > 1 ? kWarps - kWarps / team : 0; }
  static size_t smem_bytes(int B, int team_cap) {
    return (size_t)kRingBytes + kBarBytes + kFifoBytes + kTeamBarBytes + (size_t)followers(team_cap) * kPartBytes +
           ((size_t)B + 1) * 4 + 64;  // + wsum[kWarps] + flags
  }
};

// W1 fused into A2: patch the step's K/V row of kv head h (nk = this lane's 16 bytes of the
// K row, nv = its D/32 V elements) into the page under attention at in-page offset `off`
// -- iff slot_mapping says the token lives there -- and store it to the pool at
// slot_mapping[b] (csrc/cache_kernels.cu:152-205 semantics: K [D/8][BS][8], V [D][BS],
// slot < 0 = skip).  Only the row's last page runs this (peeled page variant).
template <int D, int BS>
__device__ __forceinline__ void write_new_token(__nv_bfloat16* k_cache_w, __nv_bfloat16* v_cache_w,
                                                long long kv_block_stride, long long nslot, int phys,
                                                uint8_t* kpage, uint8_t* vpage, uint4 nk,
                                                const __nv_bfloat16 (&nv)[D / 32], int h, bool store, int off) {
  const int lane = threadIdx.x & 31;
  constexpr int VPL = D / 32;  // V dims per lane
  const bool here = (nslot == (long long)phys * BS + off);
  store = store && (nslot >= 0);
  const long long wblk = nslot / BS;
  const int woff = (int)(nslot - wblk * BS);
  const size_t wbase = (size_t)wblk * kv_block_stride + (size_t)h * (D * BS);
  if (lane < D / 8) {
    if (here) *reinterpret_cast<uint4*>(kpage + (size_t)(lane * BS + off) * 16) = nk;
    if (store) *reinterpret_cast<uint4*>(k_cache_w + wbase + (size_t)(lane * BS + woff) * 8) = nk;
  }
  __nv_bfloat16* vp = reinterpret_cast<__nv_bfloat16*>(vpage);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int dd = lane * VPL + i;
    if (here) vp[dd * BS + off] = nv[i];
    if (store) v_cache_w[wbase + (size_t)dd * BS + woff] = nv[i];
  }
  fence_proxy_async();  // generic-proxy writes before a later TMA refill of this slot
  __syncwarp();
}

template <int D, int BS, int kWarps, int kStages, int kCtasPerSm = 1>
__global__ void __launch_bounds__(kWarps * 32, kCtasPerSm)
paged_decode_attn_score_kernel(const DecodeParams p) {
  using T = DecodeTraits<D, BS, kWarps, kStages>;
  static_assert(kStages < kFifo, "descriptor fifo must outrun the page ring");
  constexpr int KSTEPS = D / 16;   // k-steps of the QK^T contraction
  constexpr int MT = D / 16;       // m-tiles (16 dims) of O^T
  // 16-token MMA tiles per page.  Block size 8 (scripts/diff_page_sizes.sh:16): one tile per page whose
  // upper 8 token rows do not exist -- their ldmatrix rows alias the lower 8 (finite values), their
  // logits are masked, so their probabilities are exactly 0.  Twice the tensor work per byte on a
  // kernel whose tensor pipe is 7 % busy.
  constexpr int TT = (BS + 15) / 16;
  constexpr bool kHalfTile = (BS == 8);
  constexpr int kSlot = 8 * D + 16;  // floats per partial slot (16-byte aligned stride)

  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;  // fragment row group
  const int t = lane & 3;   // fragment column pair
  const int n_launched = gridDim.x * kWarps;

  const unsigned long long t_start = p.trace ? gtimer() : 0ull;
  // the first pass of the schedule scan needs this row's kv length: request it before anything else,
  // next to the plan words, so that the prologue pays ONE global-memory latency, not two in a row
  const int len_first = ((int)threadIdx.x < p.B) ? __ldg(p.kv_lens + threadIdx.x) : 0;
  const int chunk_pages = (p.plan != nullptr) ? max(1, __ldg(p.plan)) : p.chunk_pages;
  // team geometry: TS warps per team (power of two), member tw of team `team_in_cta`
  int TS = (p.plan != nullptr) ? __ldg(p.plan + 1) : p.team;
  TS = (TS >= 8 && kWarps >= 8 && p.team_cap >= 8) ? 8 : (TS >= 4 && kWarps >= 4 && p.team_cap >= 4) ? 4
       : (TS >= 2 && p.team_cap >= 2) ? 2 : 1;
  const int tw = warp & (TS - 1);
  const int team_in_cta = warp / TS;
  const int n_teams = gridDim.x * (kWarps / TS);
  const int w = team_in_cta * gridDim.x + blockIdx.x;  // team id, interleaved over SMs
  ChunkDesc* my_fifo = reinterpret_cast<ChunkDesc*>(smem + T::kRingBytes + T::kBarBytes) + warp * kFifo;
  uint64_t* pfull = reinterpret_cast<uint64_t*>(smem + T::kRingBytes + T::kBarBytes + T::kFifoBytes);  // [kWarps]
  uint64_t* pfree = pfull + kWarps;                                                                       // [kWarps]
  uint8_t* part_base = smem + T::kRingBytes + T::kBarBytes + T::kFifoBytes + T::kTeamBarBytes;
  int* cpre = reinterpret_cast<int*>(part_base + (size_t)T::followers(p.team_cap) * T::kPartBytes);  // [B + 1]

  // chunk ids of this warp: the first one is static (= w), the rest come from a global
  // counter, fetched only when the issue side runs out of pages (kStages pages before the
  // consumer does), so a slow SM never commits to work ahead of time.
  if (lane == 0) {
    uint64_t* bars0 = reinterpret_cast<uint64_t*>(smem + T::kRingBytes) + warp * kStages;
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(&bars0[s], 1);
    mbar_init(&pfull[warp], 1);
    mbar_init(&pfree[warp], 1);
  }
  // ---- schedule: chunks per (row, group) -> exclusive prefix sum (same in every CTA)
  // Programmatic dependent launch: let the next kernel of the stream start as SMs free
  // up; everything up to pdl_wait() below reads only the step's integer state and the KV
  // pool pages (written on earlier steps / by earlier layers, never by the preceding
  // kernel -- the caller's promise when it sets `pdl`), so the first ring of pages is in
  // flight while the preceding kernel drains.
  if (p.pdl) asm volatile("griddepcontrol.launch_dependents;");
  {
    // block-wide scan, kThreads rows per pass (a one-warp scan cost 1.2 us of prologue at
    // B = 512, the per-rank batch of the TP = 8 bench)
    int* wsum = cpre + p.B + 1;   // [kWarps], then [kWarps] = "some row is empty"
    if (threadIdx.x == 0) wsum[kWarps] = 0;
    int carry = 0;
    bool saw_empty = false;
    for (int base = 0; base < p.B; base += T::kThreads) {
      const int r = base + threadIdx.x;
      int n = 0;
      if (r < p.B) {
        const int len = (base == 0) ? len_first : __ldg(p.kv_lens + r);
        n = ((len + BS - 1) / BS + chunk_pages - 1) / chunk_pages;
        saw_empty |= (len == 0);
      }
#pragma unroll
      for (int sh = 1; sh < 32; sh <<= 1) {
        const int v = __shfl_up_sync(kFull, n, sh);
        if (lane >= sh) n += v;
      }
      if (lane == 31) wsum[warp] = n;
      __syncthreads();
      int off = carry, tot = 0;
#pragma unroll
      for (int i = 0; i < kWarps; ++i) {
        const int v = wsum[i];
        off += (i < warp) ? v : 0;
        tot += v;
      }
      if (r < p.B) cpre[r + 1] = off + n;
      carry += tot;
      __syncthreads();
    }
    if (threadIdx.x == 0) cpre[0] = 0;
    if (saw_empty) wsum[kWarps] = 1;   // after the loop's last barrier, before the one below
  }
  mbar_fence_init();
  __syncthreads();
  const bool any_empty_row = (cpre + p.B + 1)[kWarps] != 0;

  const int C = cpre[p.B] * p.GH;  // chunks in this launch
  // the host sizes the workspace for the worst plan up to max_kv_len (decode_workspace_bytes);
  // a device-side plan / kv_lens beyond that must not scribble over other memory
  if (C > p.cap_chunks) __trap();

  // (a runtime ring depth that lends idle warps' slots to busy ones was measured and is
  // slower: profiles/r01_decode_variants.txt)
  constexpr int depth = kStages;
  uint8_t* my_stages = smem + (size_t)warp * depth * T::kStageBytes;
  uint64_t* my_bars = reinterpret_cast<uint64_t*>(smem + T::kRingBytes) + warp * depth;

  // ---- chunk stream: ids -> descriptors (fifo in shared memory, written by the issue side)
  int fifo_head = 0;   // descriptors opened (issue side)
  int fifo_tail = 0;   // descriptors retired (consume side)
  int next_id = w;     // id the next open_chunk() will decode
  bool first_open = true;
  int static_next = w;  // teams: ids w, w + n_teams, ... (every member derives the same sequence)
  auto open_chunk = [&]() -> bool {  // all lanes, uniform
    if (TS > 1) {
      next_id = static_next;
      static_next += n_teams;
    } else if (!first_open) {
      int nid = 0;
      if (lane == 0) nid = atomicAdd(&p.counters[0], 1);
      next_id = __shfl_sync(kFull, nid, 0) + n_launched;
    }
    first_open = false;
    if (next_id >= C) return false;
    const int id = next_id;
    next_id = C;  // the following id is fetched on demand
    const int key = id / p.GH;
    int lo = 0, hi = p.B;  // largest b with cpre[b] <= key (cpre[B] > key)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cpre[mid] <= key) lo = mid; else hi = mid;
    }
    const int cpr = cpre[lo + 1] - cpre[lo];  // chunks per group in this row
    const int rem = id - p.GH * cpre[lo];
    const int gh = rem / cpr;
    const int c = rem - gh * cpr;
    const int kv_len = __ldg(p.kv_lens + lo);
    const int nblk = (kv_len + BS - 1) / BS;
    if (lane == 0) {
      ChunkDesc d;
      d.id = id; d.b = lo; d.gh = gh; d.blk_begin = c * chunk_pages;
      d.blk_end = min(nblk, d.blk_begin + chunk_pages); d.nblk = nblk; d.kv_len = kv_len; d.pad = 0;
      my_fifo[fifo_head % kFifo] = d;
    }
    ++fifo_head;
    __syncwarp();
    return true;
  };

  // ---- issue side: walks this member's pages of successive chunks, up to kStages ahead of the
  // consumer.  Invariant: when is_open, the cursor (is_b, is_h, is_blk) names a page of this
  // member that is not issued yet and phys_next holds its table entry (loaded one page ahead).
  int is_b = 0, is_h = 0, is_blk = 0, is_end = 0;
  bool is_open = false;
  int phys_next = 0;
  int in_flight = 0;                  // pages issued and not yet consumed
  int issue_slot = 0;                 // ring slot the next page goes to
  int cons_slot = 0;                  // ring slot of the next page to consume
  uint32_t cons_parity = 0;           // mbarrier phase parity of that slot

  // before the dependency wait (PDL) only static chunk ids may be opened: the chunk counter
  // still belongs to the preceding launch.  is_pending: the cursor stopped short of a chunk it
  // may open later (dependency wait not passed, or the descriptor fifo is full).
  bool allow_open = !p.pdl;
  bool is_pending = false;
  auto issue_advance = [&]() {
    if (is_open && is_blk < is_end) {
      phys_next = __ldg(p.block_tables + (size_t)is_b * p.table_stride + is_blk);
      return;
    }
#pragma unroll 1
    while (true) {
      if ((TS == 1 && !allow_open && !first_open) || fifo_head - fifo_tail >= kFifo - 1) {
        is_open = false;
        is_pending = true;
        return;
      }
      is_open = open_chunk();
      if (!is_open) return;
      const ChunkDesc& d = my_fifo[(fifo_head - 1) % kFifo];
      is_b = d.b; is_h = d.gh / p.HG; is_blk = d.blk_begin + tw; is_end = d.blk_end;
      if (is_blk < is_end) break;   // (a member beyond the chunk's page count sits this chunk out)
    }
    phys_next = __ldg(p.block_tables + (size_t)is_b * p.table_stride + is_blk);
  };
  auto issue_page = [&]() {  // issue the page under the cursor into the next ring slot
    const int slot = issue_slot;
    issue_slot = (issue_slot + 1 == depth) ? 0 : issue_slot + 1;
    // one elected lane (elect.sync): under `lane == 0` the compiler wraps the mbarrier / bulk-copy instructions in an
    // ELECT / R2UR / BRA.U.ANY loop each; same-box A/B +0.3-1.1 % (profiles/r02_decode_elect_ab.jsonl)
    if (elect_one_lane()) {
      uint8_t* dst = my_stages + (size_t)slot * T::kStageBytes;
      const size_t off = ((size_t)phys_next * p.kv_block_stride) + (size_t)is_h * (D * BS);
      mbar_expect_tx(&my_bars[slot], T::kStageBytes);
      tma_load_1d(dst, p.k_cache + off, T::kPageBytes, &my_bars[slot]);
      tma_load_1d(dst + T::kPageBytes, p.v_cache + off, T::kPageBytes, &my_bars[slot]);
    }
    ++in_flight;
    is_blk += TS;
  };

  // Q^T as the mma B operand: b0 = Q[head g][16j + 2t..], b1 = Q[head g][16j + 8 + 2t..]
  auto load_q = [&](uint32_t (&qf)[KSTEPS][2], int b, int gh) {
    const int h = gh / p.HG, hg = gh - h * p.HG;
    const int head0 = h * p.G + hg * 8;
    const int nheads = min(8, p.G - hg * 8);
    const bool valid = g < nheads;
    const __nv_bfloat16* qrow = p.q + (size_t)b * p.q_stride + (size_t)(head0 + (valid ? g : 0)) * D;
#pragma unroll
    for (int j = 0; j < KSTEPS; ++j) {
      qf[j][0] = valid ? __ldg(reinterpret_cast<const unsigned int*>(qrow + 16 * j + 2 * t)) : 0u;
      qf[j][1] = valid ? __ldg(reinterpret_cast<const unsigned int*>(qrow + 16 * j + 8 + 2 * t)) : 0u;
    }
  };
  uint32_t qf[KSTEPS][2];   // Q of the chunk being consumed
  uint32_t qn[KSTEPS][2];   // Q prefetched for chunk `qn_id`
  int qn_id = -1;
  // top up the ring; prefetches Q of a chunk opened on the way.  One instance of this loop serves
  // the prologue (both sides of the PDL dependency wait), the refill after every page and the
  // retry of a pending cursor (the issue path is large; few instances keep the kernel inside the
  // instruction cache).
  auto fill = [&](bool want_q) {
    if (is_pending) {
      is_pending = false;
      issue_advance();
    }
#pragma unroll 1
    while (in_flight < depth && is_open) {
      const int opened = fifo_head;
      issue_page();
      issue_advance();
      if (want_q && fifo_head != opened) {  // a new chunk was opened: prefetch its Q
        const ChunkDesc& nd = my_fifo[(fifo_head - 1) % kFifo];
        load_q(qn, nd.b, nd.gh);
        qn_id = nd.id;
      }
    }
  };

  // prologue: fill the ring
  issue_advance();
  fill(false);
  // ---- from here on the kernel reads what the preceding kernel may have produced (Q, the
  // step's K/V rows) and writes (out, partials, tickets, counters, scores, pool)
  if (p.pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    allow_open = true;
    fill(false);
  }
  if (any_empty_row && blockIdx.x == 0) {  // empty sequences: defined output (zeros)
    for (int r = warp; r < p.B; r += kWarps) {
      if (p.kv_lens[r] != 0) continue;
      __nv_bfloat16* o = p.out + (size_t)r * p.out_stride;
      for (int i = lane; i < p.Hq * D; i += 32) o[i] = __float2bfloat16(0.f);
    }
  }
  if (fifo_head > 0) {
    load_q(qn, my_fifo[0].b, my_fifo[0].gh);
    qn_id = my_fifo[0].id;
  }
  if (p.trace && lane == 0) p.trace[(warp * gridDim.x + blockIdx.x) * 8 + 0] = t_start,
                            p.trace[(warp * gridDim.x + blockIdx.x) * 8 + 1] = gtimer();

  float o[MT][4];
  float m_run[2], l_run[2];
  int tr_chunks = 0, tr_combines = 0;
  unsigned long long tr_fix = 0ull;
  int pf_sent = 0;            // follower: hand-overs written so far (mbarrier phase of pfree / pfull)
  uint32_t pf_parity = 0u;    // leader: bit f = parity of the next hand-over expected from member f

  // ================================================================ chunk loop
  while (true) {
    if (fifo_tail == fifo_head) {
      if (!is_pending) break;
      fill(true);   // a cursor parked on a full descriptor fifo
      if (fifo_tail == fifo_head) break;
    }
    const ChunkDesc d = my_fifo[fifo_tail % kFifo];
    ++fifo_tail;
    const int b = d.b, nblk = d.nblk, kv_len = d.kv_len;
    // members of the team that have pages in this chunk; the others sit it out
    const int n_active = min(TS, d.blk_end - d.blk_begin);
    if (tw >= n_active) continue;
    ++tr_chunks;
    const int h = d.gh / p.HG, hg = d.gh - h * p.HG;
    if (qn_id == d.id) {
#pragma unroll
      for (int j = 0; j < KSTEPS; ++j) { qf[j][0] = qn[j][0]; qf[j][1] = qn[j][1]; }
    } else {
      load_q(qf, b, d.gh);
    }
    // the row's last page belongs to this member iff it is one of its pages of this chunk
    const bool has_tail = (d.blk_end == nblk) && (((nblk - 1 - d.blk_begin) & (TS - 1)) == tw);
    if (p.new_k != nullptr && has_tail) {
      // this member ends with the row's last page: pull the step's K/V row towards L1 now, so
      // the loads in the tail page (no registers to spare until then) hit
      const __nv_bfloat16* ka = p.new_k + (size_t)b * p.new_k_stride + (size_t)h * D + (lane & (D / 8 - 1)) * 8;
      const __nv_bfloat16* va = p.new_v + (size_t)b * p.new_v_stride + (size_t)h * D + lane * (D / 32);
      asm volatile("prefetch.global.L1 [%0];" ::"l"(ka));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(va));
      if (lane == 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(p.slot_mapping + b));
    }
#pragma unroll
    for (int m = 0; m < MT; ++m) o[m][0] = o[m][1] = o[m][2] = o[m][3] = 0.f;
    m_run[0] = m_run[1] = -INFINITY;
    l_run[0] = l_run[1] = 0.f;

    // One page.  kTail = it is the row's LAST page: the only page that can be partially
    // filled, receive the step's token (fused cache write) or be scored; the interior pages
    // run the variant without any of that.
    // NP consecutive pages processed as ONE softmax tile of NP*BS tokens.  A warp that
    // streams alone on its SM sub-partition is bound by the dependent chain of a page
    // (ldmatrix -> HMMA -> shuffle max -> ex2 -> movmatrix -> HMMA), not by HBM (trace hook:
    // every warp needs ~0.68 us per page however many neighbours it has); two pages share one
    // max/rescale chain and give the QK^T HMMAs twice the independent accumulators.
    // kTail = the row's LAST page (NP == 1): the only page that can be partially filled,
    // receive the step's token (fused cache write) or be scored.
    auto page = [&](const int blk, auto tail_tag, auto np_tag) {
      constexpr bool kTail = decltype(tail_tag)::value;
      constexpr int NP = decltype(np_tag)::value;
      constexpr int J = NP * TT;  // 16-token tiles handled jointly
      static_assert(!kTail || NP == 1, "the tail page is processed alone");
      uint32_t slot[NP], parity[NP];
#pragma unroll
      for (int pg = 0; pg < NP; ++pg) {
        slot[pg] = (uint32_t)cons_slot;
        parity[pg] = cons_parity;
        if (++cons_slot == depth) { cons_slot = 0; cons_parity ^= 1u; }
      }
      const int page_tok0 = blk * BS;
      const bool last_partial = kTail && (page_tok0 + BS > kv_len);
      const bool write_new = kTail && (p.new_k != nullptr);
      // this step's token (W1 fused): its loads overlap the wait for the page
      uint4 nk = make_uint4(0u, 0u, 0u, 0u);
      __nv_bfloat16 nv[D / 32];
      long long nslot = -1;
      int nphys = 0;
      if constexpr (kTail) {
        if (write_new) {
          nslot = __ldg(p.slot_mapping + b);
          nphys = __ldg(p.block_tables + (size_t)b * p.table_stride + blk);
          if (lane < D / 8)
            nk = __ldg(reinterpret_cast<const uint4*>(p.new_k + (size_t)b * p.new_k_stride + (size_t)h * D + lane * 8));
          const __nv_bfloat16* vsrc = p.new_v + (size_t)b * p.new_v_stride + (size_t)h * D + lane * (D / 32);
          if constexpr (D / 32 == 4) *reinterpret_cast<uint2*>(nv) = __ldg(reinterpret_cast<const uint2*>(vsrc));
          else *reinterpret_cast<unsigned int*>(nv) = __ldg(reinterpret_cast<const unsigned int*>(vsrc));
        }
      }
      uint32_t kaddr[NP], vaddr[NP];
#pragma unroll
      for (int pg = 0; pg < NP; ++pg) {
        mbar_wait(&my_bars[slot[pg]], parity[pg]);
        kaddr[pg] = smem_u32(my_stages + (size_t)slot[pg] * T::kStageBytes);
        vaddr[pg] = kaddr[pg] + T::kPageBytes;
      }
      const uint8_t* kpage = my_stages + (size_t)slot[0] * T::kStageBytes;
      const uint8_t* vpage = kpage + T::kPageBytes;

      if constexpr (kTail) {
        if (write_new)  // W1 fused: patch the step's token into the page and store it to the pool
          write_new_token<D, BS>(p.k_cache_w, p.v_cache_w, p.kv_block_stride, nslot, nphys,
                                 const_cast<uint8_t*>(kpage), const_cast<uint8_t*>(vpage), nk, nv, h, hg == 0,
                                 kv_len - 1 - page_tok0);
        if (last_partial) {
          // zero V columns of tokens >= kv_len so 0 * garbage cannot produce NaN
          // (attention_kernels.cuh:455-466 does the same per element)
          const int valid = kv_len - page_tok0;
          __nv_bfloat16* vp = reinterpret_cast<__nv_bfloat16*>(const_cast<uint8_t*>(vpage));
          for (int i = lane; i < D * BS; i += 32)
            if ((i % BS) >= valid) vp[i] = __float2bfloat16(0.f);
          fence_proxy_async();  // generic-proxy writes before a later TMA refill of this slot
          __syncwarp();
        }
      }

      // ---------------- S^T = K_tile x Q^T for all J tiles; k-step outer so that consecutive
      // HMMAs belong to different accumulators (the asm statements keep program order)
      float sa[J][4], sb[J][4];
#pragma unroll
      for (int jt = 0; jt < J; ++jt) {
        sa[jt][0] = sa[jt][1] = sa[jt][2] = sa[jt][3] = 0.f;
        sb[jt][0] = sb[jt][1] = sb[jt][2] = sb[jt][3] = 0.f;
      }
      {
        // ldmatrix.x4 row address: matrix mi = lane/8 -> (token half = mi&1, chunk = 2j + mi>>1)
        const int mi = lane >> 3, r = lane & 7;
        const uint32_t koff = (uint32_t)(((mi >> 1) * BS + (kHalfTile ? 0 : (mi & 1) * 8) + r) * 16);
#pragma unroll
        for (int j = 0; j < KSTEPS; ++j) {
          uint32_t a[J][4];
#pragma unroll
          for (int jt = 0; jt < J; ++jt)
            ldmatrix_x4(a[jt], kaddr[jt / TT] + koff + (uint32_t)((jt % TT) * 16 * 16 + 2 * j * BS * 16));
#pragma unroll
          for (int jt = 0; jt < J; ++jt) {
            if (j & 1) mma_bf16_16816(sb[jt], a[jt], qf[j][0], qf[j][1]);
            else       mma_bf16_16816(sa[jt], a[jt], qf[j][0], qf[j][1]);
          }
        }
      }
      // x[jt][0]=(tok g, head 2t) [1]=(tok g, head 2t+1) [2]=(tok g+8, head 2t) [3]=(tok g+8, 2t+1)
      float x[J][4];
#pragma unroll
      for (int jt = 0; jt < J; ++jt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) x[jt][i] = (sa[jt][i] + sb[jt][i]) * p.scale_log2;
        if (last_partial) {
          const int tok = page_tok0 + jt * 16 + g;
          if (tok >= kv_len) { x[jt][0] = -INFINITY; x[jt][1] = -INFINITY; }
          if (tok + 8 >= kv_len) { x[jt][2] = -INFINITY; x[jt][3] = -INFINITY; }
        }
        if constexpr (kHalfTile) { x[jt][2] = -INFINITY; x[jt][3] = -INFINITY; }   // token rows 8..15 of an 8-token page
      }
      // ---------------- online softmax over the J*16 tokens (per head = per column; rows live
      // in lanes with equal t)
      float mx0 = fmaxf(x[0][0], x[0][2]), mx1 = fmaxf(x[0][1], x[0][3]);
#pragma unroll
      for (int jt = 1; jt < J; ++jt) {
        mx0 = fmaxf(mx0, fmaxf(x[jt][0], x[jt][2]));
        mx1 = fmaxf(mx1, fmaxf(x[jt][1], x[jt][3]));
      }
#pragma unroll
      for (int sh = 4; sh <= 16; sh <<= 1) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(kFull, mx0, sh));
        mx1 = fmaxf(mx1, __shfl_xor_sync(kFull, mx1, sh));
      }
      const float mn0 = fmaxf(m_run[0], mx0), mn1 = fmaxf(m_run[1], mx1);
      const float al0 = fast_exp2(m_run[0] - mn0), al1 = fast_exp2(m_run[1] - mn1);
      m_run[0] = mn0; m_run[1] = mn1;
      uint32_t pb[J][2];  // P^T as the mma B operand: the two 8x8 (tok x head) tiles, transposed
      float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
      for (int jt = 0; jt < J; ++jt) {
        const float p0 = fast_exp2(x[jt][0] - mn0), p1 = fast_exp2(x[jt][1] - mn1);
        const float p2 = fast_exp2(x[jt][2] - mn0), p3 = fast_exp2(x[jt][3] - mn1);
        ps0 += p0 + p2;
        ps1 += p1 + p3;
        pb[jt][0] = movmatrix_trans(pack_bf16x2(p0, p1));  // tokens 0..7
        pb[jt][1] = movmatrix_trans(pack_bf16x2(p2, p3));  // tokens 8..15
      }
      l_run[0] = l_run[0] * al0 + ps0;
      l_run[1] = l_run[1] * al1 + ps1;
      if (__any_sync(kFull, (al0 != 1.f) || (al1 != 1.f))) {
#pragma unroll
        for (int m = 0; m < MT; ++m) { o[m][0] *= al0; o[m][1] *= al1; o[m][2] *= al0; o[m][3] *= al1; }
      }
      // ---------------- O^T += V^T_tile x P^T
      {
        // matrix mi -> (dim half = mi&1, token half = mi>>1); row r = dim within the half
        const int mi = lane >> 3, r = lane & 7;
        const uint32_t voff = (uint32_t)((((mi & 1) * 8 + r) * BS + (kHalfTile ? 0 : (mi >> 1) * 8)) * 2);
#pragma unroll
        for (int jt = 0; jt < J; ++jt) {
#pragma unroll
          for (int m = 0; m < MT; ++m) {
            uint32_t a[4];
            ldmatrix_x4(a, vaddr[jt / TT] + voff + (uint32_t)(((jt % TT) * 16 + m * 16 * BS) * 2));
            mma_bf16_16816(o[m], a, pb[jt][0], pb[jt][1]);
          }
        }
      }

      // ---------------- fused value_l2 score of the row's last (just filled) block
      if (kTail && p.seq_lens != nullptr && hg == 0 && (kv_len % BS) == 0 &&
          (__ldg(p.seq_lens + b) % BS) == 0) {
        const __nv_bfloat16* vp = reinterpret_cast<const __nv_bfloat16*>(vpage);
        const float sc = warp_value_l2_head_score<D, BS>(
            [&](int tok, int dd) { return __bfloat162float(vp[dd * BS + tok]); });
        if (lane == 0)
          score_store(p.score, __ldg(p.block_tables + (size_t)b * p.table_stride + blk), h, sc);
      }

      // ---------------- refill the consumed slots (the issue side may open the next chunk here)
      __syncwarp();
      in_flight -= NP;
      fill(true);
    };
    {
      const int body_end = d.blk_end - (has_tail ? 1 : 0);
#pragma unroll 1
      for (int blk = d.blk_begin + tw; blk < body_end; blk += TS) page(blk, std::false_type{}, std::integral_constant<int, 1>{});
      if (has_tail) page(nblk - 1, std::true_type{}, std::integral_constant<int, 1>{});
    }
    const int tix = (warp * gridDim.x + blockIdx.x) * 8;   // trace row of this warp
    if (p.trace && lane == 0) p.trace[tix + 2] = gtimer();

    // ---------------- chunk end
    const int head0 = h * p.G + hg * 8;       // first q head of this group
    const int nheads = min(8, p.G - hg * 8);  // valid q heads in the group
    // l over the 8 row-lanes of each column
#pragma unroll
    for (int sh = 4; sh <= 16; sh <<= 1) {
      l_run[0] += __shfl_xor_sync(kFull, l_run[0], sh);
      l_run[1] += __shfl_xor_sync(kFull, l_run[1], sh);
    }
    // ---- team hand-over through shared memory: members 1 .. n_active-1 pass (m, l, O) in the
    // mma fragment layout ([register][lane]: conflict-free, no cross-lane traffic) to member 0
    if (n_active > 1) {
      constexpr int NR = MT * 4;   // O registers per lane
      if (tw > 0) {
        const int fidx = warp - team_in_cta - 1;   // followers before this one in the CTA
        float* pb = reinterpret_cast<float*>(part_base + (size_t)fidx * T::kPartBytes);
        if (pf_sent > 0) mbar_wait(&pfree[warp], (uint32_t)(pf_sent - 1) & 1u);  // previous hand-over consumed
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
          for (int i = 0; i < 4; ++i) pb[(m * 4 + i) * 32 + lane] = o[m][i];
        if (g == 0) {
          float* ml = pb + NR * 32 + t * 4;
          ml[0] = m_run[0]; ml[1] = l_run[0]; ml[2] = m_run[1]; ml[3] = l_run[1];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&pfull[warp]);
        ++pf_sent;
        continue;
      }
#pragma unroll 1
      for (int f = 1; f < n_active; ++f) {
        const int fw = warp + f;
        const float* pb = reinterpret_cast<const float*>(part_base + (size_t)(fw - team_in_cta - 1) * T::kPartBytes);
        mbar_wait(&pfull[fw], (pf_parity >> f) & 1u);
        pf_parity ^= 1u << f;
        const float* ml = pb + NR * 32 + t * 4;
        const float mf0 = ml[0], lf0 = ml[1], mf1 = ml[2], lf1 = ml[3];
        const float M0 = fmaxf(m_run[0], mf0), M1 = fmaxf(m_run[1], mf1);
        const float a0 = fast_exp2(m_run[0] - M0), b0 = fast_exp2(mf0 - M0);
        const float a1 = fast_exp2(m_run[1] - M1), b1 = fast_exp2(mf1 - M1);
        m_run[0] = M0; m_run[1] = M1;
        l_run[0] = l_run[0] * a0 + lf0 * b0;
        l_run[1] = l_run[1] * a1 + lf1 * b1;
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          o[m][0] = o[m][0] * a0 + pb[(m * 4 + 0) * 32 + lane] * b0;
          o[m][1] = o[m][1] * a1 + pb[(m * 4 + 1) * 32 + lane] * b1;
          o[m][2] = o[m][2] * a0 + pb[(m * 4 + 2) * 32 + lane] * b0;
          o[m][3] = o[m][3] * a1 + pb[(m * 4 + 3) * 32 + lane] * b1;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&pfree[fw]);
      }
    }
    const int n_chunks = (nblk + chunk_pages - 1) / chunk_pages;  // chunks of this group
    // o[m][0]=(dim 16m+g, head 2t) [1]=(dim 16m+g, head 2t+1) [2]=(dim 16m+g+8, 2t) [3]=(.., 2t+1)
    if (n_chunks == 1) {  // the whole group was this chunk: final result
      const float inv0 = 1.f / (l_run[0] + 1e-6f), inv1 = 1.f / (l_run[1] + 1e-6f);
      __nv_bfloat16* ob = p.out + (size_t)b * p.out_stride + (size_t)head0 * D;
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        if (2 * t < nheads) {
          ob[(size_t)(2 * t) * D + 16 * m + g] = __float2bfloat16_rn(o[m][0] * inv0);
          ob[(size_t)(2 * t) * D + 16 * m + g + 8] = __float2bfloat16_rn(o[m][2] * inv0);
        }
        if (2 * t + 1 < nheads) {
          ob[(size_t)(2 * t + 1) * D + 16 * m + g] = __float2bfloat16_rn(o[m][1] * inv1);
          ob[(size_t)(2 * t + 1) * D + 16 * m + g + 8] = __float2bfloat16_rn(o[m][3] * inv1);
        }
      }
      continue;
    }
    // ---- split group: publish the partial, last arriver combines in chunk order
    {
      float* po = p.part + (size_t)d.id * kSlot;
      float* pml = po + 8 * D;
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        po[(size_t)(2 * t) * D + 16 * m + g] = o[m][0];
        po[(size_t)(2 * t) * D + 16 * m + g + 8] = o[m][2];
        po[(size_t)(2 * t + 1) * D + 16 * m + g] = o[m][1];
        po[(size_t)(2 * t + 1) * D + 16 * m + g + 8] = o[m][3];
      }
      if (g == 0) {
        pml[(2 * t) * 2 + 0] = m_run[0];
        pml[(2 * t) * 2 + 1] = l_run[0];
        pml[(2 * t + 1) * 2 + 0] = m_run[1];
        pml[(2 * t + 1) * 2 + 1] = l_run[1];
      }
    }
    __syncwarp();
    int* tk = p.tickets + (size_t)b * p.GH + d.gh;
    int ticket = 0;
    const unsigned long long tf0 = p.trace ? gtimer() : 0ull;
    if (lane == 0) ticket = atom_add_acq_rel(tk, 1);
    ticket = __shfl_sync(kFull, ticket, 0);
    if (p.trace) tr_fix += gtimer() - tf0;
    if (ticket != n_chunks - 1) continue;
    // lane 0's acquire must order the OTHER lanes' loads of the peers' partials too: shuffles
    // are not memory-ordering operations, a warp barrier is (bar.warp.sync orders memory
    // accesses among the participating threads)
    __syncwarp();
    ++tr_combines;
    if (lane == 0) *tk = 0;  // leave the workspace clean for the next launch
    {
      const int id0 = d.id - d.blk_begin / chunk_pages;  // first chunk of the group
      const int hh = lane >> 2;             // head of the group this lane combines
      const int d0 = (lane & 3) * (D / 4);  // its D/4 contiguous dims
      // pass 1: group max per head (lanes sharing a head read the same words: broadcast)
      float M = -INFINITY;
#pragma unroll 8
      for (int s = 0; s < n_chunks; ++s)
        M = fmaxf(M, __ldcg(p.part + (size_t)(id0 + s) * kSlot + 8 * D + hh * 2));
      // pass 2: weighted sum, 4 chunks (32 independent 16-byte loads per lane) in flight
      float L = 0.f;
      float acc[D / 4];
#pragma unroll
      for (int i = 0; i < D / 4; ++i) acc[i] = 0.f;
      constexpr int CU = 4;
      for (int s0 = 0; s0 < n_chunks; s0 += CU) {
        float4 v[CU][D / 16];
        float ms[CU], ls[CU];
#pragma unroll
        for (int u = 0; u < CU; ++u) {
          const int s = min(s0 + u, n_chunks - 1);
          const float* slot_s = p.part + (size_t)(id0 + s) * kSlot;
          ms[u] = __ldcg(slot_s + 8 * D + hh * 2);
          ls[u] = __ldcg(slot_s + 8 * D + hh * 2 + 1);
          const float4* src = reinterpret_cast<const float4*>(slot_s + (size_t)hh * D + d0);
#pragma unroll
          for (int i = 0; i < D / 16; ++i) v[u][i] = __ldcg(src + i);
        }
#pragma unroll
        for (int u = 0; u < CU; ++u) {
          const float wgt = (s0 + u < n_chunks) ? fast_exp2(ms[u] - M) : 0.f;
          L += wgt * ls[u];
#pragma unroll
          for (int i = 0; i < D / 16; ++i) {
            acc[4 * i + 0] += wgt * v[u][i].x;
            acc[4 * i + 1] += wgt * v[u][i].y;
            acc[4 * i + 2] += wgt * v[u][i].z;
            acc[4 * i + 3] += wgt * v[u][i].w;
          }
        }
      }
      if (hh < nheads) {
        const float inv = 1.f / (L + 1e-6f);
        uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)b * p.out_stride + (size_t)(head0 + hh) * D + d0);
#pragma unroll
        for (int i = 0; i < D / 32; ++i) {
          uint4 v;
          v.x = pack_bf16x2(acc[8 * i + 0] * inv, acc[8 * i + 1] * inv);
          v.y = pack_bf16x2(acc[8 * i + 2] * inv, acc[8 * i + 3] * inv);
          v.z = pack_bf16x2(acc[8 * i + 4] * inv, acc[8 * i + 5] * inv);
          v.w = pack_bf16x2(acc[8 * i + 6] * inv, acc[8 * i + 7] * inv);
          dst[i] = v;
        }
      }
    }
  }
  // ---- exit: the last warp of the launch re-arms the chunk counter
  if (lane == 0) {
    const int old = atom_add_acq_rel(&p.counters[1], 1);
    if (old == n_launched - 1) {
      p.counters[0] = 0;
      p.counters[1] = 0;
    }
    if (p.trace) {
      const int tix = (warp * gridDim.x + blockIdx.x) * 8;
      p.trace[tix + 3] = gtimer();
      p.trace[tix + 4] = tr_chunks;
      p.trace[tix + 5] = tr_combines;
      unsigned int smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      p.trace[tix + 6] = smid;
      p.trace[tix + 7] = tr_fix;
    }
  }
}

// ---------------------------------------------------------------------------- host
// Workspace layout (fixed offsets, so the zeroed regions stay zero whatever the batch, and so
// that nothing depends on the chunk plan of one particular launch):
//   [counters: 2 i32, 256 B][tickets: kMaxGroups i32][partials: one slot per chunk id,
//    (8*D + 16) f32 = 8 heads x D of O, then 8 x (m, l)]
constexpr int kMaxGroups = 1 << 20;
constexpr size_t kOffTickets = 256;
constexpr size_t kOffPartials = kOffTickets + (size_t)kMaxGroups * sizeof(int);
static size_t slot_bytes(int D) { return (size_t)(8 * D + 16) * sizeof(float); }
#ifdef PC_DEBUG
static unsigned long long* g_decode_trace = nullptr;
void set_decode_trace(void* buf) { g_decode_trace = static_cast<unsigned long long*>(buf); }
static int g_force_chunk_pages = 0, g_force_team = 0;   // planner overrides (calibration sweeps)
static int g_plan_epoch = 0;
void set_decode_plan_override(int chunk_pages, int team) {
  g_force_chunk_pages = chunk_pages;
  g_force_team = team;
  ++g_plan_epoch;   // invalidates the worst-case memo
}
#else
constexpr int g_plan_epoch = 0;
#endif

struct DecodePlan {
  int chunk_pages;
  int team;
  long long max_chunks;
};

// kernel configuration per (D, BS): the ring (warps x stages x (K + V page)) must fit 227 KB
// ctas = CTAs per SM: with TWO half-size CTAs (4 warps, 96 KiB of rings each) the prologue of the
// next layer's CTA (schedule scan + first table entries + first page latency: 3-4.5 us during which
// an SM with ONE resident CTA pulls nothing) overlaps the streaming of its sibling.
struct DecodeCfg { int warps, stages, ctas; };
// Measured (profiles/r02_decode_two_cta.txt): step 0.940-0.944 ms against 0.947-0.948 at the bench
// shape, kernel alone 28.1 against 27.9 us, slightly slower at evict-freq 4 and at the TP-8 per-rank
// shape -- within noise, so the one-CTA configuration stays and this one exists in the PC_DEBUG
// library only.
#ifdef PC_DEBUG
static int g_decode_two_cta = 0;
void set_decode_two_cta(int v) { g_decode_two_cta = v; ++g_plan_epoch; }   // (invalidates the worst-case memo)
#else
constexpr int g_decode_two_cta = 0;
#endif
static DecodeCfg decode_cfg(int D, int BS) {
  if (D == 128 && BS == 16 && g_decode_two_cta) return {4, 3, 2};
  if (D == 128 && BS == 32) return {4, 3, 1};
  if (D == 64 && BS == 16) return {8, 6, 1};
  if (BS == 8) return {8, 6, 1};      // (128, 8): 4 KiB stages, (64, 8): 2 KiB
  return {8, 3, 1};   // (128, 16), (64, 32)
}
// dynamic shared memory of a launch (mirrors DecodeTraits::smem_bytes)
static size_t decode_smem_bytes(int D, int BS, int B, int team_cap) {
  const DecodeCfg c = decode_cfg(D, BS);
  const size_t ring = (size_t)c.warps * c.stages * 2 * BS * D * 2;
  const size_t followers = team_cap > 1 ? c.warps - c.warps / team_cap : 0;
  return ring + (size_t)c.warps * c.stages * 8 + (size_t)c.warps * kFifo * sizeof(ChunkDesc) + 2 * c.warps * 8 +
         followers * (32 * (D / 4) * 4 + 64) + ((size_t)B + 1) * 4 + 64;
}
constexpr size_t kSmemLimit = 227 * 1024;
static size_t decode_smem_limit(int D, int BS) { return decode_cfg(D, BS).ctas == 2 ? (size_t)113 * 1024 : kSmemLimit; }
// largest team whose hand-over buffers still fit next to the ring and the O(B) schedule
static int decode_team_cap(int D, int BS, int B) {
  const DecodeCfg c = decode_cfg(D, BS);
  for (int t = c.warps; t > 1; t >>= 1)
    if (decode_smem_bytes(D, BS, B, t) <= decode_smem_limit(D, BS)) return t;
  return 1;
}

// The launch plan: chunks per group (n, as chunk_pages) and warps per team (T), picked by a
// small cost model calibrated on B200 (profiles/r02_decode_team_sweep.txt: every (T, n) of a
// grid timed for ten launch shapes; profiles/r01_decode_policy_sweep.txt for the split costs):
//  * one warp streams a page (8 KiB at D 128 / block 16) in ~0.6 us whatever its neighbours do
//    (dependent chain + a 3-deep ring): a launch with fewer chunks than warps is latency-bound
//    unless the idle warps join the busy ones' chunks (teams, ~0.5 us per hand-over);
//  * an SM pulls at most ~77 GB/s, so HBM is saturated only while most SMs stream: chunks
//    are spread over SMs first, and an uneven chunks-per-SM count is paid in full;
//  * every chunk of a SPLIT group (n > 1) costs an fp32 partial (8 heads x D, written and
//    re-read), a release fence + ticket (2-5 us of warp time under load), and the last arriver
//    combines n partials: ~4 + 0.6 n us.  Large launches are NOT split for balance any more:
//    64 x 8192 tokens runs 334 us whole against 370 us in 64-page chunks.
// T > 1 launches are kept to one wave (chunks <= teams: the teams' chunk ids are static);
// T = 1 launches may run several waves off the global chunk counter.
struct PlanModel {
  double hbm_bytes_per_us = 6.4e6;    // whole chip
  double sm_bytes_per_us = 7.7e4;     // one SM
  double page_us_base = 0.25, page_us_per_8k = 0.35;   // one page through one warp
  double split_us = 4.0, split_us_per_chunk = 0.6;     // global split-KV combine
  double team_us = 0.5;
};

static double plan_cost(const PlanModel& m, long long groups, long long max_blocks, long long cp, int T, int sms,
                        int warps, int page_bytes) {
  const long long n = (max_blocks + cp - 1) / cp;
  const long long C = groups * n;
  const long long workers = (long long)sms * warps;
  const double page_us = m.page_us_base + m.page_us_per_8k * page_bytes / 8192.0;
  const long long waves = T == 1 ? (C + workers - 1) / workers : 1;
  // (a split chunk also stalls its warp ~3 us in the release fence + ticket)
  const double t_warp = (double)waves * ((double)((cp + T - 1) / T) * page_us + (n > 1 ? 3.0 : 0.0));
  const double t_sm = (double)((C + sms - 1) / sms) * cp * page_bytes / m.sm_bytes_per_us;
  const long long busy = C < sms ? C : sms;
  double hbm = m.hbm_bytes_per_us * sms / 148.0;
  if (busy * m.sm_bytes_per_us < hbm) hbm = busy * m.sm_bytes_per_us;
  const double t_hbm = (double)groups * max_blocks * page_bytes / hbm;
  // the per-warp chain and the bandwidth bound overlap imperfectly while a launch ramps up
  // and drains: a quarter of the smaller one is exposed, saturating at ~7.5 us
  const double t_bw = t_sm > t_hbm ? t_sm : t_hbm;
  const double lo = t_warp < t_bw ? t_warp : t_bw;
  double t = (t_warp > t_bw ? t_warp : t_bw) + 0.25 * lo * 30.0 / (30.0 + lo);
  if (n > 1) t += m.split_us + m.split_us_per_chunk * n;
  if (T > 1) t += m.team_us;
  return t;
}

static DecodePlan plan_decode(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms, int team_cap) {
  DecodePlan pl;
  const int warps = decode_cfg(D, BS).warps * decode_cfg(D, BS).ctas;   // warps per SM
  const int G = Hq / Hkv;
  const int HG = (G + 7) / 8;
  const long long groups = (long long)B * Hkv * HG;
  const long long max_blocks = max_kv_len > 0 ? (max_kv_len + BS - 1) / BS : 1;
  const long long workers = (long long)num_sms * warps;
  const PlanModel m;
  const int page_bytes = 2 * BS * D * 2;
  long long cp = max_blocks;
  int team = 1;
  double best = plan_cost(m, groups, max_blocks, cp, 1, num_sms, warps, page_bytes);
  for (int T = 1; T <= team_cap && T <= decode_cfg(D, BS).warps; T <<= 1) {
    const long long teams = workers / T;
    long long last_c = -1;
    for (long long nn = 1; nn <= max_blocks; nn = nn < 16 ? nn + 1 : nn + nn / 8) {
      const long long c = (max_blocks + nn - 1) / nn;
      if (c == last_c) continue;
      last_c = c;
      if (nn > 1 && c < 4) break;                                          // never shorter than 4 pages
      if (nn > 1 && c < 2LL * T) break;                                    // >= 2 pages per member when splitting
      const long long C = groups * ((max_blocks + c - 1) / c);
      if (T > 1 && C > teams) break;                                       // one wave only
      if (T == 1 && C > 8 * workers && nn > 1) break;
      const double t = plan_cost(m, groups, max_blocks, c, T, num_sms, warps, page_bytes);
      if (t < best - 1e-9) { best = t; cp = c; team = T; }
    }
  }
#ifdef PC_DEBUG
  if (g_force_chunk_pages > 0) cp = g_force_chunk_pages;
  if (g_force_team > 0) team = g_force_team < team_cap ? g_force_team : team_cap;
#endif
  if (cp > max_blocks) cp = max_blocks;
  if (cp < 1) cp = 1;
  pl.chunk_pages = (int)cp;
  pl.team = team;
  pl.max_chunks = groups * ((max_blocks + cp - 1) / cp);
  return pl;
}

// Chunk slots a workspace must hold so that EVERY plan the planner can produce for a
// longest row of 1 .. max_kv_len tokens fits.  chunks-per-group is not monotone in the
// length (90 blocks -> 9-page chunks -> 10 per group; 88 blocks -> 8-page chunks -> 11), and
// under CUDA-graph replay the plan is read from device memory each step, so sizing by the
// plan of max_kv_len alone is not enough.
static long long worst_case_chunks(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms) {
  struct Key { int B, Hq, Hkv, D, BS, max_kv_len, num_sms, epoch; long long chunks; };
  static thread_local Key memo = {0, 0, 0, 0, 0, 0, 0, -1, 0};
  if (memo.B == B && memo.Hq == Hq && memo.Hkv == Hkv && memo.D == D && memo.BS == BS &&
      memo.max_kv_len == max_kv_len && memo.num_sms == num_sms && memo.epoch == g_plan_epoch)
    return memo.chunks;
  const int max_blocks = max_kv_len > 0 ? (max_kv_len + BS - 1) / BS : 1;
  const int cap = decode_team_cap(D, BS, B);
  long long worst = 1;
  for (int n = 1; n <= max_blocks; ++n) {
    const long long c = plan_decode(B, Hq, Hkv, D, BS, n * BS, num_sms, cap).max_chunks;
    if (c > worst) worst = c;
  }
  memo = {B, Hq, Hkv, D, BS, max_kv_len, num_sms, g_plan_epoch, worst};
  return worst;
}

size_t decode_workspace_bytes(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms) {
  if (Hkv <= 0 || Hq < Hkv) return 0;
  return kOffPartials + (size_t)worst_case_chunks(B, Hq, Hkv, D, BS, max_kv_len, num_sms) * slot_bytes(D);
}
void decode_plan(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, int num_sms, int* out2) {
  out2[0] = out2[1] = 0;
  if (Hkv <= 0 || Hq < Hkv) return;
  const DecodePlan pl = plan_decode(B, Hq, Hkv, D, BS, max_kv_len, num_sms, decode_team_cap(D, BS, B));
  out2[0] = pl.chunk_pages;
  out2[1] = pl.team;
}

template <int D, int BS, int kWarps, int kStages, int kCtasPerSm = 1>
static int launch_decode(const DecodeParams& p, int num_sms, long long max_teams, cudaStream_t stream) {
  using T = DecodeTraits<D, BS, kWarps, kStages>;
  auto kern = paged_decode_attn_score_kernel<D, BS, kWarps, kStages, kCtasPerSm>;
  const size_t smem = T::smem_bytes(p.B, p.team_cap);
  if (smem != decode_smem_bytes(D, BS, p.B, p.team_cap) || decode_cfg(D, BS).warps != kWarps)
    return set_error(PC_ERR_INVALID, "decode: internal shared-memory accounting mismatch");
  if (smem > decode_smem_limit(D, BS)) return set_error(PC_ERR_INVALID, "decode: batch too large for the in-kernel schedule");
  // largest dynamic smem opted in so far, per device (monotone; a racing thread at worst
  // repeats the idempotent attribute call)
  static std::atomic<size_t> smem_set[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || smem_set[dev].load(std::memory_order_acquire) < smem) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_cuda_error("cudaFuncSetAttribute(decode)", e);
    if (dev >= 0 && dev < 64) {
      size_t cur = smem_set[dev].load(std::memory_order_relaxed);
      while (cur < smem && !smem_set[dev].compare_exchange_weak(cur, smem, std::memory_order_release)) {
      }
    }
  }
  const long long slots = (long long)num_sms * kCtasPerSm;
  long long ctas = max_teams < slots ? max_teams : slots;  // spread over SMs first
  if (ctas < 1) ctas = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3(T::kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.pdl ? 1 : 0;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
  if (e != cudaSuccess) return set_cuda_error("paged_decode_attn_score_kernel", e);
  return check_launch("paged_decode_attn_score_kernel");
}

int paged_decode_attn_score(const DecodeArgs& a, cudaStream_t stream) {
  if (a.B <= 0) return 0;
  if (a.Hkv <= 0 || a.Hq % a.Hkv != 0)
    return set_error(PC_ERR_INVALID, "num_heads must be a multiple of num_kv_heads");
  if ((a.out_stride % 8) != 0 || (reinterpret_cast<uintptr_t>(a.out) & 15) != 0 ||
      (a.q_stride % 2) != 0 || (reinterpret_cast<uintptr_t>(a.q) & 3) != 0)
    return set_error(PC_ERR_INVALID, "decode: out rows must be 16-byte aligned, query rows 4-byte aligned");
  const int num_sms = device_sm_count();
  if (num_sms <= 0) return PC_ERR_CUDA;
  if (a.D != 64 && a.D != 128 || a.BS != 8 && a.BS != 16 && a.BS != 32)
    // mirrors TORCH_CHECK(false, "Unsupported head size / block size") in
    // csrc/attention/paged_attention_v1.cu:108-110,160-162
    return set_error(PC_ERR_UNSUPPORTED, "Unsupported head size / block size (supported: head 64|128, block 8|16|32)");
  const int team_cap = decode_team_cap(a.D, a.BS, a.B);
  const DecodePlan pl = plan_decode(a.B, a.Hq, a.Hkv, a.D, a.BS, a.max_kv_len, num_sms, team_cap);
  // with a device-side plan (graph replay) any plan up to max_kv_len may be in force
  const long long need_chunks =
      a.plan ? worst_case_chunks(a.B, a.Hq, a.Hkv, a.D, a.BS, a.max_kv_len, num_sms) : pl.max_chunks;
  if (a.workspace_bytes < kOffPartials + (size_t)need_chunks * slot_bytes(a.D))
    return set_error(PC_ERR_INVALID, "decode workspace too small (see pc_paged_decode_workspace_bytes)");
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  DecodeParams p;
  p.q = static_cast<const __nv_bfloat16*>(a.q);
  p.k_cache = static_cast<const __nv_bfloat16*>(a.k_cache);
  p.v_cache = static_cast<const __nv_bfloat16*>(a.v_cache);
  p.block_tables = a.block_tables;
  p.kv_lens = a.kv_lens;
  p.out = static_cast<__nv_bfloat16*>(a.out);
  p.counters = reinterpret_cast<int*>(ws);
  p.tickets = reinterpret_cast<int*>(ws + kOffTickets);
  p.part = reinterpret_cast<float*>(ws + kOffPartials);
  {
    const size_t cap = (a.workspace_bytes - kOffPartials) / slot_bytes(a.D);
    p.cap_chunks = cap > 0x7fffffffull ? 0x7fffffff : (int)cap;
  }
  p.seq_lens = a.score.n_pools > 0 ? a.seq_lens : nullptr;
  p.score = a.score;
  p.B = a.B; p.Hq = a.Hq; p.Hkv = a.Hkv; p.G = a.Hq / a.Hkv; p.HG = (p.G + 7) / 8; p.GH = a.Hkv * p.HG;
  if ((long long)p.B * p.GH > kMaxGroups)
    return set_error(PC_ERR_INVALID, "decode: batch x kv heads exceeds the ticket table (1M groups)");
  p.q_stride = a.q_stride; p.out_stride = a.out_stride;
  p.table_stride = a.table_stride; p.kv_block_stride = a.kv_block_stride;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  p.chunk_pages = pl.chunk_pages;
  p.team = pl.team;
  p.team_cap = team_cap;
  p.plan = a.plan;
  p.pdl = a.pdl;
#ifdef PC_DEBUG
  p.trace = g_decode_trace;
#else
  p.trace = nullptr;
#endif
  p.new_k = static_cast<const __nv_bfloat16*>(a.new_k);
  p.new_v = static_cast<const __nv_bfloat16*>(a.new_v);
  p.slot_mapping = a.slot_mapping;
  p.k_cache_w = static_cast<__nv_bfloat16*>(const_cast<void*>(a.k_cache));
  p.v_cache_w = static_cast<__nv_bfloat16*>(const_cast<void*>(a.v_cache));
  p.new_k_stride = a.new_k_stride; p.new_v_stride = a.new_v_stride;
  if (p.new_k != nullptr) {
    if (!p.new_v || !p.slot_mapping)
      return set_error(PC_ERR_INVALID, "decode: fused cache write needs key, value and slot_mapping");
    if ((a.new_k_stride % 8) || (a.new_v_stride % 4) || (reinterpret_cast<uintptr_t>(a.new_k) & 15) ||
        (reinterpret_cast<uintptr_t>(a.new_v) & 7))
      return set_error(PC_ERR_INVALID, "decode: fused cache write needs 16-byte aligned key rows, 8-byte aligned value rows");
  }

  // a device-side plan may change between replays of a captured launch: full grid
  const long long grid_teams = a.plan ? (long long)num_sms * decode_cfg(a.D, a.BS).ctas : pl.max_chunks;
#ifdef PC_DEBUG
  if (a.D == 128 && a.BS == 16 && decode_cfg(128, 16).ctas == 2)
    return launch_decode<128, 16, 4, 3, 2>(p, num_sms, grid_teams, stream);
#endif
  if (a.D == 128 && a.BS == 16) return launch_decode<128, 16, 8, 3>(p, num_sms, grid_teams, stream);
  if (a.D == 128 && a.BS == 32) return launch_decode<128, 32, 4, 3>(p, num_sms, grid_teams, stream);
  if (a.D == 64 && a.BS == 16) return launch_decode<64, 16, 8, 6>(p, num_sms, grid_teams, stream);
  if (a.D == 128 && a.BS == 8) return launch_decode<128, 8, 8, 6>(p, num_sms, grid_teams, stream);
  if (a.D == 64 && a.BS == 8) return launch_decode<64, 8, 8, 6>(p, num_sms, grid_teams, stream);
  return launch_decode<64, 32, 8, 3>(p, num_sms, grid_teams, stream);
}

// ---------------------------------------------------------------------------------------
// Scratch for pc_paged_decode_attn_score_static: the split-KV workspace of the kernel above
// (counters, tickets, partials) as a file-scope device array, for callers that cannot allocate.
// PC_DECODE_STATIC_SCRATCH_BYTES (include/pagedevict_b200.h) per GPU that loads the library,
// held until exit.  Every call on a device shares it.
static __device__ __align__(256) unsigned char g_decode_static_scratch[PC_DECODE_STATIC_SCRATCH_BYTES];

// Device address of the array for the CURRENT device (in host code the symbol's name is a
// host-side stand-in, never a device pointer): looked up once per device with that device
// current, kept afterwards.  Checks that this launch fits and clears what the kernel needs at
// zero (chunk counter + the arrival tickets of this launch's groups) on `stream`.  Shapes the
// kernel does not support are left to paged_decode_attn_score's own argument checks.
int decode_static_scratch(int B, int Hq, int Hkv, int D, int BS, int max_kv_len, cudaStream_t stream,
                          void** out_ptr, size_t* out_bytes) {
  static std::atomic<void*> addr[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_cuda_error("cudaGetDevice", e);
  if (dev < 0 || dev >= 64) return set_error(PC_ERR_INVALID, "pc_paged_decode_attn_score_static: device ordinal above 63");
  void* ptr = addr[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    e = cudaGetSymbolAddress(&ptr, g_decode_static_scratch);
    if (e != cudaSuccess) return set_cuda_error("cudaGetSymbolAddress(decode static scratch)", e);
    if (ptr == nullptr || (reinterpret_cast<uintptr_t>(ptr) & 255) != 0)
      return set_error(PC_ERR_CUDA, "pc_paged_decode_attn_score_static: static scratch has no usable device address");
    addr[dev].store(ptr, std::memory_order_release);   // (a racing thread stores the same address)
  }
  *out_ptr = ptr;
  *out_bytes = (size_t)PC_DECODE_STATIC_SCRATCH_BYTES;
  const bool shape_ok = B > 0 && Hkv > 0 && Hq % Hkv == 0 && (D == 64 || D == 128) && (BS == 8 || BS == 16 || BS == 32);
  if (!shape_ok) return 0;
  const int num_sms = device_sm_count();
  if (num_sms <= 0) return PC_ERR_CUDA;
  const DecodePlan pl = plan_decode(B, Hq, Hkv, D, BS, max_kv_len, num_sms, decode_team_cap(D, BS, B));
  const size_t need = kOffPartials + (size_t)pl.max_chunks * slot_bytes(D);
  if (need > (size_t)PC_DECODE_STATIC_SCRATCH_BYTES) {
    char msg[224];
    snprintf(msg, sizeof(msg),
             "pc_paged_decode_attn_score_static: this launch needs %zu bytes of scratch, the static array holds %zu "
             "(PC_DECODE_STATIC_SCRATCH_BYTES); use pc_paged_decode_attn_score with a caller workspace",
             need, (size_t)PC_DECODE_STATIC_SCRATCH_BYTES);
    return set_error(PC_ERR_INVALID, msg);
  }
  const int G = Hq / Hkv;
  long long groups = (long long)B * Hkv * ((G + 7) / 8);
  if (groups > kMaxGroups) groups = kMaxGroups;      // (rejected by paged_decode_attn_score)
  e = cudaMemsetAsync(ptr, 0, kOffTickets + (size_t)groups * sizeof(int), stream);
  if (e != cudaSuccess) return set_cuda_error("cudaMemsetAsync(decode static scratch)", e);
  return 0;
}

}  // namespace pc
