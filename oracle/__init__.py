"""CPU oracle for the PagedCompression hot path.  TEST INFRASTRUCTURE ONLY.

Everything in this package restates (in numpy / torch-CPU / plain C) the algorithm
of the reference fork krishnateja95/vllm-Paged-Compression for ONE path: prompt-time
streamingLLM eviction -> reshape_and_cache -> decode-time value_l2 scoring + block
eviction + block-table compaction -> paged decode attention.  Each function cites
the reference file:line it follows.

Who may import this package: tests/, __graft_entry__.smoke(), and bench.py's
`cpu_baseline` / `--impl reference` legs -- as the CHECKER or the timed CPU baseline,
never as the thing shipped.  Nothing under vllm_paged_compression_b200/ imports it;
the product fails loudly when its CUDA library is missing.

Pinning status (see DESIGN.md "Oracle"):
  * integer domain (pruned lengths, gather indices, slot mappings, decode triggers,
    streamingLLM victims, compaction, host block-table bookkeeping): PINNED against
    the reference's own Python classes executed in the build container; vectors in
    tests/golden/ref_policy.json (generator: tests/golden/make_golden.py).
  * reshape_and_cache / paged_attention_v1: PINNED against the reference's own CPU
    kernels (csrc/cpu/{cache,attention}.cpp) compiled to oracle/_ref/ and against
    tests/golden/ref_ops.npz produced from them.
  * value_l2 decode scoring: the reference's L2 branch is dead code with layout /
    dtype bugs; the oracle implements the intended fp32
    semantics with a FIXED reduction order.  Pinned only to tolerance against
    KVCachePruner.get_block_score evaluated in fp32 on a correctly de-paged V block
    ("parity partially unpinned" for score bits; argmin/tie-break is pinned).
"""
