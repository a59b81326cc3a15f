# compute-sanitizer over a reduced -m gpu subset: the kernels with hand-rolled synchronisation
# (decode: acq_rel tickets, in-kernel counter re-arm, generic->async proxy fences, team hand-over
# through shared-memory mbarriers; prefill: tcgen05 / TMA / mbarrier pipeline; cache write: TMA + bulk
# store; evict_select; token prune; zero slots)
set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
SUB='test_paged_decode_streamk_schedule or test_decode_layer_fused_cache_write or test_fused_last_block_score_bit_exact or test_prefill_attention_many_short_items_persistent_walk or test_prefill_attention_tile_edges_and_strided_rows or test_prune_prompt_and_cache or test_evict_select_compact_layers_matches_per_layer or test_prompt_token_prune_bit_exact or test_paged_decode_empty_sequence_and_determinism'
for tool in memcheck racecheck synccheck; do
  (time timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --launch-timeout 0 \
     python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -p no:cacheprovider -k "$SUB") > $O/r02_sanitizer_$tool.log 2>&1
  echo "rc=$?" >> $O/r02_sanitizer_$tool.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|rc=" $O/r02_sanitizer_$tool.log | tail -5
done
