set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
(time PC_TEST_PREFILL_PP=1 timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest23.log 2>&1
tail -15 $O/r02_pytest23.log
PF="timeout 120 python -m benchtools.microbench_prefill"
: > $O/r02_pf23.jsonl
for args in "--reqs 16" "--reqs 64" "--reqs 16 --len 2048" "--reqs 4 --len 8192 --evict-freq 0"; do
  echo "PP $args" >> $O/r02_pf23.jsonl; $PF $args --pp 1 >> $O/r02_pf23.jsonl 2>&1
  echo "TC $args" >> $O/r02_pf23.jsonl; $PF $args --pp 0 >> $O/r02_pf23.jsonl 2>&1
done
cat $O/r02_pf23.jsonl | cut -c1-60,100-230
