cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
(time PC_TEST_PREFILL_DQ=1 timeout 150 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest34.log 2>&1
tail -5 $O/r02_pytest34.log
PF="timeout 40 python -m benchtools.microbench_prefill"
: > $O/r02_pf34.jsonl
for pass in 1 2; do
for args in "--reqs 16" "--reqs 64" "--reqs 16 --len 2048" "--reqs 64 --len 2048" "--reqs 4 --len 8192 --evict-freq 0" "--reqs 16 --D 64"; do
  echo "TC $args" >> $O/r02_pf34.jsonl; $PF $args --dq 0 >> $O/r02_pf34.jsonl 2>&1
  echo "DQ $args" >> $O/r02_pf34.jsonl; $PF $args --dq 1 >> $O/r02_pf34.jsonl 2>&1
done
done
grep -v "^Traceback\|^  " $O/r02_pf34.jsonl | cut -c1-40,140-200
