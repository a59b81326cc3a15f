set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest30.log 2>&1
tail -5 $O/r02_pytest30.log
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
PF="timeout 40 python -m benchtools.microbench_prefill"
: > $O/r02_pf30.jsonl
for pass in 1 2; do
for args in "--reqs 16" "--reqs 64" "--reqs 16 --len 2048" "--reqs 64 --len 2048" "--reqs 4 --len 8192 --evict-freq 0"; do
  echo "X0 $args" >> $O/r02_pf30.jsonl; $PF $args --xitem 0 >> $O/r02_pf30.jsonl 2>&1
  echo "X1 $args" >> $O/r02_pf30.jsonl; $PF $args --xitem 1 >> $O/r02_pf30.jsonl 2>&1
done
done
cat $O/r02_pf30.jsonl | cut -c1-60,100-230
