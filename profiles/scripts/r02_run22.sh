set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
PF="timeout 600 python -m benchtools.microbench_prefill"
: > $O/r02_pf22.jsonl
for rep in 1 2; do
for args in "--reqs 16" "--reqs 64" "--reqs 16 --len 2048" "--reqs 64 --len 2048" "--reqs 4 --len 8192 --evict-freq 0"; do
  echo "NEW $args" >> $O/r02_pf22.jsonl
  $PF $args >> $O/r02_pf22.jsonl 2>&1
  echo "R1 $args" >> $O/r02_pf22.jsonl
  PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_r1prefill.so $PF $args >> $O/r02_pf22.jsonl 2>&1
done; done
