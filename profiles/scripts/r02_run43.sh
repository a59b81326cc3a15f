# next item's first K / V tile prefetched into L2 by the dual-query-tile kernel's Q producer, against the previous
# commit (libpagedevict_b200_prev.so); then the whole GPU suite on the new library
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf43.jsonl
for pass in 1 2; do
for args in "--reqs 64" "--reqs 256 --len 256" "--reqs 64 --len 2048"; do
  echo "PREV $args" >> $O/r02_pf43.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_prev.so $PF $args >> $O/r02_pf43.jsonl 2>&1
  echo "NEW $args" >> $O/r02_pf43.jsonl; $PF $args >> $O/r02_pf43.jsonl 2>&1
done
done
grep -v "^Traceback\|^  " $O/r02_pf43.jsonl | cut -c1-60,150-200
(time timeout 600 python -m pytest tests -m gpu -x -q) > $O/r02_pytest43.log 2>&1
tail -4 $O/r02_pytest43.log
