# dual-query-tile kernel with two Q sets + 3/3 K/V stages against the 1 Q set + 5/5 version (saved as
# libpagedevict_b200_prev.so): prefill parity tests first, then a same-box alternating A/B
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layer.py -m gpu -x -q -k "prefill or layer or prune") > $O/r02_pytest36.log 2>&1
tail -4 $O/r02_pytest36.log
PF="timeout 60 python -m benchtools.microbench_prefill"
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
: > $O/r02_pf36.jsonl
for pass in 1 2; do
for args in "--reqs 64" "--reqs 16" "--reqs 64 --len 512" "--reqs 256 --len 256" "--reqs 64 --len 2048" "--reqs 16 --len 4096"; do
  echo "PREV $args" >> $O/r02_pf36.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_prev.so $PF $args >> $O/r02_pf36.jsonl 2>&1
  echo "NEW $args" >> $O/r02_pf36.jsonl; $PF $args >> $O/r02_pf36.jsonl 2>&1
done
done
for args in "--reqs 4 --len 8192 --evict-freq 0" "--reqs 8 --len 6144"; do
  for dq in 0 1; do
    echo "DBG dq=$dq $args" >> $O/r02_pf36.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_debug.so $PF $args --dq $dq >> $O/r02_pf36.jsonl 2>&1
  done
done
grep -v "^Traceback\|^  " $O/r02_pf36.jsonl | cut -c1-60,150-200
