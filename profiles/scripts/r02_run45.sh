# K/V producer of the dual-query-tile kernel with K one tile ahead of V, against the validated library
# (libpagedevict_b200_prev.so): prefill parity tests + same-box alternating A/B
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
(timeout 100 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layer.py -m gpu -x -q -k "prefill or cfg2") > $O/r02_pytest45.log 2>&1
tail -2 $O/r02_pytest45.log
PF="timeout 40 python -m benchtools.microbench_prefill"
: > $O/r02_pf45.jsonl
for pass in 1 2; do
for args in "--reqs 64" "--reqs 64 --len 2048"; do
  echo "PREV $args" >> $O/r02_pf45.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_prev.so $PF $args >> $O/r02_pf45.jsonl 2>&1
  echo "NEW $args" >> $O/r02_pf45.jsonl; $PF $args >> $O/r02_pf45.jsonl 2>&1
done
done
grep -v "^Traceback\|^  " $O/r02_pf45.jsonl | cut -c1-60,150-200
