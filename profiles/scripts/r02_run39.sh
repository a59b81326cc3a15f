# division-free item iterator in the dual-query-tile kernel (+ head size 64 of the one-tile kernel back on lane 0)
# against the previous commit (libpagedevict_b200_prev.so): prefill parity tests, same-box A/B, clock trace
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
(time timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layer.py -m gpu -x -q -k "prefill or prune or cfg") > $O/r02_pytest39.log 2>&1
tail -3 $O/r02_pytest39.log
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf39.jsonl
for pass in 1 2; do
for args in "--reqs 64" "--reqs 256 --len 256" "--reqs 64 --len 2048" "--reqs 16 --D 64" "--reqs 64 --len 512"; do
  echo "PREV $args" >> $O/r02_pf39.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_prev.so $PF $args >> $O/r02_pf39.jsonl 2>&1
  echo "NEW $args" >> $O/r02_pf39.jsonl; $PF $args >> $O/r02_pf39.jsonl 2>&1
done
done
grep -v "^Traceback\|^  " $O/r02_pf39.jsonl | cut -c1-60,150-200
PC_LIB_PATH=$D/libpagedevict_b200_debug.so timeout 120 python profiles/tools_trace_prefill_dq.py --reqs 64 --len 1024 > $O/r02_prefill_dq_trace_1024_iter.txt 2>&1
head -60 $O/r02_prefill_dq_trace_1024_iter.txt
