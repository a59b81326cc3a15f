# clock trace of CTA 0 of the dual-query-tile kernel (PC_DEBUG library), the changed parity cases, and the
# product library (K/V producer descriptor prefetch, rule <= 6144 tokens) against the saved previous one
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
(time timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest37.log 2>&1
tail -3 $O/r02_pytest37.log
PC_LIB_PATH=$D/libpagedevict_b200_debug.so timeout 120 python profiles/tools_trace_prefill_dq.py --reqs 64 --len 1024 > $O/r02_prefill_dq_trace_1024.txt 2>&1
PC_LIB_PATH=$D/libpagedevict_b200_debug.so timeout 120 python profiles/tools_trace_prefill_dq.py --reqs 256 --len 256 --dump-items 3 > $O/r02_prefill_dq_trace_256.txt 2>&1
head -75 $O/r02_prefill_dq_trace_1024.txt
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf37.jsonl
for args in "--reqs 64" "--reqs 256 --len 256" "--reqs 64 --len 2048"; do
  echo "PREV $args" >> $O/r02_pf37.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_prev.so $PF $args >> $O/r02_pf37.jsonl 2>&1
  echo "NEW $args" >> $O/r02_pf37.jsonl; $PF $args >> $O/r02_pf37.jsonl 2>&1
done
grep -v "^Traceback\|^  " $O/r02_pf37.jsonl | cut -c1-60,150-200
