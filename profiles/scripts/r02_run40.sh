# (a) decode kernel with the page issue under elect.sync (libpagedevict_b200_decelect.so, -DPC_DECODE_ELECT) against the
# product library: decode parity tests on the variant, same-box alternating A/B at the bench / TP-8 / small shapes;
# (b) dual-query-tile prefill kernel at head size 64 (PC_DEBUG library, --dq 1) against the one-tile kernel
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
D=$GRAFT_REPO_ROOT/vllm_paged_compression_b200
(time PC_LIB_PATH=$D/libpagedevict_b200_decelect.so timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "decode") > $O/r02_pytest40.log 2>&1
tail -3 $O/r02_pytest40.log
MB="timeout 60 python -m benchtools.microbench"
: > $O/r02_dec40.jsonl
for pass in 1 2; do
for args in "--B 64 --ctx 592" "--B 512 --ctx 592 --Hq 4 --Hkv 1" "--B 64 --ctx 864" "--B 16 --ctx 2048" "--B 256 --ctx 4096 --Hq 64 --Hkv 8"; do
  echo "PROD $args" >> $O/r02_dec40.jsonl; $MB $args >> $O/r02_dec40.jsonl 2>&1
  echo "ELECT $args" >> $O/r02_dec40.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_decelect.so $MB $args >> $O/r02_dec40.jsonl 2>&1
done
done
grep -v "^Traceback\|^  " $O/r02_dec40.jsonl | cut -c1-50,95-190
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf40.jsonl
for args in "--reqs 16 --D 64" "--reqs 64 --D 64" "--reqs 64 --D 64 --len 2048" "--reqs 64 --D 64 --len 256"; do
  for dq in 0 1; do
    echo "DBG dq=$dq $args" >> $O/r02_pf40.jsonl; PC_LIB_PATH=$D/libpagedevict_b200_debug.so $PF $args --dq $dq >> $O/r02_pf40.jsonl 2>&1
  done
done
grep -v "^Traceback\|^  " $O/r02_pf40.jsonl | cut -c1-60,150-200
