# product library with the dual-query-tile prefill kernel on its rule: whole GPU suite, same-box A/B around the
# 2048-token threshold (PC_DEBUG library), the default bench line, ncu --set full of the kernel at the bench shape
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 600 python -m pytest tests -m gpu -x -q) > $O/r02_pytest35.log 2>&1
tail -4 $O/r02_pytest35.log
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf35.jsonl
for args in "--reqs 64" "--reqs 16 --len 3072" "--reqs 16 --len 4096" "--reqs 64 --len 512" "--reqs 256 --len 256" "--reqs 32 --Hq 8 --Hkv 1 --len 2048"; do
  for dq in 0 1; do
    echo "dq=$dq $args" >> $O/r02_pf35.jsonl
    PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so $PF $args --dq $dq >> $O/r02_pf35.jsonl 2>&1
  done
done
grep -v "^Traceback\|^  " $O/r02_pf35.jsonl | cut -c1-60,150-215
(time timeout 600 python bench.py) > $O/r02_bench35.json 2> $O/r02_bench35.err
tail -c 1500 $O/r02_bench35.err; cut -c1-300 $O/r02_bench35.json
timeout 300 ncu --set full --clock-control none --import-source on -k regex:prefill_attn_dq -s 3 -c 1 -f -o $O/r02_prefill_dq \
  python -m benchtools.microbench_prefill --reqs 64 --reps 2 > $O/r02_ncu35.log 2>&1
tail -3 $O/r02_ncu35.log; ls -la $O/*.ncu-rep
