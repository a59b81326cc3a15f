set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
PF="timeout 600 python -m benchtools.microbench_prefill"
: > $O/r02_pf12.jsonl
$PF --reqs 16 --modes >> $O/r02_pf12.jsonl 2>&1
$PF --reqs 64 --modes >> $O/r02_pf12.jsonl 2>&1
$PF --reqs 64 --len 2048 --modes >> $O/r02_pf12.jsonl 2>&1
unset PC_LIB_PATH
cat $O/r02_pf12.jsonl | cut -c1-260
timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_attn_tc -s 4 -c 1 -f -o $O/r02_prefill_b64 python -m benchtools.microbench_prefill --reqs 64 --reps 3 > $O/r02_ncu_pf12.log 2>&1
tail -2 $O/r02_ncu_pf12.log
