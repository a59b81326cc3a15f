set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prune_prompt_and_cache or reshape_and_cache") > $O/r02_pytest11.log 2>&1
tail -5 $O/r02_pytest11.log
CW="timeout 300 python -m benchtools.microbench_cache_write"
: > $O/r02_cw11.jsonl
$CW >> $O/r02_cw11.jsonl 2>&1
$CW --reqs 64 >> $O/r02_cw11.jsonl 2>&1
$CW --len 8192 --reqs 8 --evict-freq 4 >> $O/r02_cw11.jsonl 2>&1
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
for hc in -1 8; do
  $CW --Hkv 1 --reqs 512 --hc $hc >> $O/r02_cw11.jsonl 2>&1
  $CW --Hkv 2 --reqs 256 --hc $hc >> $O/r02_cw11.jsonl 2>&1
  $CW --Hkv 4 --reqs 128 --hc $hc >> $O/r02_cw11.jsonl 2>&1
  $CW --D 64 --reqs 64 --hc $hc >> $O/r02_cw11.jsonl 2>&1
  $CW --block 32 --reqs 64 --hc $hc >> $O/r02_cw11.jsonl 2>&1
  $CW --block 32 --D 64 --reqs 64 --hc $hc >> $O/r02_cw11.jsonl 2>&1
done
unset PC_LIB_PATH
