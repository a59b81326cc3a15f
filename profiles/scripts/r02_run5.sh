set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prune_prompt_and_cache or reshape_and_cache") > $O/r02_pytest5.log 2>&1
tail -3 $O/r02_pytest5.log
# ---- cache write: product rule at the prefill batch sizes, then heads-per-CTA sweep (debug lib)
CW="python -m benchtools.microbench_cache_write"
: > $O/r02_cw5.jsonl
$CW >> $O/r02_cw5.jsonl 2>&1
$CW --reqs 64 >> $O/r02_cw5.jsonl 2>&1
$CW --reqs 64 --no-score >> $O/r02_cw5.jsonl 2>&1
$CW --Hkv 1 --reqs 64 >> $O/r02_cw5.jsonl 2>&1
$CW --Hkv 1 --reqs 512 >> $O/r02_cw5.jsonl 2>&1
$CW --block 32 --reqs 64 >> $O/r02_cw5.jsonl 2>&1
$CW --len 8192 --reqs 8 --evict-freq 4 >> $O/r02_cw5.jsonl 2>&1
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
for hc in 1 2 4 8; do
  $CW --hc $hc >> $O/r02_cw5.jsonl 2>&1
  $CW --hc $hc --reqs 64 >> $O/r02_cw5.jsonl 2>&1
done
unset PC_LIB_PATH
cat $O/r02_cw5.jsonl
# ---- e2e step anatomy
timeout 900 python profiles/tools_e2e_breakdown.py --io-group 0,1,2,4 > $O/r02_e2e_breakdown.jsonl 2> $O/r02_e2e_breakdown.err
cat $O/r02_e2e_breakdown.jsonl; tail -3 $O/r02_e2e_breakdown.err
# ---- ncu --set full of the decode kernel at the three per-rank launch shapes
B="--steps 8 --warmup 3 --no-legs --no-cpu-baseline --no-prefill --no-e2e --no-strong"
for n in 1 4 8; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_decode_attn -s 400 -c 3 \
     -f -o $O/r02_decode_tp$n python bench.py $B --tp-emulate $n > $O/r02_ncu_tp$n.log 2>&1
  tail -2 $O/r02_ncu_tp$n.log
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cache_write_block -s 20 -c 2 \
     -f -o $O/r02_cache_write python -m benchtools.microbench_cache_write --reqs 64 > $O/r02_ncu_cw.log 2>&1
ls -la $O/*.ncu-rep
