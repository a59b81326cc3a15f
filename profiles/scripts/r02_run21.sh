set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest21.log 2>&1
tail -4 $O/r02_pytest21.log
PF="timeout 600 python -m benchtools.microbench_prefill"
: > $O/r02_pf21.jsonl
$PF --reqs 16 >> $O/r02_pf21.jsonl 2>&1
$PF --reqs 64 >> $O/r02_pf21.jsonl 2>&1
$PF --reqs 64 --len 2048 >> $O/r02_pf21.jsonl 2>&1
$PF --reqs 4 --len 8192 --evict-freq 0 >> $O/r02_pf21.jsonl 2>&1
cut -c100-230 $O/r02_pf21.jsonl
