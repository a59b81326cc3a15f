set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill") > $O/r02_pytest13.log 2>&1
tail -5 $O/r02_pytest13.log
PF="timeout 600 python -m benchtools.microbench_prefill"
: > $O/r02_pf13.jsonl
$PF --reqs 16 >> $O/r02_pf13.jsonl 2>&1
$PF --reqs 64 >> $O/r02_pf13.jsonl 2>&1
$PF --reqs 16 --len 2048 >> $O/r02_pf13.jsonl 2>&1
$PF --reqs 4 --len 8192 --evict-freq 0 >> $O/r02_pf13.jsonl 2>&1
$PF --reqs 16 --D 64 >> $O/r02_pf13.jsonl 2>&1
cut -c1-200 $O/r02_pf13.jsonl
