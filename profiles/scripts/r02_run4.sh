set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r02_pytest4.log 2>&1
tail -5 gpurun_out/r02_pytest4.log
python -m benchtools.microbench_cache_write > gpurun_out/r02_cw.jsonl 2>&1
python -m benchtools.microbench_cache_write --Hkv 1 --reqs 64 >> gpurun_out/r02_cw.jsonl 2>&1
python -m benchtools.microbench_cache_write --block 32 >> gpurun_out/r02_cw.jsonl 2>&1
python -m benchtools.microbench_cache_write --len 8192 --reqs 2 --evict-freq 4 >> gpurun_out/r02_cw.jsonl 2>&1
cat gpurun_out/r02_cw.jsonl
B="--steps 20 --warmup 5 --no-cpu-baseline --no-legs --no-prefill"
python bench.py $B > gpurun_out/r02_b4_n1.json 2> gpurun_out/r02_b4.err
python bench.py $B --tp-emulate 8 > gpurun_out/r02_b4_tp8.json 2>> gpurun_out/r02_b4.err
tail -3 gpurun_out/r02_b4.err
python bench.py --sweep > gpurun_out/r02_sweep.jsonl 2> gpurun_out/r02_sweep.err
tail -2 gpurun_out/r02_sweep.err
