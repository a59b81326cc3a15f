set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "decode or fused") > $O/r02_pytest30.log 2>&1
tail -4 $O/r02_pytest30.log
B="--steps 64 --warmup 8 --no-legs --no-cpu-baseline --no-prefill --no-e2e"
python bench.py $B > $O/r02_b30_a.json 2>> $O/r02_b30.err
python bench.py $B > $O/r02_b30_b.json 2>> $O/r02_b30.err
python bench.py $B --tp-emulate 8 > $O/r02_b30_tp8.json 2>> $O/r02_b30.err
tail -3 $O/r02_b30.err
