set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_layer.py -m gpu -x -q -k "graph_runner or bench_shapes") > $O/r02_pytest31.log 2>&1
tail -4 $O/r02_pytest31.log
B="--steps 64 --warmup 8 --no-legs --no-cpu-baseline --no-prefill"
python bench.py $B > $O/r02_b31.json 2>> $O/r02_b31.err
python bench.py $B --no-e2e --tp-emulate 8 > $O/r02_b31_tp8.json 2>> $O/r02_b31.err
tail -3 $O/r02_b31.err
