set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_layer.py -m gpu -x -q -k "graph_runner or bench_shapes") > $O/r02_pytest25.log 2>&1
tail -5 $O/r02_pytest25.log
