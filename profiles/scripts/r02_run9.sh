set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layer.py -m gpu -x -q -k "token_prune or tokenprune or head_policies") > $O/r02_pytest9.log 2>&1
tail -30 $O/r02_pytest9.log
