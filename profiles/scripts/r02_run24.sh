set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layer.py -m gpu -x -q -k "8 or bs8 or prefill") > $O/r02_pytest24.log 2>&1
tail -25 $O/r02_pytest24.log
