cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q) > $O/r02_pytest33_alone.log 2>&1
tail -6 $O/r02_pytest33_alone.log
(time timeout 900 python -m pytest tests -m gpu -x -q) > $O/r02_pytest33_all.log 2>&1
tail -6 $O/r02_pytest33_all.log
