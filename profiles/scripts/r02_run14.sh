set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > $O/r02_pytest14.log 2>&1
tail -5 $O/r02_pytest14.log
(time python bench.py > $O/r02_b14_n1.json) 2> $O/r02_b14.err
tail -3 $O/r02_b14.err
(time python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_b14_ref.json) 2> $O/r02_b14_ref.err
tail -3 $O/r02_b14_ref.err
