set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time python bench.py > $O/r02_b15_n1.json) 2> $O/r02_b15.err
tail -3 $O/r02_b15.err
(time python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_b15_ref.json) 2> $O/r02_b15_ref.err
tail -3 $O/r02_b15_ref.err
