set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m benchtools.microbench_block_ops > $O/r02_block_ops.jsonl 2> $O/r02_block_ops.err
cat $O/r02_block_ops.jsonl; tail -3 $O/r02_block_ops.err
: > $O/r02_block8.jsonl
for bs in 8 16 32; do
  timeout 300 python -m benchtools.microbench --B 64 --ctx 1024 --block $bs >> $O/r02_block8.jsonl 2>>$O/r02_block8.err
  timeout 300 python -m benchtools.microbench --B 64 --ctx 8192 --block $bs >> $O/r02_block8.jsonl 2>>$O/r02_block8.err
  timeout 300 python -m benchtools.microbench_cache_write --reqs 64 --block $bs >> $O/r02_block8.jsonl 2>>$O/r02_block8.err
done
cut -c1-260 $O/r02_block8.jsonl; tail -3 $O/r02_block8.err
