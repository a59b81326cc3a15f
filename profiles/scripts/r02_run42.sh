# final prefill shape table (product library, its own dispatch rule) + the reference arm line of the final code
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
PF="timeout 60 python -m benchtools.microbench_prefill"
: > $O/r02_pf42.jsonl
for args in "--reqs 256 --len 256" "--reqs 64 --len 512" "--reqs 16" "--reqs 64" "--reqs 16 --len 2048" "--reqs 64 --len 2048" "--reqs 16 --len 4096" "--reqs 8 --len 6144" "--reqs 4 --len 8192 --evict-freq 0" "--reqs 8 --len 8192" "--reqs 64 --D 64" "--reqs 64 --len 2048 --Hq 8 --Hkv 1" "--reqs 64 --evict-freq 4" "--reqs 64 --evict-freq 0"; do
  echo "$args" >> $O/r02_pf42.jsonl; $PF $args >> $O/r02_pf42.jsonl 2>&1
done
grep -v "^Traceback\|^  " $O/r02_pf42.jsonl | cut -c1-60,150-215
(time timeout 300 python bench.py --impl reference) > $O/r02_bench42_reference.json 2> $O/r02_bench42_reference.err
cut -c1-300 $O/r02_bench42_reference.json
