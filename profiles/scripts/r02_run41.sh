# final validation of the round-2 code: whole GPU suite, default bench line, ncu --set full of the dual-query-tile
# prefill kernel at the bench shape, ncu launch list of the bench command
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 600 python -m pytest tests -m gpu -x -q) > $O/r02_pytest41.log 2>&1
tail -4 $O/r02_pytest41.log
(time timeout 600 python bench.py) > $O/r02_bench41.json 2> $O/r02_bench41.err
tail -c 600 $O/r02_bench41.err; cut -c1-200 $O/r02_bench41.json
timeout 200 ncu --set full --clock-control none --import-source on -k regex:prefill_attn_dq -s 3 -c 1 -f -o $O/r02_prefill_dq_final \
  python -m benchtools.microbench_prefill --reqs 64 --reps 2 > $O/r02_ncu41.log 2>&1
tail -2 $O/r02_ncu41.log
B="--steps 8 --warmup 3 --no-legs --no-cpu-baseline --no-prefill"
python bench.py $B > $O/r02_b41.json 2> $O/r02_b41.err && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:paged_decode|evict_select|cache_write|prefill_attn|zero_cache|token_" -c 6000 --csv --log-file $O/r02_launches41.csv python bench.py $B > $O/r02_ncu41b.log 2>&1
tail -2 $O/r02_ncu41b.log
