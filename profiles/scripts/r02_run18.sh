set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
B="--steps 8 --warmup 3 --no-legs --no-cpu-baseline --no-prefill"
python bench.py $B > $O/r02_b18.json 2> $O/r02_b18.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:paged_decode|evict_select|cache_write|prefill_attn|zero_cache|token_" -c 6000 --csv --log-file $O/r02_launches.csv python bench.py $B > $O/r02_ncu18.log 2>&1
tail -2 $O/r02_ncu18.log
