set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(time timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "torch_library") > $O/r02_pytest17.log 2>&1
tail -4 $O/r02_pytest17.log
# ---- ragged U(ctx/2, ctx) decode sweep, GQA 4:1 (8B shape) and 8:1
: > $O/r02_sweep_decode_ragged.jsonl
for B in 1 4 16 64 256; do for ctx in 1024 2048 4096 8192 16384 32768; do
  timeout 300 python -m benchtools.microbench --B $B --ctx $ctx --Hq 32 --Hkv 8 --ragged >> $O/r02_sweep_decode_ragged.jsonl 2>> $O/r02_sweep17.err
done; done
for B in 16 64 256; do for ctx in 2048 8192; do
  timeout 300 python -m benchtools.microbench --B $B --ctx $ctx --Hq 64 --Hkv 8 --ragged >> $O/r02_sweep_decode_ragged.jsonl 2>> $O/r02_sweep17.err
done; done
wc -l $O/r02_sweep_decode_ragged.jsonl
# ---- ncu launch list of the bench command (after it exited 0 without ncu)
B="--steps 8 --warmup 3 --no-legs --no-cpu-baseline --no-prefill"
python bench.py $B > $O/r02_b17.json 2> $O/r02_b17.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/r02_launches.csv python bench.py $B > $O/r02_ncu17.log 2>&1
tail -2 $O/r02_ncu17.log
# ---- ncu --set full: decode kernel in the step, stream cache write, prefill (64 prompts)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_decode_attn -s 400 -c 3 -f -o $O/r02_decode_final python bench.py --steps 8 --warmup 3 --no-legs --no-cpu-baseline --no-prefill --no-e2e > $O/r02_ncu17_dec.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cache_write_stream -s 6 -c 2 -f -o $O/r02_cache_write_stream python -m benchtools.microbench_cache_write --reqs 64 --reps 6 > $O/r02_ncu17_cw.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_attn_tc -s 4 -c 1 -f -o $O/r02_prefill_b64_final python -m benchtools.microbench_prefill --reqs 64 --reps 3 > $O/r02_ncu17_pf.log 2>&1
ls -la $O/*.ncu-rep
