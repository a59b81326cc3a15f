set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
(time PC_DEBUG_TWO_CTA=1 timeout 600 python - <<'PY'
import ctypes, subprocess, sys, os
PY
) 2>/dev/null
B="--steps 64 --warmup 8 --no-legs --no-cpu-baseline --no-prefill --no-e2e"
for rep in 1 2; do
python bench.py $B > $O/r02_b29_one_$rep.json 2>> $O/r02_b29.err
python bench.py $B --debug-knob decode_two_cta=1 > $O/r02_b29_two_$rep.json 2>> $O/r02_b29.err
done
python bench.py $B --tp-emulate 8 > $O/r02_b29_tp8_one.json 2>> $O/r02_b29.err
python bench.py $B --tp-emulate 8 --debug-knob decode_two_cta=1 > $O/r02_b29_tp8_two.json 2>> $O/r02_b29.err
python bench.py $B --evict-freq 4 > $O/r02_b29_f4_one.json 2>> $O/r02_b29.err
python bench.py $B --evict-freq 4 --debug-knob decode_two_cta=1 > $O/r02_b29_f4_two.json 2>> $O/r02_b29.err
tail -3 $O/r02_b29.err
