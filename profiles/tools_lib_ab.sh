for v in libpagedevict_b200.so var_REFILL.so; do export PC_LIB_PATH=$PWD/vllm_paged_compression_b200/$v; echo "== $v";
for s in "64 592" "64 1024" "64 8192" "1 32768"; do set -- $s; python -m benchtools.microbench --B $1 --ctx $2 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['B'],j['ctx'],round(j['us'],1),round(j['kv_read_GBps']))"; done
python bench.py --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print(j['value'], j['ms_per_step'], j['roofline']['us_per_launch'])
"; done
