# usage (GPU box): bash profiles/tools_cfg_ab.sh "0 1 2 3"  -> microbench + bench step per PC_DECODE_CFG
for c in $1; do export PC_DECODE_CFG=$c; echo "== cfg $c";
for s in "64 592" "64 1024" "64 8192" "1 32768" "16 4096" "256 4096"; do set -- $s; python -m vllm_paged_compression_b200.microbench --B $1 --ctx $2 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['B'],j['ctx'],round(j['us'],1),round(j['kv_read_GBps']))"; done
python bench.py --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print(j['value'], j['ms_per_step'], j['roofline']['us_per_launch'])
    else: print(l[:200])
"; done
