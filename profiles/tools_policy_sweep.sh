# usage: bash profiles/tools_policy_sweep.sh  -- decode chunk-policy sweep (env overrides of csrc/decode_attn.cu)
run() { python -m benchtools.microbench --B $1 --ctx $2 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print('  B%d ctx%d %.1f us %.0f GB/s'%(r['B'],r['ctx'],r['us'],r['kv_read_GBps']))"; }
for cp in 10 13 19 37; do echo "== PC_DECODE_CHUNK_PAGES=$cp"; PC_DECODE_CHUNK_PAGES=$cp run 64 592; done
for pw in 4 8 16; do for mp in 16 32 64; do
  echo "== per_warp=$pw min_pages=$mp"
  for s in "1 8192" "4 4096" "16 2048" "64 1024" "64 8192" "256 4096"; do
    PC_DECODE_CHUNKS_PER_WARP=$pw PC_DECODE_MIN_CHUNK_PAGES=$mp run $s
  done
done; done
