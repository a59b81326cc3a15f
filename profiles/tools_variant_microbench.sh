# usage (GPU box): bash profiles/tools_variant_microbench.sh var_A.so var_B.so ...   ("old" = old_wt/ tree)
run() { for s in "64 592" "64 1024" "64 8192" "1 32768" "256 4096" "16 4096"; do set -- $s; python -m benchtools.microbench --B $1 --ctx $2 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['B'],j['ctx'],round(j['us'],1),round(j['kv_read_GBps']))"; done; }
R=$PWD
for v in "$@"; do
  echo "== $v"
  if [ "$v" = old ]; then (cd old_wt && unset PC_LIB_PATH && run); else PC_LIB_PATH=$R/vllm_paged_compression_b200/$v run; fi
done
