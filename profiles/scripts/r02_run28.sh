cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
for sh in "1 1024" "1 8192" "4 2048" "16 2048" "64 576"; do
  echo "=== B ctx $sh" >> $O/r02_trace28.txt
  timeout 200 python profiles/tools_trace_decode.py $sh >> $O/r02_trace28.txt 2>&1
done
