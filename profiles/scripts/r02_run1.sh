set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
lscpu | grep -E "^CPU\(s\)|NUMA|Model name" 
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r02_pytest1.log 2>&1
tail -5 gpurun_out/r02_pytest1.log
python profiles/tools_pcie.py > gpurun_out/r02_pcie.txt 2>&1
python profiles/tools_pcie.py --bind >> gpurun_out/r02_pcie.txt 2>&1
cat gpurun_out/r02_pcie.txt
(time python bench.py --steps 20 --warmup 5) > gpurun_out/r02_bench1.json 2> gpurun_out/r02_bench1.err
tail -3 gpurun_out/r02_bench1.err
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
python profiles/tools_trace_decode.py 64 592 32 8 > gpurun_out/r02_trace_b64.txt 2>&1
python profiles/tools_trace_decode.py 512 592 4 1 > gpurun_out/r02_trace_b512.txt 2>&1
python profiles/tools_trace_decode.py 256 592 8 2 > gpurun_out/r02_trace_b256.txt 2>&1
unset PC_LIB_PATH
for s in "64 592 32 8" "128 592 16 4" "256 592 8 2" "512 592 4 1" "512 592 8 1" "16 2048 32 8" "4 4096 32 8"; do set -- $s; python -m benchtools.microbench --B $1 --ctx $2 --Hq $3 --Hkv $4; done > gpurun_out/r02_micro1.jsonl 2>&1
