set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r02_pytest3.log 2>&1
tail -5 gpurun_out/r02_pytest3.log
B="--steps 20 --warmup 5 --no-cpu-baseline"
python bench.py $B > gpurun_out/r02_b3_n1.json 2> gpurun_out/r02_b3.err
python bench.py $B --no-legs --no-prefill --tp-emulate 8 > gpurun_out/r02_b3_tp8.json 2>> gpurun_out/r02_b3.err
python bench.py --steps 64 --warmup 8 --no-legs --no-prefill --no-cpu-baseline > gpurun_out/r02_b3_n1_k64.json 2>> gpurun_out/r02_b3.err
tail -5 gpurun_out/r02_b3.err
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
timeout 900 python profiles/tools_team_sweep.py 64,592,32,8 64,848,32,8 64,1056,32,8 512,592,4,1 16,2048,32,8 16,1024,32,8 4,4096,32,8 1,32768,32,8 1,4096,32,8 64,8192,32,8 > gpurun_out/r02_team_sweep.txt 2>&1
