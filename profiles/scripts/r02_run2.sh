set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r02_pytest2.log 2>&1
tail -5 gpurun_out/r02_pytest2.log
export PC_LIB_PATH=$GRAFT_REPO_ROOT/vllm_paged_compression_b200/libpagedevict_b200_debug.so
timeout 900 python profiles/tools_team_sweep.py 64,592,32,8 64,848,32,8 64,1056,32,8 512,592,4,1 256,592,8,2 16,2048,32,8 16,1024,32,8 4,4096,32,8 1,32768,32,8 1,4096,32,8 > gpurun_out/r02_team_sweep.txt 2>&1
unset PC_LIB_PATH
B="--steps 20 --warmup 5 --no-legs --no-cpu-baseline --no-prefill --no-e2e"
python bench.py $B > gpurun_out/r02_b2_n1.json 2> gpurun_out/r02_b2_n1.err
python bench.py $B --no-pdl > gpurun_out/r02_b2_n1_nopdl.json 2>> gpurun_out/r02_b2_n1.err
for n in 2 4 8; do
python bench.py $B --tp-emulate $n > gpurun_out/r02_b2_tp$n.json 2>> gpurun_out/r02_b2_n1.err
python bench.py $B --tp-emulate $n --force-pdl > gpurun_out/r02_b2_tp${n}_pdl.json 2>> gpurun_out/r02_b2_n1.err
python bench.py $B --tp-emulate $n --no-pdl > gpurun_out/r02_b2_tp${n}_nopdl.json 2>> gpurun_out/r02_b2_n1.err
done
tail -5 gpurun_out/r02_b2_n1.err
