set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L
(time timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q) > $O/r02_pytest16_tp.log 2>&1
tail -5 $O/r02_pytest16_tp.log
(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > $O/r02_b16_n2.json) 2> $O/r02_b16_n2.err
tail -5 $O/r02_b16_n2.err
(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/r02_b16_n2_ref.json) 2> $O/r02_b16_n2_ref.err
tail -3 $O/r02_b16_n2_ref.err
