set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
O=gpurun_out
(timeout 900 python -m pytest tests/test_gpu_layer.py -m gpu -x -q -k "decode_graph_runner_matches or pinned_layer_io") > $O/r02_pytest6.log 2>&1
tail -3 $O/r02_pytest6.log
T="timeout 900 python profiles/tools_e2e_breakdown.py"
: > $O/r02_e2e_breakdown6.jsonl
$T --io-group 0 >> $O/r02_e2e_breakdown6.jsonl 2>> $O/r02_e2e_breakdown6.err
$T --io-group 0 --after-value >> $O/r02_e2e_breakdown6.jsonl 2>> $O/r02_e2e_breakdown6.err
$T --io-group 0 --no-io-prefetch >> $O/r02_e2e_breakdown6.jsonl 2>> $O/r02_e2e_breakdown6.err
$T --io-group 0 --no-io-prefetch --after-value >> $O/r02_e2e_breakdown6.jsonl 2>> $O/r02_e2e_breakdown6.err
$T --io-group 8 >> $O/r02_e2e_breakdown6.jsonl 2>> $O/r02_e2e_breakdown6.err
cat $O/r02_e2e_breakdown6.jsonl; tail -3 $O/r02_e2e_breakdown6.err
B="--steps 20 --warmup 5 --no-cpu-baseline --no-legs --no-prefill"
python bench.py $B > $O/r02_b6_n1.json 2> $O/r02_b6.err
python bench.py $B --tp-emulate 8 > $O/r02_b6_tp8.json 2>> $O/r02_b6.err
tail -3 $O/r02_b6.err
