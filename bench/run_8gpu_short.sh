#!/bin/bash
# The short 8-GPU session: multi-GPU tests, headline bench (eager baseline only), all-reduce sweep, MLP.
set -x
N=${1:-8}
OUT=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > $OUT/r2f_multi_n$N.log 2>&1; tail -3 $OUT/r2f_multi_n$N.log
timeout 600 $TR --master-port 29731 bench.py --gpus $N --steps 20 --warmup 5 --baseline-variants eager > $OUT/r2f_bench_n$N.json 2> $OUT/r2f_bench_n$N.err; echo rc=$?; grep "bench +" $OUT/r2f_bench_n$N.err
timeout 300 $TR --master-port 29732 bench/allreduce_sweep.py --max-mb 256 --out $OUT/r2f_sweep_n$N.json > $OUT/r2f_sweep_n$N.log 2>&1; grep -c '"ok": true' $OUT/r2f_sweep_n$N.log; grep CHECKS $OUT/r2f_sweep_n$N.log
timeout 200 $TR --master-port 29734 bench/mlp.py --epochs 3 --steps-per-execution 8 2>/dev/null | grep bench > $OUT/r2f_mlp_n$N.json; cat $OUT/r2f_mlp_n$N.json
