#!/bin/bash
# One 8-GPU session: multi-GPU tests, headline bench, all-reduce sweep (+ block-cap variants), tuner packing.
set -x
N=${1:-8}
OUT=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > $OUT/r2_multi_n$N.log 2>&1; tail -3 $OUT/r2_multi_n$N.log
timeout 900 $TR --master-port 29731 bench.py --gpus $N --steps 20 --warmup 5 > $OUT/r2_bench_n$N.json 2> $OUT/r2_bench_n$N.err; grep "bench +" $OUT/r2_bench_n$N.err
timeout 600 $TR --master-port 29732 bench/allreduce_sweep.py --max-mb 1024 --out $OUT/r2_sweep_full.json > $OUT/r2_sweep_n$N.log 2>&1; grep -c '"ok": true' $OUT/r2_sweep_n$N.log; grep CHECKS $OUT/r2_sweep_n$N.log
for cap in 64 128; do
  TFC_COMM_MAX_BLOCKS=$cap timeout 300 $TR --master-port 29733 bench/allreduce_sweep.py --max-mb 256 --out $OUT/r2_sweep_cap$cap.json > $OUT/r2_sweep_cap${cap}_n$N.log 2>&1
done
for spe in 1 8; do timeout 300 $TR --master-port 29734 bench/mlp.py --epochs 3 --steps-per-execution $spe 2>/dev/null | grep bench >> $OUT/r2_mlp_n$N.json; done
cat $OUT/r2_mlp_n$N.json
timeout 600 python bench/tuner8.py --trials 8 --model mlp --epochs 40 > $OUT/r2_tuner8_mlp.json 2> $OUT/r2_tuner8_mlp.err; cat $OUT/r2_tuner8_mlp.json
timeout 600 python bench/tuner8.py --trials 8 --model resnet18 --epochs 5 > $OUT/r2_tuner8_resnet18.json 2> $OUT/r2_tuner8_resnet18.err; cat $OUT/r2_tuner8_resnet18.json
timeout 300 python bench/tuner8.py --trials 8 --model mlp --epochs 40 --cold > $OUT/r2_tuner8_mlp_cold.json 2>/dev/null; cat $OUT/r2_tuner8_mlp_cold.json
