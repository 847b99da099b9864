#!/bin/bash
# Multi-GPU visit: gpu_multi.sh <n_gpus> [configs...]: headline bench through torchrun (as the driver launches it) and the
# sharded BASELINE configs (C3 strong-scaling view, C4 full config, C5 sweep) through scripts/bench_configs.py.
N=$1; shift
CFGS=${@:-c3 c4full c5}
OUT=gpurun_out/r1_multi_$N
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $OUT/smi.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR scripts/bench_configs.py $CFGS > $OUT/configs.jsonl 2> $OUT/configs.err; echo "configs rc=$?"; cat $OUT/configs.jsonl; tail -5 $OUT/configs.err
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
