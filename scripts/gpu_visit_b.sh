#!/bin/bash
# Visit B: all GPU tests incl. the full-size property tests, headline bench, C2 at 1M / 20M, ncu of the linear-gap DP kernel.
set -u
OUT=gpurun_out/r1_visit_b
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
free -g > $OUT/mem.txt; nproc >> $OUT/mem.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/pytest_gpu.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
timeout 600 python scripts/bench_configs.py c2 c2big > $OUT/configs.log 2>&1; echo "configs rc=$?"; tail -4 $OUT/configs.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:score_linear_kernel -s 3 -c 1 -o $OUT/prof_linear -f python scripts/bench_configs.py c2big > $OUT/ncu_linear.log 2>&1
echo "ncu linear rc=$?"; tail -2 $OUT/ncu_linear.log
