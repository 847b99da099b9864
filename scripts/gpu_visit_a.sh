#!/bin/bash
# Validation visit: GPU parity tests, smoke, headline bench, reference arm, secondary configs, ncu of the fused contextual kernel.
set -u
OUT=gpurun_out/r1_visit_a
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"; cat $OUT/bench_ref.json
timeout 600 python scripts/bench_configs.py c1 c2 c4 gotoh > $OUT/configs.log 2>&1; echo "configs rc=$?"; cat $OUT/configs.log | tail -12
timeout 300 python scripts/bench_configs.py c4 > $OUT/c4_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ctx_ -s 3 -c 2 -o $OUT/prof_ctx -f python scripts/bench_configs.py c4 > $OUT/ncu_ctx.log 2>&1
echo "ncu ctx rc=$?"; tail -2 $OUT/ncu_ctx.log
