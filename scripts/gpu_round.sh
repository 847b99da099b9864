#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, launch list and a full ncu capture of the DP kernel.
# Usage (from the repo root, under gpurun): bash scripts/gpu_round.sh <tag> [kernel-regex] [bench sentences]
set -u
TAG=${1:-r1}
KREGEX=${2:-score_general}
NSENT=${3:-50000000}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
python bench.py --sentences $NSENT > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"; cat $OUT/bench_ref.json
PCMD="python bench.py --sentences 2000000 --steps 2 --warmup 3 --no-cpu-baseline"
$PCMD > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $PCMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$PCMD > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 5 -c 5 -o $OUT/prof $PCMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full.log
