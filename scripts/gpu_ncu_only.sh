#!/bin/bash
# Full ncu capture of the DP kernels of one search (all length classes). Usage: gpu_ncu_only.sh <tag> <regex> <count> [skip]
TAG=${1:-x}; KREGEX=${2:-wsb_score}; CNT=${3:-8}; SKIP=${4:-24}
OUT=gpurun_out/$TAG; mkdir -p $OUT
PCMD="python bench.py --sentences 2000000 --steps 2 --warmup 3 --no-cpu-baseline"
$PCMD > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s $SKIP -c $CNT -o $OUT/prof $PCMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full.log | cut -c1-300
