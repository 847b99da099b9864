#!/bin/bash
# ncu --set full of one kernel of a scripts/bench_configs.py config: gpu_ncu_cfg.sh <tag> <config> <kernel-regex> [skip]
TAG=$1; CFG=$2; K=$3; SKIP=${4:-3}
mkdir -p gpurun_out/$TAG
timeout 300 python scripts/bench_configs.py $CFG > gpurun_out/$TAG/plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o gpurun_out/$TAG/prof -f python scripts/bench_configs.py $CFG > gpurun_out/$TAG/ncu.log 2>&1
tail -2 gpurun_out/$TAG/ncu.log
