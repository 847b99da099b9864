#!/bin/bash
# Round 2, visit P: ncu capture (source-level) of the traceback kernel and the pruned top-k kernels in the C3 bench at 6.25 M sentences.
set -u
OUT=gpurun_out/${1:-r2_p}
mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"trace_warp_kernel|topk_" -c 8 -o $OUT/fixed_full python bench.py --sentences 6250000 --steps 2 --warmup 1 --no-cpu-baseline --no-secondary --no-clocks > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
