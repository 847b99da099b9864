#!/bin/bash
# Round 2, visit N: what the WSB main launch is sensitive to: L1 size (VX_WSB_PAD_SMEM), PDL (VX_NO_PDL), split launches.
set -u
OUT=gpurun_out/${1:-r2_n}
mkdir -p $OUT
run() {
	name=$1; shift
	env "$@" timeout 600 python bench.py --sentences 50000000 --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_$name.json 2> $OUT/bench_$name.err
}
run base VX_X=0
run pad12k VX_WSB_PAD_SMEM=12288
run pad25k VX_WSB_PAD_SMEM=25600
run nopdl VX_NO_PDL=1
run split28_nopdl VX_NO_PDL=1 VX_WSB_SPLIT=28
run split32 VX_WSB_SPLIT=32
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        print("%-36s dev %.3f ms frac %.3f launches/step %.1f dp %.3f" % (f.split("/")[-1], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"] / d["steps"], d["phases_ms"]["dp_ms"]))
    except Exception as e:
        print(f, "failed", e)
PY
