#!/bin/bash
# Round 2, visit M: the 8-warp WSB launch split at a length class (shorter classes get a smaller shared-memory slice -> more L1
# for the S-row gather): VX_WSB_SPLIT=<rows>.
set -u
OUT=gpurun_out/${1:-r2_m}
mkdir -p $OUT
for split in 0 20 24 28; do
	for n in 50000000 6250000; do
		VX_WSB_SPLIT=$split timeout 600 python bench.py --sentences $n --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_s${split}_$n.json 2> $OUT/bench_s${split}_$n.err
	done
done
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        print("%-36s dev %.3f ms frac %.3f parity %s launches/step %.1f dp %.3f" % (f.split("/")[-1], d["ms_per_step"], d["roofline"]["frac"], d["sharded_parity"], d["gpu_launches"] / d["steps"], d["phases_ms"]["dp_ms"]))
    except Exception as e:
        print(f, "failed", e)
PY
