#!/bin/bash
# Round 2, visit O: long length classes at one warp less but inside the 196 KB shared-memory carve-out (VX_WSB_SMEM_MAX=196608).
set -u
OUT=gpurun_out/${1:-r2_o2}
mkdir -p $OUT
for mx in 0 196608; do
	for n in 50000000 6250000; do
		if [ $mx = 0 ]; then
			timeout 600 python bench.py --sentences $n --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_m${mx}_$n.json 2> $OUT/bench_m${mx}_$n.err
		else
			VX_WSB_SMEM_MAX=$mx timeout 600 python bench.py --sentences $n --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_m${mx}_$n.json 2> $OUT/bench_m${mx}_$n.err
		fi
	done
done
timeout 600 python scripts/bench_configs.py win > $OUT/win_m0.jsonl 2> $OUT/win_m0.err
VX_WSB_SMEM_MAX=196608 timeout 600 python scripts/bench_configs.py win > $OUT/win_m196608.jsonl 2> $OUT/win_m196608.err
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        print("%-36s dev %.3f ms frac %.3f launches/step %.1f dp %.3f parity %s" % (f.split("/")[-1], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"] / d["steps"], d["phases_ms"]["dp_ms"], d["sharded_parity"]))
    except Exception as e:
        print(f, "failed", e)
for f in sorted(glob.glob(sys.argv[1] + "/win_*.jsonl")):
    for ln in open(f):
        d = json.loads(ln)
        print(f.split("/")[-1], {k: (round(v, 3) if isinstance(v, float) else v) for k, v in d.items() if "ms" in k})
PY
