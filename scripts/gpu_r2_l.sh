#!/bin/bash
# Round 2, visit L: WSB main launch with a long queue (two long-capable warps per CTA) vs separate launches per class.
set -u
OUT=gpurun_out/${1:-r2_l}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -2 $OUT/pytest_gpu.log
for n in 50000000 6250000 1000000; do
	timeout 600 python bench.py --sentences $n --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_$n.json 2> $OUT/bench_$n.err; echo "bench $n rc=$?"
	VX_WSB_NO_HYBRID=1 timeout 600 python bench.py --sentences $n --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-clocks > $OUT/bench_nohybrid_$n.json 2> $OUT/bench_nohybrid_$n.err
done
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        print("%-36s dev %.3f ms e2e %.3f ms frac %.3f parity %s launches/step %.1f phases %s" % (f.split("/")[-1], d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["sharded_parity"], d["gpu_launches"] / d["steps"], {k: round(v, 3) for k, v in d["phases_ms"].items() if "gather" not in k}))
    except Exception as e:
        print(f, "failed", e)
PY
