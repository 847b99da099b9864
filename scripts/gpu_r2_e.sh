#!/bin/bash
# Round 2, visit E: WSB kernel variants A/B (pipelined history loops vs not), at 50 M and at the 6.25 M shard size.
set -u
OUT=gpurun_out/${1:-r2_e}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest.log
for v in pipe nopipe; do
	for n in 50000000 6250000; do
		E=""; [ $v = nopipe ] && E="VX_WSB_NO_PIPE=1"
		env $E timeout 600 python bench.py --sentences $n --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > $OUT/bench_${v}_$n.json 2> $OUT/bench_${v}_$n.err; echo "bench $v $n rc=$?"
	done
done
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        print("%-40s value %.4g e2e %.4g dp_ms %.3f frac %.4f parity %s phases %s" % (f.split("/")[-1], d["value"], d["e2e"]["value"], d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d["sharded_parity"], {k: round(v, 3) for k, v in d["phases_ms"].items()}))
    except Exception as e:
        print(f, "failed", e)
PY
