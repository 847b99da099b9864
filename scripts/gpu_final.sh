#!/bin/bash
# Closing check on one fresh B200: smoke, every GPU test, the default bench line, the reference arm.
set -u
OUT=gpurun_out/${1:-r2_close}
mkdir -p $OUT
python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/smoke.log
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest_gpu.log
s=$(date +%s); python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$? in $(( $(date +%s) - s )) s"
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"
python - $OUT <<'PY'
import json, sys
d = json.load(open(sys.argv[1] + "/bench.json"))
print("value %.4g e2e %.4g frac %.4f parity %s clocks %s launches %s cpu %s" % (d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["sharded_parity"], d.get("clocks"), d["gpu_launches"], d["cpu_baseline"]["value"]))
print({k: (v.get("device_ms") or v.get("batch_wall_ms") or v.get("phases_ms")) for k, v in d["secondary"].items()})
r = json.load(open(sys.argv[1] + "/bench_ref.json"))
print("reference arm", r["value"], r["impl"])
PY
