#!/bin/bash
# Round 2, visit C (1 GPU): shard tests, whole GPU suite, bench N=1.
set -u
OUT=gpurun_out/r2_c
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -x -q > $OUT/pytest_shard.log 2>&1; echo "shard rc=$?"; tail -15 $OUT/pytest_shard.log
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -5 $OUT/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -3 $OUT/bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_c/bench.json"))
print("value %.4g e2e %.4g frac %.4f parity %s phases %s scaling %s" % (d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["sharded_parity"], d["phases_ms"], d["scaling"]))
PY
