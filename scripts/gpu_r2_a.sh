#!/bin/bash
# Round 2, visit A: FADD2 pipe rates, bit-exact parity of the packed-add WSB kernel, bench A/B (FADD2 vs scalar adds).
set -u
OUT=gpurun_out/r2_a
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 120 scripts/ubench/_bin/pipes > $OUT/pipes.txt 2>&1; echo "pipes rc=$?"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_parity.log 2>&1; echo "pytest parity rc=$?"; tail -3 $OUT/pytest_parity.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_f2.json 2> $OUT/bench_f2.err; echo "bench f2 rc=$?"
VX_WSB_NO_F2=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_scalar.json 2> $OUT/bench_scalar.err; echo "bench scalar rc=$?"
python - <<'PY'
import json
for n in ("f2", "scalar"):
    try:
        d = json.load(open("gpurun_out/r2_a/bench_%s.json" % n))
        print(n, "value %.4g e2e %.4g dp_ms %.3f frac %.4f clocks %s" % (d["value"], d["e2e"]["value"], d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d.get("clocks")))
    except Exception as e:
        print(n, "failed", e)
PY
timeout 1500 python -m pytest tests/test_gpu_full_size.py -m gpu -x -q > $OUT/pytest_full.log 2>&1; echo "pytest full rc=$?"; tail -3 $OUT/pytest_full.log
