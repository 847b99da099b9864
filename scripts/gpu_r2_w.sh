#!/bin/bash
# Round 2, visit W: general-gap scratch kernel with U rows per pass over the history: parity, long queries, long windows.
set -u
OUT=gpurun_out/${1:-r2_w}
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest_gpu.log
timeout 600 python scripts/bench_configs.py long win > $OUT/long_win.jsonl 2> $OUT/err; echo "rc=$?"
python - $OUT/long_win.jsonl <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    d = json.loads(ln)
    print(d["config"][:75], round(d["wall_ms"], 3), "ms")
PY
