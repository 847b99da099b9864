#!/bin/bash
# Round 2, visit R: queries of 33 .. 64 tokens on the linear / affine kernels: parity + throughput.
set -u
OUT=gpurun_out/${1:-r2_r}
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 900 python scripts/bench_configs.py long > $OUT/long.jsonl 2> $OUT/long.err; echo "long rc=$?"
python - $OUT <<'PY'
import json, sys
for ln in open(sys.argv[1] + "/long.jsonl"):
    d = json.loads(ln)
    print(d["config"][:70], round(d["wall_ms"], 3), "ms", round(d["sentences_per_s_e2e"] / 1e9, 3), "G/s", d["launches"])
PY
