#!/bin/bash
# Round 2, visit K: whole GPU suite + C5 (10 M sentences, one GPU) + C5 shard after the row-stream rewrite of the linear kernels.
set -u
OUT=gpurun_out/${1:-r2_k}
mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 $OUT/pytest_gpu.log
timeout 900 python scripts/bench_configs.py c5 c5s > $OUT/c5.jsonl 2> $OUT/c5.err; echo "c5 rc=$?"
python - $OUT <<'PY'
import json, sys
for ln in open(sys.argv[1] + "/c5.jsonl"):
    d = json.loads(ln)
    print(d["config"][:60], round(d["batch_wall_ms"], 2), round(d["batch_device_ms"], 2), d.get("launches"))
PY
