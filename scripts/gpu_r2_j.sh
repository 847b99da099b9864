#!/bin/bash
# Round 2, visit J: C5 shard with the multi-setting kernel as one 4 + 4 pass vs two passes of four settings (VX_MULTI_SPLIT=1).
set -u
OUT=gpurun_out/${1:-r2_j}
mkdir -p $OUT
python scripts/bench_configs.py c5s > $OUT/c5s.jsonl 2> $OUT/err
VX_MULTI_SPLIT=1 python scripts/bench_configs.py c5s > $OUT/c5s_split.jsonl 2>> $OUT/err
python - $OUT <<'PY'
import json, sys
for f in ("c5s", "c5s_split"):
    d = json.loads(open(sys.argv[1] + "/" + f + ".jsonl").readline())
    print(f, round(d["batch_wall_ms"], 2), round(d["batch_device_ms"], 2))
PY
