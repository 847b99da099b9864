#!/bin/bash
# Round 2, visit V: general-gap scratch kernel (queries of 17-32 tokens) with fewer persistent warps, so that the column
# histories of all warps fit in L2.
set -u
OUT=gpurun_out/${1:-r2_v}
mkdir -p $OUT
for b in 4 3 2 1; do
	VX_SCRATCH_BLOCKS_PER_SM=$b timeout 600 python scripts/bench_configs.py long 2> $OUT/err_$b | grep "WSB" > $OUT/long_$b.jsonl
	python - $OUT/long_$b.jsonl $b <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    d = json.loads(ln)
    print("blocks/SM", sys.argv[2], d["config"][:60], round(d["wall_ms"], 3), "ms")
PY
done
