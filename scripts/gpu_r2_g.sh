#!/bin/bash
# Round 2, visit G: linear / affine / multi-query kernels with the quad-unrolled row stream (for_each_row): parity suite, then
# the configurations those kernels serve.  `VX_...` switches are read once per process.
set -u
OUT=gpurun_out/${1:-r2_g}
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 900 python scripts/bench_configs.py mq > $OUT/mq.jsonl 2> $OUT/mq.err; echo "mq rc=$?"
timeout 900 python scripts/bench_configs.py c2 c2big gotoh > $OUT/lin.jsonl 2> $OUT/lin.err; echo "lin rc=$?"
timeout 900 python scripts/bench_configs.py c5 > $OUT/c5.jsonl 2> $OUT/c5.err; echo "c5 rc=$?"
timeout 900 python scripts/bench_configs.py tag long > $OUT/tag_long.jsonl 2> $OUT/tag_long.err; echo "tag long rc=$?"
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/*.jsonl")):
    for ln in open(f):
        try:
            d = json.loads(ln)
        except Exception:
            continue
        keys = [k for k in ("dp_ms", "device_ms", "wall_ms", "single_calls_wall_ms", "batch_wall_ms", "speedup", "sentences_per_s", "identical_results", "sweep_ms", "device_sweep_ms") if k in d]
        print(d.get("config", "?")[:70], {k: (round(d[k], 3) if isinstance(d[k], float) else d[k]) for k in keys})
PY
