#!/bin/bash
# Round 2, visit H: multi-query kernels with chunk-interleaved S (every load instruction of a window's lanes covers one
# contiguous piece): parity, the multi-query configurations, one ncu capture of the packed kernel.
set -u
OUT=gpurun_out/${1:-r2_h}
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 900 python scripts/bench_configs.py mq > $OUT/mq.jsonl 2> $OUT/mq.err; echo "mq rc=$?"
VX_MQ_NO_PACK=1 timeout 900 python scripts/bench_configs.py mq > $OUT/mq_unpacked.jsonl 2> $OUT/mq_unpacked.err; echo "mq unpacked rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:score_linear_mq -c 2 -o $OUT/mq2_full python scripts/bench_configs.py mq > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/*.jsonl")):
    for ln in open(f):
        try:
            d = json.loads(ln)
        except Exception:
            continue
        keys = [k for k in ("device_ms", "wall_ms", "single_calls_wall_ms", "batch_wall_ms", "speedup", "identical_results") if k in d]
        print(f.split("/")[-1], d.get("config", "?")[:60], {k: (round(d[k], 3) if isinstance(d[k], float) else d[k]) for k in keys})
PY
