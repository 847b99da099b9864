#!/bin/bash
# Round 2, visit I: ncu capture of the multi-setting sweep kernel (C5) + parity after the row-stream clean-up.
set -u
OUT=gpurun_out/${1:-r2_i}
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -2 $OUT/pytest_gpu.log
timeout 900 python scripts/bench_configs.py c5s > $OUT/c5s.jsonl 2> $OUT/c5s.err; echo "c5s rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:score_linear_multi -c 2 -o $OUT/multi_full python scripts/bench_configs.py c5s > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
cat $OUT/c5s.jsonl | cut -c1-400
