#!/bin/bash
mkdir -p gpurun_out/c5
cat > gpurun_out/c5/run.py <<'PY'
import sys
sys.argv = ["x"]
sys.path.insert(0, "scripts")
import bench_configs as b
b.run_sweep("c5 small", 10_000_000, n_queries=2)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c5/launches.csv env PYTHONPATH=. python gpurun_out/c5/run.py > gpurun_out/c5/ncu.log 2>&1
tail -2 gpurun_out/c5/ncu.log
