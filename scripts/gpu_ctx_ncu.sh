#!/bin/bash
# ncu --set full of the contextual similarity kernels through the vx_debug_ctx_sim hook
mkdir -p gpurun_out/ctxncu
cat > gpurun_out/ctxncu/ctx_run.py <<'PY'
import numpy as np
from tests.test_gpu_ctx_mma import ctx_sim, unit
rng = np.random.default_rng(1)
n, d, t, tp = 2_000_000, 128, 16, 16
x, q = unit(rng, n, d), unit(rng, t, d)
for k in (0, 1):
	out, ms = ctx_sim(x, q, tp, k)
	print(k, ms)
PY
timeout 120 env PYTHONPATH=. python gpurun_out/ctxncu/ctx_run.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ctx_ -o gpurun_out/ctxncu/prof -f env PYTHONPATH=. python gpurun_out/ctxncu/ctx_run.py > gpurun_out/ctxncu/ncu.log 2>&1
tail -3 gpurun_out/ctxncu/ncu.log
