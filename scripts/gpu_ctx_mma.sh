#!/bin/bash
# quick GPU check of the tensor-core contextual similarity kernel, every step under its own timeout
mkdir -p gpurun_out/ctxmma
timeout 120 python -m pytest tests/test_gpu_ctx_mma.py -x -q 2>&1 | tail -15 | tee gpurun_out/ctxmma/pytest.log
timeout 120 python - <<'PY' 2>&1 | tee gpurun_out/ctxmma/perf.log
import numpy as np
from tests.test_gpu_ctx_mma import ctx_sim, unit
rng = np.random.default_rng(1)
for n, d, t, tp in [(5_000_000, 128, 8, 8), (5_000_000, 128, 16, 16), (2_000_000, 256, 12, 12)]:
	x, q = unit(rng, n, d), unit(rng, t, d)
	for k in (0, 1):
		out, ms = ctx_sim(x, q, tp, k)
		gb = (n * d * 4 + n * tp * 4) / 1e9
		sub = slice(0, 200000)
		ref = np.clip(x[sub].astype(np.float64) @ q.astype(np.float64).T, 0, 1)
		err = np.abs(out[sub, :t] - ref).max()
		print("n=%d d=%d tp=%d kernel=%d  %.3f ms  %.0f GB/s  max err %.2e" % (n, d, tp, k, ms, gb / ms * 1e3, err))
PY
