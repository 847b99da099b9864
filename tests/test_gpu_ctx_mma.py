"""Contextual similarity GEMM: the tcgen05 3xTF32 kernel (vx_ctx_mma.cu) against the FP32 SIMT kernel
(ctx_sim_kernel) and a float64 numpy product, through the C-ABI test hook vx_debug_ctx_sim.

Tolerance (north_star: the similarity GEMM agrees within 1e-5): 2e-6 absolute on unit vectors.
"""

import ctypes as C

import numpy as np
import pytest

from vectorian_b200 import _cabi

pytestmark = pytest.mark.gpu


def ctx_sim(x, qhat, tp, kernel):
	lib = _cabi.lib()
	n, d = x.shape
	out = np.empty((n, tp), dtype=np.float32)
	ms = C.c_float(0)
	rc = lib.vx_debug_ctx_sim(x.ctypes.data, n, d, qhat.ctypes.data, qhat.shape[0], tp, kernel,
		out.ctypes.data, C.addressof(ms))
	if rc != 0:
		raise RuntimeError(lib.vx_last_error().decode())
	return out, ms.value


def unit(rng, n, d):
	x = rng.standard_normal((n, d)).astype(np.float32)
	x /= np.linalg.norm(x, axis=1, keepdims=True)
	return np.ascontiguousarray(x)


@pytest.mark.parametrize("n,d,t,tp", [
	(128, 64, 3, 4), (1000, 128, 8, 8), (4096 + 77, 128, 12, 12), (20000, 128, 13, 16),
	(3000, 256, 16, 16), (128 * 148 * 3 + 5, 64, 5, 8), (777, 192, 7, 8), (50, 96, 2, 4), (2500, 768, 9, 12), (129, 32, 1, 4),
	(5000, 300, 8, 8), (600, 100, 4, 4), (300, 20, 3, 4)])
def test_mma_matches_simt_and_float64(n, d, t, tp):
	rng = np.random.default_rng(n + d + t)
	x, q = unit(rng, n, d), unit(rng, t, d)
	x[::7] = q[rng.integers(0, t, size=len(x[::7]))]    # exact matches: S = 1 up to rounding, clipped at 1
	x[1::11] *= -1.0                                      # negative products: clipped at 0
	ref = np.zeros((n, tp), dtype=np.float64)
	ref[:, :t] = np.clip(x.astype(np.float64) @ q.astype(np.float64).T, 0.0, 1.0)
	simt, _ = ctx_sim(x, q, tp, 0)
	mma, _ = ctx_sim(x, q, tp, 1)
	assert np.all(mma[:, t:] == 0.0)
	assert np.abs(simt - ref).max() < 2e-6
	assert np.abs(mma - ref).max() < 2e-6
	assert np.abs(mma - simt).max() < 2e-6


def test_mma_rejects_other_shapes():
	rng = np.random.default_rng(0)
	x, q = unit(rng, 256, 102), unit(rng, 4, 102)    # rows not 16-byte aligned: no TMA
	with pytest.raises(RuntimeError):
		ctx_sim(x, q, 4, 1)
	simt, _ = ctx_sim(x, q, 4, 0)
	assert np.abs(simt - np.clip(x @ q.T, 0, 1)).max() < 2e-6
