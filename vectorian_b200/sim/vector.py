"""Vector similarity strategies with the reference's names and call signature.

Mirrors vectorian/sim/vector.py:17-176.  ``CosineSim`` is the one the device computes itself
(vx_sim_static: normalise + contract + clip on the GPU); every other ``VectorSim`` is evaluated on
the host through its ``compute(a, b, out)`` and the resulting matrix is uploaded (vx_sim_upload),
which is what the reference does too: it calls back into Python once per query
(vectorian/core/cpp/metric/static.cpp:47-50).
"""

import numpy as np


class VectorSim:
	"""A strategy to compute a scalar similarity from pairs of vectors (sim/vector.py:17-46)."""

	def __call__(self, a, b, out):
		self.compute(a, b, out)

	def compute(self, a, b, out):
		raise NotImplementedError()

	@property
	def name(self):
		raise NotImplementedError()

	@property
	def on_device(self):
		"""True if libvectorian_b200 evaluates this similarity itself."""
		return False


class CosineSim(VectorSim):
	"""Cosine of the angle between two vectors (sim/vector.py:63-82); negative values are clipped later."""

	def compute(self, a, b, out):
		np.matmul(a.normalized, b.normalized.T, out=out)

	@property
	def name(self):
		return "cosine"

	@property
	def on_device(self):
		return True


class FuzzyJaccardSim(VectorSim):
	"""sim/vector.py:85-95"""

	def compute(self, a, b, out):
		for i, a_v in enumerate(a.unmodified):
			for j, b_v in enumerate(b.unmodified):
				out[i, j] = np.sum(np.minimum(a_v, b_v)) / np.sum(np.maximum(a_v, b_v))

	@property
	def name(self):
		return "fuzzy-jaccard"


class PNormSim(VectorSim):
	"""sim/vector.py:141-163: max(0, 1 - |a - b|_p * scale)"""

	def __init__(self, p=2, scale=1):
		self._p = p
		self._scale = scale

	def compute(self, a, b, out):
		d = a.unmodified[:, np.newaxis] - b.unmodified[np.newaxis, :]
		d = np.sum(np.power(np.abs(d), self._p), axis=-1)
		d = np.power(d, 1 / self._p)
		out[:, :] = np.maximum(0, 1 - d * self._scale)

	@property
	def name(self):
		return f"p-norm({self._p})"
