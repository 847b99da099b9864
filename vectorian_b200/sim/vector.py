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


class ImprovedSqrtCosineSim(VectorSim):
	"""sim/vector.py:98-131 (Sohangir & Wang 2017), vectors made non-negative first."""

	@staticmethod
	def _to_non_negative(x):
		t = np.repeat(x, 2, axis=-1)
		t[:, 1::2] = -t[:, 1::2]
		return np.maximum(0, t)

	def compute(self, a, b, out):
		a_pos = self._to_non_negative(a.unmodified)
		b_pos = self._to_non_negative(b.unmodified)
		num = np.sum(np.sqrt(a_pos[:, np.newaxis] * b_pos[np.newaxis, :]), axis=-1)
		x = np.sqrt(np.sum(a_pos, axis=-1))
		y = np.sqrt(np.sum(b_pos, axis=-1))
		denom = x[:, np.newaxis] * y[np.newaxis, :]
		with np.errstate(divide='ignore', invalid='ignore'):
			out[:, :] = num / denom
		np.nan_to_num(out, copy=False, nan=0)

	@property
	def name(self):
		return "improved-sqrt-cosine"


class PNormDistance(VectorSim):
	"""A p-norm DISTANCE (sim/vector.py:134-160); combine with kernel.DistanceToSimilarity."""

	def __init__(self, p=2):
		self._p = p

	def compute(self, a, b, out):
		d = a.unmodified[:, np.newaxis] - b.unmodified[np.newaxis, :]
		d = np.sum(np.power(np.abs(d), self._p), axis=-1)
		d = np.power(d, 1 / self._p)
		out[:, :] = d

	@property
	def name(self):
		return f"p-norm({self._p})"


class EuclideanDistance(PNormDistance):
	def __init__(self):
		super().__init__(p=2)


class ModifiedVectorSim(VectorSim):
	"""A VectorSim whose output is modified by unary operators (sim/vector.py:176-200)."""

	def __init__(self, source, *operators):
		from .kernel import Kernel
		self._source = source
		self._kernel = Kernel(operators)

	def compute(self, a, b, out):
		self._source(a, b, out)
		self._kernel(out)

	@property
	def name(self):
		return self._kernel.name(self._source.name)
