"""Gap cost classes with the signatures of ``pyalign.gaps``.

Mirrors vectorian/core/lib/pyalign/pyalign/gaps.py:89-286 of the reference
(same class names, constructor arguments, ``costs(n)`` tables in fp32,
``to_special_case()`` used for solver selection, ``to_tuple()``).
Plotting helpers are UI and out of scope.
"""

import numpy as np


class GapCost:
	"""A gap cost w_k that depends only on the gap length k."""

	def to_tuple(self):
		raise NotImplementedError()

	def to_special_case(self):
		return {}

	def costs(self, n):
		raise NotImplementedError()

	@property
	def title(self):
		raise NotImplementedError()


class ConstantGapCost(GapCost):
	"""w_k = u, w_0 = 0 (gaps.py:89-119)."""

	def __init__(self, u):
		self._u = u
		assert u >= 0

	def to_tuple(self):
		return 'constant', self._u

	def to_special_case(self):
		if self._u == 0:
			return {'linear': 0}
		return {}

	def costs(self, n):
		w = np.full((n,), self._u, dtype=np.float32)
		w[0] = 0
		return w

	@property
	def title(self):
		return f"w(k) = {self._u:.1f}, w(0) = 0"


class LinearGapCost(GapCost):
	"""w_k = u k (gaps.py:122-155)."""

	def __init__(self, u):
		self._u = u
		assert u >= 0

	def to_tuple(self):
		return 'linear', self._u

	def to_special_case(self):
		return {'linear': self._u}

	def costs(self, n):
		return np.linspace(0., (n - 1) * self._u, n, dtype=np.float32)

	@property
	def title(self):
		return f"w(k) = {self._u:.1f} * k"


class AffineGapCost(GapCost):
	"""w_k = u + v k, w_0 = 0 (gaps.py:158-197)."""

	def __init__(self, open=None, extend=0, u=None, v=None):
		if open is not None:
			assert u is None and v is None
			self._u = open - extend
			self._v = extend
		else:
			assert extend == 0
			self._u = u
			self._v = v
		assert self._u >= 0
		assert self._v >= 0

	def to_tuple(self):
		return 'affine', self._u, self._v

	def to_special_case(self):
		return {'affine': (self._u, self._v)}

	def costs(self, n):
		w = self._u + np.linspace(0., (n - 1) * self._v, n, dtype=np.float32)
		w[0] = 0
		return w

	@property
	def title(self):
		return f"w(k) = {self._u:.1f} + {self._v:.1f} * k, w(0) = 0"


class LogarithmicGapCost(GapCost):
	"""w_k = u + v ln(k), w_0 = 0 (gaps.py:200-231)."""

	def __init__(self, u, v):
		self._u = u
		self._v = v
		assert u >= 0
		assert v >= 0

	def to_tuple(self):
		return 'logarithmic', self._u, self._v

	def costs(self, n):
		assert n > 0
		w = np.full((n,), self._u, dtype=np.float32)
		w[1:] += self._v * np.log(np.arange(1, n), dtype=np.float32)
		w[0] = 0
		return w

	@property
	def title(self):
		return f"w(k) = {self._u:.1f} + {self._v:.1f} * ln(k), w(0) = 0"


class ExponentialGapCost(GapCost):
	"""w_k = 1 - u^(-k v) (gaps.py:234-258)."""

	def __init__(self, u, v):
		self._u = u
		self._v = v
		assert u >= 0
		assert v >= 0

	def to_tuple(self):
		return 'exponential', self._u, self._v

	def costs(self, n):
		c = np.empty((n,), dtype=np.float32)
		for i in range(n):
			c[i] = 1 - (self._u ** -(i * self._v))
		return c

	@property
	def title(self):
		return f"w(k) = 1 - {self._u:.1f}^(-k * {self._v:.1f})"


class UserFuncGapCost(GapCost):
	"""A custom gap cost defined by a Python function (gaps.py:261-279)."""

	def __init__(self, costs_fn):
		self._costs_fn = costs_fn

	def to_tuple(self):
		return 'user', self._costs_fn

	def costs(self, n):
		c = np.empty((n,), dtype=np.float32)
		c[0] = 0
		for i in range(1, n):
			c[i] = self._costs_fn(i)
		return c

	@property
	def title(self):
		return f"w(k) = {self._costs_fn}, w(0) = 0"


def smooth_gap_cost(k):
	"""1 - 2^(-x/k): w_k = 0.5 at gap length k (gaps.py:282-292)."""
	if k > 0:
		return ExponentialGapCost(2, 1 / k)
	return ConstantGapCost(0)


def resolve_gap_costs(gap, max_len_s, max_len_t):
	"""Solver selection of the reference, as a plain dict for vx_align_params.

	Follows GapCostOptions (pyalign/algorithm/factory.h:33-65), to_gap_cost_options
	(factory.h:67-91) and FactoryCreator::resolve_gap_type (factory.h:930-965):
	both sides linear -> linear solver, both affine -> Gotoh, otherwise general.
	Where the reference dereferences an empty optional for a mixed special/general
	pair (factory.h:956-963, undefined behaviour) the special-cased side is
	materialised through ``costs(n)``.
	"""
	if isinstance(gap, dict):
		if not all(k in ("s", "t") for k in gap.keys()):
			raise ValueError(gap)
		gap_s, gap_t = gap.get("s"), gap.get("t")
	else:
		gap_s = gap_t = gap

	def special(g):
		if g is None:
			return {'linear': 0.0}
		return g.to_special_case()

	sp_s, sp_t = special(gap_s), special(gap_t)

	if 'linear' in sp_s and 'affine' not in sp_s and 'linear' in sp_t and 'affine' not in sp_t:
		return dict(gap_kind=0, lin=(float(sp_s['linear']), float(sp_t['linear'])))
	if 'affine' in sp_s and 'affine' in sp_t:
		# AffineGapCost(u, v): w(k) = u + v k  ->  (open0, ext) = (u, v)   factory.h:41-52
		return dict(gap_kind=1, affine_s=tuple(map(float, sp_s['affine'])),
			affine_t=tuple(map(float, sp_t['affine'])))

	def table(g, n):
		if g is None:
			return np.zeros((n,), dtype=np.float32)   # zero_gap_tensor  factory.h:11-17
		w = np.ascontiguousarray(g.costs(n), dtype=np.float32)
		if w.shape != (n,):
			raise ValueError(f"expected gap cost tensor length of {n}, got {w.shape[0]}")
		return w

	return dict(gap_kind=2, w_s=table(gap_s, max_len_s + 1), w_t=table(gap_t, max_len_t + 1))
