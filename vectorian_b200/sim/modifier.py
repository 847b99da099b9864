"""Token-similarity modifiers: similarity matrices of several token similarities combined per query
(vectorian/sim/modifier.py:8-117, applied by ModifiedSimilarityMatrixFactory::create,
vectorian/core/cpp/metric/modifier.cpp:18-74).

The operands are [V x T] matrices, a few MB per query; the combination runs on the host with numpy exactly where
the reference runs it (the C++ core calls back into Python with the operand matrices), and the result is uploaded
for the device DP (vx_sim_upload).  Each operand matrix already carries the identity fix-up and the clip; the
combined matrix is used as it is, like in the reference.
"""

from typing import List

import numpy as np

from .kernel import Kernel, UnaryOperator
from .token import TokenSim


class TokenSimilarityModifier(TokenSim):
	@property
	def is_modifier(self):
		return True

	@property
	def operands(self):
		raise NotImplementedError()


class UnaryTokenSimilarityModifier(TokenSimilarityModifier):
	def __init__(self, operand: TokenSim, operators: List[UnaryOperator]):
		self._operand = operand
		self._kernel = Kernel(operators)

	def __call__(self, operands, out):
		data = operands[0]
		out["similarity"][:] = data["similarity"]
		self._kernel(out["similarity"])
		for k in data.keys():
			if k != "similarity":
				out[k][:] = data[k]

	@property
	def operands(self):
		return [self._operand]

	@property
	def name(self):
		return self._kernel.name(self._operand.name)

	@property
	def embeddings(self):
		return list(self._operand.embeddings)


class MixedTokenSimilarity(TokenSimilarityModifier):
	def __init__(self, metrics: List[TokenSim], weights):
		if len(metrics) != len(weights):
			raise ValueError("one weight per metric")
		self._metrics = list(metrics)
		self._weights = list(weights)

	@property
	def operands(self):
		return self._metrics

	def __call__(self, operands, out):
		for k in out.keys():
			avg = np.average([x[k] for x in operands], axis=0, weights=self._weights)
			assert avg.shape == out[k].shape
			out[k][:] = avg

	@property
	def name(self):
		total = np.sum(self._weights)
		return '(' + ' + '.join(f'{w / total} * {m.name}' for m, w in zip(self._metrics, self._weights)) + ')'

	@property
	def embeddings(self):
		return [e for m in self._metrics for e in m.embeddings]


class ExtremumTokenSimilarity(TokenSimilarityModifier):
	_name = None

	def __init__(self, metrics: List[TokenSim]):
		self._metrics = list(metrics)

	@property
	def operands(self):
		return self._metrics

	def __call__(self, operands, out):
		# the reference selects with argmax for BOTH subclasses (sim/modifier.py:87-93: `_operator` is never used)
		sim = np.array([x["similarity"] for x in operands])
		sel = np.argmax(sim, axis=0)
		out["similarity"][:] = np.take_along_axis(sim, sel[None], axis=0)[0]

	@property
	def name(self):
		return f'{self._name}({", ".join(x.name for x in self._metrics)})'

	@property
	def embeddings(self):
		return [e for m in self._metrics for e in m.embeddings]


class MaximumTokenSimilarity(ExtremumTokenSimilarity):
	_name = 'maximum'


class MinimumTokenSimilarity(ExtremumTokenSimilarity):
	_name = 'minimum'
