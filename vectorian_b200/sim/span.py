"""Span similarity = token similarity + an optimizer (alignment) (vectorian/sim/span.py:9-71)."""

from ..alignment import Optimizer, LocalAlignment, ConstantGapCost
from .token import TokenSim


class SpanSim:
	def create_index(self, partition):
		raise NotImplementedError()

	def to_args(self, index):
		raise NotImplementedError()


class OptimizedSpanSim(SpanSim):
	def __init__(self, token_sim: TokenSim, optimizer: Optimizer = None, tag_weights: dict = None, **kwargs):
		if not isinstance(token_sim, TokenSim):
			raise TypeError(token_sim)
		if optimizer is None:
			optimizer = LocalAlignment(gap={'s': ConstantGapCost(0), 't': ConstantGapCost(0)})
		if not isinstance(optimizer, Optimizer):
			raise TypeError(optimizer)
		self._token_sim = token_sim
		self._optimizer = optimizer
		self._tag_weights = tag_weights
		self._options = kwargs

	@property
	def token_sim(self):
		return self._token_sim

	@property
	def optimizer(self):
		return self._optimizer

	def create_index(self, partition, **kwargs):
		from ..index import BruteForceIndex
		return BruteForceIndex(partition, self, **kwargs)

	def to_args(self, index):
		if not self._tag_weights:
			if self._options:
				raise ValueError(f"illegal option(s): {', '.join(self._options.keys())}")
			return {
				'metric': 'alignment-isolated',
				'token_metric': self._token_sim,
				'alignment': self._optimizer.to_args(index.partition)
			}
		else:
			return {
				'metric': 'alignment-tag-weighted',
				'token_metric': self._token_sim,
				'alignment': self._optimizer.to_args(index.partition),
				'pos_mismatch_penalty': self._options.get('pos_mismatch_penalty', 0),
				'similarity_threshold': self._options.get('similarity_threshold', 0),
				'tag_weights': self._tag_weights
			}
