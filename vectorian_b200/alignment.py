"""Alignment optimizers with the signatures of vectorian/alignment.py:9-187.

``to_args`` produces the same option dict the reference hands to its C++ core
(``{'algorithm': 'pyalign', 'options': {'locality', 'gap_cost'}}``, alignment.py:31-38); the host
index turns it into ``vx_align_params`` (solver choice: pyalign/algorithm/factory.h:930-965).
Optimal-transport optimizers (WMD / WRD, alignment.py:190-313) are out of scope.
"""

from typing import Dict, Union

from ._cabi import VX_LOCAL, VX_GLOBAL, VX_SEMIGLOBAL
from .gaps import *  # noqa: F401,F403  (the reference does `from pyalign.gaps import *`)
from .gaps import GapCost, ConstantGapCost


class Locality:
	"""core.pyalign.Locality (pyalign/algorithm/enum.h:27-31)"""
	LOCAL = VX_LOCAL
	GLOBAL = VX_GLOBAL
	SEMIGLOBAL = VX_SEMIGLOBAL


class Optimizer:
	def to_description(self, partition):
		raise NotImplementedError()

	def to_args(self, partition):
		raise NotImplementedError()


def coalesce_default_gap(gap):
	return ConstantGapCost(0) if gap is None else gap


class Alignment(Optimizer):
	_locality = None
	_name = None

	def __init__(self, gap: Union[GapCost, Dict[str, GapCost]] = None):
		gap = coalesce_default_gap(gap)
		self._gap = gap
		if isinstance(gap, dict):
			if not all(k in ("s", "t") for k in gap.keys()):
				raise ValueError(gap)

	@property
	def gap(self):
		return self._gap

	def _make_args(self, locality, gaps):
		return {'algorithm': 'pyalign', 'options': {'locality': locality, 'gap_cost': gaps}}

	def to_description(self, partition):
		return {self._name: {'gap': self._gap}}

	def to_args(self, partition):
		return self._make_args(self._locality, self._gap)


class GlobalAlignment(Alignment):
	"""Needleman-Wunsch (alignment.py:47-97)."""
	_locality = Locality.GLOBAL
	_name = 'GlobalAlignment'


class SemiGlobalAlignment(Alignment):
	"""End-gaps-free alignment (alignment.py:100-130)."""
	_locality = Locality.SEMIGLOBAL
	_name = 'SemiGlobalAlignment'


class LocalAlignment(Alignment):
	"""Smith-Waterman / Waterman-Smith-Beyer (alignment.py:133-187)."""
	_locality = Locality.LOCAL
	_name = 'LocalAlignment'


class OptimalTransport(Optimizer):
	pass


class WordMoversDistance(OptimalTransport):
	"""Out of scope (optimal transport, alignment.py:190-283): construction raises."""

	def __init__(self, *args, **kwargs):
		raise NotImplementedError("WordMoversDistance is outside the B200 hot path (SURVEY.md section 2.1)")


class WordRotatorsDistance(OptimalTransport):
	def __init__(self, *args, **kwargs):
		raise NotImplementedError("WordRotatorsDistance is outside the B200 hot path (SURVEY.md section 2.1)")
