"""Pins the oracle's DP + traceback against the reference's own known-answer tests.

Reference: vectorian/core/lib/pyalign/pyalign/tests/test_{nw,sw,wsb,gotoh,batch}.py.
Those tests enumerate ALL optimal alignments; Vectorian asks for ONE
(cpp/metric/alignment.h:404-406), so the score must match exactly and the single
traceback must be one of the listed optimal edge sets (tests/__init__.py:13-24).
"""

import numpy as np
import pytest

from vectorian_b200 import _cabi, gaps


def equality(s, t, eq, ne):
	return np.array([[eq if a == b else ne for b in t] for a in s], dtype=np.float32)


def dict_sim(s, t, d):
	return np.array([[d.get((a, b), d.get((b, a), 0)) for b in t] for a in s], dtype=np.float32)


def params(locality, gap, len_s, len_t):
	r = gaps.resolve_gap_costs(gap, len_s, len_t)
	kind = r.pop("gap_kind")
	return _cabi.make_align_params(locality, kind, **r)


def check(oracle, sim, locality, gap, score, *edge_sets, places=7):
	ap, keep = params(locality, gap, *sim.shape)
	got, target = oracle.align(sim, ap)
	assert round(abs(got - score), places) == 0
	assert oracle.edges_from_target(target) in [list(map(list, e)) for e in edge_sets]
	return got, target


def test_nw_aatcg_aacg(oracle):  # test_nw.py:9-27
	check(oracle, equality("AATCG", "AACG", 1, -1), _cabi.VX_GLOBAL, gaps.LinearGapCost(2),
		2.0, [[0, 0], [1, 1], [3, 2], [4, 3]])


def test_sw_aatcg_aacg(oracle):  # test_sw.py:10-29
	got, target = check(oracle, equality("AATCG", "AACG", 1, -1), _cabi.VX_LOCAL, gaps.LinearGapCost(2),
		2, [[0, 0], [1, 1]], [[3, 2], [4, 3]])
	# the reverse-scan seed (solver.h:2042-2048) picks the later optimum
	assert oracle.edges_from_target(target) == [[3, 2], [4, 3]]


def test_wsb_cg_ccga(oracle):  # test_wsb.py:10-30
	got, target = check(oracle, equality("CG", "CCGA", 1, -1), _cabi.VX_GLOBAL, gaps.LogarithmicGapCost(3, 1),
		-3.7, [[0, 0], [1, 1]], [[0, 0], [1, 3]], places=1)
	assert np.float32(got) == np.float32(-3.6931472)


def test_gotoh_cc_acct_minimize(oracle):  # test_gotoh.py:13-31, minimize == maximize of the negated costs
	check(oracle, -equality("CC", "ACCT", 0, 1), _cabi.VX_GLOBAL, gaps.AffineGapCost(5, 1),
		-7.0, [[0, 0], [1, 1]], [[0, 2], [1, 3]])


def test_gotoh_ggtgtga_tcgcgt(oracle):  # test_gotoh.py:37-78
	d = {("T", "C"): -1, ("A", "T"): 0, ("G", "T"): -2, ("A", "A"): 6, ("T", "T"): 6, ("G", "G"): 6}
	check(oracle, dict_sim("GGTGTGA", "TCGCGT", d), _cabi.VX_GLOBAL, gaps.AffineGapCost(11, 1),
		-2, [[0, 0], [1, 1], [3, 2], [4, 3], [5, 4], [6, 5]])


def test_gotoh_aaaggg(oracle):  # test_gotoh.py:80-113
	check(oracle, equality("AAAGGG", "TTAAAAGGGGTT", 0, -1), _cabi.VX_GLOBAL, gaps.AffineGapCost(5, 1),
		-14,
		[[0, 3], [1, 4], [2, 5], [3, 6], [4, 7], [5, 8]],
		[[0, 0], [1, 1], [2, 2], [3, 9], [4, 10], [5, 11]])


def test_gotoh_taaatttgc(oracle):  # test_gotoh.py:115-143
	check(oracle, equality("TAAATTTGC", "TCGCCTTAC", 10, -30), _cabi.VX_GLOBAL, gaps.AffineGapCost(40, 1),
		-60, [[0, 0], [1, 7], [8, 8]], [[0, 6], [1, 7], [8, 8]])


@pytest.mark.parametrize("open_cost,expected", [(25, -44), (30, -54)])
def test_gotoh_agat_ctct(oracle, open_cost, expected):  # test_gotoh.py:145-170
	check(oracle, equality("AGAT", "CTCT", 10, -30), _cabi.VX_GLOBAL, gaps.AffineGapCost(open_cost, 1),
		expected, [[3, 3]])


def test_batch_problems(oracle):  # test_batch.py:10-40 (the batch equals the per-problem answers)
	g = gaps.LinearGapCost(2)
	check(oracle, equality("AATCG", "AACG", 1, -1), _cabi.VX_GLOBAL, g, 2, [[0, 0], [1, 1], [3, 2], [4, 3]])
	check(oracle, equality("AATGC", "AACG", 1, -1), _cabi.VX_GLOBAL, g, 0, [[0, 0], [1, 1], [2, 2], [3, 3]])
	check(oracle, equality("AATCG", "AGTT", 1, -1), _cabi.VX_GLOBAL, g, -2,
		[[0, 0], [1, 1], [2, 2], [3, 3]], [[0, 0], [1, 1], [2, 2], [4, 3]])
