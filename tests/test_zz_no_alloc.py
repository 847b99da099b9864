"""vx_search_topk_noalloc (no counterpart in the reference): the hits of
vx_search through the entry point whose top-k scratch is a static device array and which allocates nothing.

Collected last on purpose."""

import numpy as np
import pytest

from vectorian_b200 import _cabi, engine

from .helpers import GAPS, align_params, small_problem

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def problem():
	corpus, emb, q_id, q_vec = small_problem(n_sentences=6000, t=12, seed=31)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"])
	de = engine.DeviceEmbedding(emb)
	sims = [engine.DeviceSim.static(de, q_id, q_vec)]
	from vectorian_b200 import synthetic
	q2 = synthetic.make_query(7, emb.shape[0], seed=77)
	sims.append(engine.DeviceSim.static(de, q2, emb[q2].copy()))
	yield dc, sims
	for s in sims:
		s.close()
	de.close()
	dc.close()


def _same_hits(a, b):
	assert len(a) == len(b), (len(a), len(b))
	for name in ("score", "score_max", "doc", "slice"):
		np.testing.assert_array_equal(a[name], b[name], err_msg=name)   # bit-exact: the bar of the other DP tests


@pytest.mark.parametrize("locality,gap", [(_cabi.VX_LOCAL, "smooth5"), (_cabi.VX_GLOBAL, "linear0.2"), (_cabi.VX_SEMIGLOBAL, "affine")])
def test_noalloc_equals_search_on_consecutive_calls(problem, locality, gap):
	dc, sims = problem
	# (query, k, min_score): the second call is a different query with fewer winners, the third finds nothing, the
	# fourth repeats the first -- all through the same static scratch, back to back
	calls = [(0, 100, 0.0), (1, 20, 0.05), (0, 50, 5.0), (0, 100, 0.0)]
	expected = []
	for qi, k, min_score in calls:
		ap, keep = align_params(locality, GAPS[gap], dc.max_len, sims[qi].t)
		qp, _ = _cabi.make_query_params(max_matches=k, min_score=min_score)
		expected.append(dc.search(sims[qi], ap, qp)[0].copy())
	got = []
	for qi, k, min_score in calls:
		ap, keep = align_params(locality, GAPS[gap], dc.max_len, sims[qi].t)
		qp, _ = _cabi.make_query_params(max_matches=k, min_score=min_score)
		got.append(dc.search_topk_noalloc(sims[qi], ap, qp).copy())
	assert len(expected[0]) == 100 and len(expected[2]) == 0
	for e, g in zip(expected, got):
		_same_hits(g, e)


def test_noalloc_allocates_no_device_memory(problem):
	import torch
	dc, sims = problem
	ap, keep = align_params(_cabi.VX_LOCAL, GAPS["smooth5"], dc.max_len, sims[0].t)
	qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
	first = dc.search_topk_noalloc(sims[0], ap, qp).copy()   # kernels loaded, symbol address resolved
	free0 = torch.cuda.mem_get_info()[0]
	for _ in range(3):
		_same_hits(dc.search_topk_noalloc(sims[0], ap, qp), first)
	assert torch.cuda.mem_get_info()[0] >= free0


def test_noalloc_rejects_what_does_not_fit(problem):
	dc, sims = problem
	ap, keep = align_params(_cabi.VX_LOCAL, GAPS["smooth5"], dc.max_len, sims[0].t)
	qp, _ = _cabi.make_query_params(max_matches=_cabi.VX_NOALLOC_MAX_K + 1, min_score=0.0)
	with pytest.raises(_cabi.VxError, match="VX_NOALLOC_MAX_K"):
		dc.search_topk_noalloc(sims[0], ap, qp)
	qp, keep2 = _cabi.make_query_params(max_matches=10, min_score=0.0, t_weights=np.ones(sims[0].t, np.float32),
		t_pos=np.zeros(sims[0].t, np.int8))
	with pytest.raises(_cabi.VxError, match="use vx_search"):
		dc.search_topk_noalloc(sims[0], ap, qp)
	# the largest allowed k works
	qp, _ = _cabi.make_query_params(max_matches=_cabi.VX_NOALLOC_MAX_K, min_score=0.0)
	_same_hits(dc.search_topk_noalloc(sims[0], ap, qp), dc.search(sims[0], ap, qp)[0])
