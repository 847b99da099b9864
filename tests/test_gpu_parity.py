"""Parity of the CUDA path (through the C ABI) against the oracle on the same seeded inputs.

Bar (BASELINE.json north_star): token/index work bit-exact; with the SAME similarity matrix the
DP scores are bit-exact too (fp32 add/compare only); the similarity GEMM agrees within 1e-5.
"""

import numpy as np
import pytest

from vectorian_b200 import _cabi, gaps
from tests import helpers
from tests.helpers import GAPS, LOCALITIES, align_params

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
	from vectorian_b200 import engine as e
	_cabi.lib()
	return e


def make_both(engine, oracle, corpus, window_size=1, window_step=1, with_pos=False, boost=None):
	pos = corpus.get("token_pos") if with_pos else None
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], window_size, window_step, token_pos=pos)
	oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], window_size, window_step, token_pos=pos, boost=boost)
	if boost is not None:
		dc.set_boost(boost)
	return dc, oc


def run_case(engine, oracle, dc, oc, s, locality, gap, t, k=50, min_score=0.0, contextual=False, sim=None, **qkw):
	ap, keep = align_params(locality, gap, dc.max_len, t)
	qp, keep2 = _cabi.make_query_params(max_matches=k, min_score=min_score, **qkw)
	if sim is None:
		sim = engine.DeviceSim.upload(s)
	g_all = dc.score_all(sim, ap, qp)
	o_all = oc.score_all(s, ap, qp, n_threads=4, contextual=contextual)
	np.testing.assert_array_equal(g_all, o_all)
	gpu = dc.search(sim, ap, qp)
	ref = oc.search(s, ap, qp, n_threads=4, contextual=contextual)
	wd, wsl, _, _ = oc.windows()
	helpers.assert_hits_match(gpu, ref, o_all, wd, wsl)
	return gpu


def test_kat_through_cabi(engine, oracle):
	"""The pyalign known answers (tests/test_oracle_kat.py) through vx_search: a one-sentence corpus
	whose token ids 0..L-1 index the rows of an uploaded similarity matrix."""
	from tests.test_oracle_kat import equality
	cases = [
		(equality("AATCG", "AACG", 1, -1), _cabi.VX_GLOBAL, gaps.LinearGapCost(2), 2.0, [[0, 0], [1, 1], [3, 2], [4, 3]]),
		(equality("AATCG", "AACG", 1, -1), _cabi.VX_LOCAL, gaps.LinearGapCost(2), 2.0, [[3, 2], [4, 3]]),
		(equality("CG", "CCGA", 1, -1), _cabi.VX_GLOBAL, gaps.LogarithmicGapCost(3, 1), -3.6931472, [[0, 0], [1, 3]]),
		(equality("AGAT", "CTCT", 10, -30), _cabi.VX_GLOBAL, gaps.AffineGapCost(25, 1), -44.0, [[3, 3]]),
		(equality("GGTGTGA"[:6], "TCGCGT", 0, -1), _cabi.VX_SEMIGLOBAL, gaps.AffineGapCost(5, 1), None, None),
	]
	for sim, locality, gap, score, edges in cases:
		len_s, len_t = sim.shape
		dc = engine.DeviceCorpus(np.arange(len_s, dtype=np.int32), [0, len_s], [0], [len_s], [0, 1])
		ap, keep = align_params(locality, gap, len_s, len_t)
		qp, _ = _cabi.make_query_params(max_matches=1, min_score=-1e30)
		hits, target, dist = dc.search(engine.DeviceSim.upload(sim), ap, qp)
		o_score, o_target = oracle.align(sim, ap)
		assert len(hits) == 1
		assert np.float32(hits["score"][0]) == np.float32(o_score) / np.float32(len_t)
		np.testing.assert_array_equal(target[0], o_target)
		if score is not None:
			assert np.float32(hits["score"][0]) == np.float32(score) / np.float32(len_t)
			assert oracle.edges_from_target(target[0]) == edges


@pytest.mark.parametrize("gap_name", list(GAPS.keys()))
@pytest.mark.parametrize("locality", LOCALITIES)
def test_all_solvers_bit_exact(engine, oracle, gap_name, locality):
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=2500, t=8, seed=11)
	s = oracle.sim_static(emb, q_id, q_vec)
	dc, oc = make_both(engine, oracle, corpus)
	run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 8)


@pytest.mark.parametrize("t", [1, 2, 3, 5, 7, 12, 13, 16])
def test_query_lengths(engine, oracle, t):
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=1500, t=t, seed=23 + t)
	s = oracle.sim_static(emb, q_id, q_vec)
	dc, oc = make_both(engine, oracle, corpus)
	for locality in LOCALITIES:
		for gap_name in ("linear0.2", "affine", "smooth5"):
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], t)


def test_long_query_uses_exact_path(engine, oracle):
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=400, t=21, seed=5)
	s = oracle.sim_static(emb, q_id, q_vec)
	dc, oc = make_both(engine, oracle, corpus)
	for locality in LOCALITIES:
		run_case(engine, oracle, dc, oc, s, locality, GAPS["smooth5"], 21, k=10)
		run_case(engine, oracle, dc, oc, s, locality, GAPS["affine"], 21, k=10)


def test_similarity_gemm(engine, oracle):
	"""vx_sim_static vs the oracle: 1e-5 (summation order differs, as between BLAS builds)."""
	for t, dim in ((5, 300), (8, 300), (12, 128), (16, 50)):
		corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=10, vocab=5000, dim=dim, t=t, seed=t)
		emb[17, :] = 0  # a vanishing vector: 0/0 -> nan -> 0
		q_vec = emb[q_id].copy()
		q_id[-1] = -1  # a query token outside the vocabulary gets no identity fix-up
		de = engine.DeviceEmbedding(emb)
		sim = engine.DeviceSim.static(de, q_id, q_vec)
		got = sim.download()
		want = oracle.sim_static(emb, q_id, q_vec)
		np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
		assert np.all(got >= 0) and np.all(got <= 1)
		for j, i in enumerate(q_id):
			if i >= 0:
				assert got[i, j] == 1.0
		assert np.all(got[17] == 0)


def test_end_to_end_with_device_gemm(engine, oracle):
	"""Whole path on the device (GEMM -> DP -> top-k -> traceback) vs. the oracle, 1e-5 relative."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=3000, vocab=3000, dim=300, t=12, seed=3)
	dc, oc = make_both(engine, oracle, corpus)
	sim = engine.DeviceSim.static(engine.DeviceEmbedding(emb), q_id, q_vec)
	s = oracle.sim_static(emb, q_id, q_vec)
	ap, keep = align_params(_cabi.VX_LOCAL, gaps.smooth_gap_cost(5), dc.max_len, 12)
	qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
	g_all = dc.score_all(sim, ap, qp)
	o_all = oc.score_all(s, ap, qp, n_threads=4)
	np.testing.assert_allclose(g_all, o_all, rtol=1e-5, atol=1e-7)
	gpu = dc.search(sim, ap, qp)
	ref = oc.search(s, ap, qp, n_threads=4)
	wd, wsl, _, _ = oc.windows()
	helpers.assert_hits_match(gpu, ref, o_all, wd, wsl, rel_tol=1e-5)


def test_windows_and_boost(engine, oracle):
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=1200, t=6, seed=41)
	s = oracle.sim_static(emb, q_id, q_vec)
	rng = np.random.default_rng(0)
	boost = rng.uniform(0.5, 1.5, size=len(corpus["span_start"])).astype(np.float32)
	for ws, st in ((1, 1), (3, 1), (3, 2), (2, 5), (8, 8)):
		dc, oc = make_both(engine, oracle, corpus, ws, st, boost=boost)
		wd, wsl, wtok, wlen = oc.windows()
		assert dc.n_windows == len(wd)
		assert dc.n_cells == int(wlen.sum())
		assert dc.max_len == int(wlen.max())
		run_case(engine, oracle, dc, oc, s, _cabi.VX_LOCAL, GAPS["linear0.2"], 6)
		run_case(engine, oracle, dc, oc, s, _cabi.VX_GLOBAL, GAPS["smooth5"], 6)


def test_tag_weighted_and_submatch(engine, oracle):
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=800, t=7, seed=77, with_pos=True)
	s = oracle.sim_static(emb, q_id, q_vec)
	dc, oc = make_both(engine, oracle, corpus, with_pos=True)
	rng = np.random.default_rng(1)
	t_weights = rng.uniform(0.2, 2.0, size=7).astype(np.float32)
	t_pos = rng.integers(0, 17, size=7).astype(np.int8)
	for locality in LOCALITIES:
		for gap_name in ("linear0.2", "affine", "smooth5"):
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 7,
				t_weights=t_weights, t_pos=t_pos, pos_mismatch_penalty=0.4, similarity_threshold=0.15)
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 7, submatch_weight=0.5)


def test_tag_weighted_on_the_fast_kernels(engine, oracle):
	"""Tag-weighted similarity with weights whose sums are exact in fp32 (multiples of 2^-10; the default 1 among them)
	takes the fast DP kernels -- the weighting is applied to the gathered S row (TagWeightedSlice::similarity,
	cpp/slice/static.h:236-263) --, other weights keep the per-window scorer; both are bit-identical to the oracle,
	for every locality x gap kind, with windows, a boost and a long (> 64 tokens) sentence in the corpus."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=1500, t=12, seed=78, with_pos=True)
	s = oracle.sim_static(emb, q_id, q_vec)
	rng = np.random.default_rng(2)
	t_pos = rng.integers(0, 17, size=12).astype(np.int8)
	dyadic = (rng.integers(1, 12, size=12) / 4.0).astype(np.float32)            # 0.25 .. 2.75
	ones = np.ones(12, np.float32)
	ragged = rng.uniform(0.2, 2.0, size=12).astype(np.float32)                  # not exact: per-window scorer
	boost = rng.uniform(0.5, 1.5, size=len(corpus["span_start"])).astype(np.float32)
	dc, oc = make_both(engine, oracle, corpus, with_pos=True, boost=boost)
	for name, w in (("dyadic", dyadic), ("ones", ones), ("ragged", ragged)):
		for locality in LOCALITIES:
			for gap_name in ("linear0.2", "affine", "smooth5"):
				run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 12,
					t_weights=w, t_pos=t_pos, pos_mismatch_penalty=0.4, similarity_threshold=0.15)
	dc.close()
	# windows of three sentences, step two
	dc, oc = make_both(engine, oracle, corpus, window_size=3, window_step=2, with_pos=True)
	for gap_name in ("linear0.2", "smooth5"):
		run_case(engine, oracle, dc, oc, s, _cabi.VX_LOCAL, GAPS[gap_name], 12,
			t_weights=dyadic, t_pos=t_pos, pos_mismatch_penalty=0.25, similarity_threshold=0.0)
	dc.close()


def test_edge_cases(engine, oracle):
	# ragged: empty document, empty spans, single-token sentences, a sentence longer than 64 tokens
	token_id = np.arange(200, dtype=np.int32) % 50
	doc_tok_off = [0, 0, 120, 200]
	span_start = [0, 1, 1, 2, 100, 0, 80]
	span_end = [1, 1, 2, 100, 120, 80, 80]
	doc_span_off = [0, 0, 5, 7]
	corpus = dict(token_id=token_id, doc_tok_off=doc_tok_off, span_start=span_start, span_end=span_end, doc_span_off=doc_span_off)
	rng = np.random.default_rng(5)
	s = rng.uniform(0, 1, size=(50, 5)).astype(np.float32)
	dc, oc = make_both(engine, oracle, corpus)
	assert dc.n_windows == 5
	for locality in LOCALITIES:
		for gap_name in ("none", "linear0.2", "affine", "smooth5"):
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 5, k=3)
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], 5, k=100, min_score=0.3)
	# empty query -> no matches (matcher_impl.h:80-83); empty corpus -> no matches
	ap, _ = align_params(_cabi.VX_LOCAL, None, dc.max_len, 0)
	qp, _ = _cabi.make_query_params(max_matches=10, min_score=0.0)
	hits, _, _ = dc.search(engine.DeviceSim.upload(np.zeros((50, 0), dtype=np.float32)), ap, qp)
	assert len(hits) == 0
	empty = engine.DeviceCorpus(np.zeros(0, np.int32), [0], [], [], [0])
	assert empty.n_windows == 0
	hits, _, _ = empty.search(engine.DeviceSim.upload(s), ap, qp)
	assert len(hits) == 0


def test_error_behaviour(engine):
	"""Same failures as the reference, as VxError (the reference raises RuntimeError/ValueError)."""
	tok = np.zeros(10, np.int32)
	with pytest.raises(_cabi.VxError, match="window size"):
		engine.DeviceCorpus(tok, [0, 10], [0], [10], [0, 1], window_size=0)
	with pytest.raises(_cabi.VxError, match="window step"):
		engine.DeviceCorpus(tok, [0, 10], [0], [10], [0, 1], window_step=0)
	dc = engine.DeviceCorpus(tok, [0, 10], [0], [10], [0, 1])
	sim = engine.DeviceSim.upload(np.ones((1, 3), np.float32))
	ap, keep = _cabi.make_align_params(_cabi.VX_LOCAL, _cabi.VX_GAP_GENERAL, w_s=np.zeros(4), w_t=np.zeros(4))
	qp, _ = _cabi.make_query_params(max_matches=5, min_score=0.0)
	with pytest.raises(_cabi.VxError, match="expected gap cost tensor length of 11, got 4"):
		dc.search(sim, ap, qp)
	ap2, _ = _cabi.make_align_params(7, _cabi.VX_GAP_LINEAR)
	with pytest.raises(_cabi.VxError, match="invalid locality"):
		dc.search(sim, ap2, qp)
	qp0, _ = _cabi.make_query_params(max_matches=0, min_score=0.0)
	ap3, _ = _cabi.make_align_params(_cabi.VX_LOCAL, _cabi.VX_GAP_LINEAR)
	with pytest.raises(_cabi.VxError, match="max_matches"):
		dc.search(sim, ap3, qp0)
	with pytest.raises(_cabi.VxError, match="fewer rows"):
		engine.DeviceCorpus(np.full(10, 5, np.int32), [0, 10], [0], [10], [0, 1]).search(sim, ap3, qp)


def test_many_ties_topk(engine, oracle):
	"""All windows identical -> every score ties; the order must be (doc desc, slice desc)."""
	n = 5000
	token_id = np.zeros(n * 4, np.int32)
	starts = (np.arange(n) % 100) * 4
	doc_span_off = np.arange(0, n + 1, 100)
	doc_tok_off = doc_span_off * 4
	corpus = dict(token_id=token_id, doc_tok_off=doc_tok_off, span_start=starts, span_end=starts + 4, doc_span_off=doc_span_off)
	s = np.full((1, 3), 0.5, np.float32)
	dc, oc = make_both(engine, oracle, corpus)
	ap, _ = align_params(_cabi.VX_LOCAL, None, 4, 3)
	qp, _ = _cabi.make_query_params(max_matches=130, min_score=0.0)
	hits, target, dist = dc.search(engine.DeviceSim.upload(s), ap, qp)
	assert len(hits) == 130
	assert np.all(hits["score"] == hits["score"][0])
	want = [(49 - i // 100, 99 - i % 100) for i in range(130)]
	assert [(int(h["doc"]), int(h["slice"])) for h in hits] == want


@pytest.mark.parametrize("t", [17, 24, 27, 32, 33, 47, 50, 64])
def test_queries_of_17_to_64_tokens_on_the_score_kernels(engine, oracle, t):
	"""Queries longer than 16 tokens (a whole sentence as the query) take the linear (up to 64 tokens) / affine (up to 48) /
	general-scratch (up to 32) score kernels instead of the one-thread-per-window scorer, which keeps the rest: bit-identical
	scores, identical top-k and tracebacks, for every locality x gap kind, incl. windows longer than 64 tokens and a boost."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=900, t=t, seed=90 + t)
	s = oracle.sim_static(emb, q_id, q_vec)
	rng = np.random.default_rng(t)
	boost = rng.uniform(0.5, 1.5, size=len(corpus["span_start"])).astype(np.float32)
	dc, oc = make_both(engine, oracle, corpus, boost=boost)
	for locality in LOCALITIES:
		for gap_name in ("linear0.2", "linear_st", "affine", "smooth5"):
			run_case(engine, oracle, dc, oc, s, locality, GAPS[gap_name], t)
	dc.close()
	dc, oc = make_both(engine, oracle, corpus, window_size=4, window_step=3)     # windows of up to 4 x 64 tokens
	run_case(engine, oracle, dc, oc, s, _cabi.VX_LOCAL, GAPS["linear0.2"], t)
	run_case(engine, oracle, dc, oc, s, _cabi.VX_SEMIGLOBAL, GAPS["affine"], t)
	dc.close()


@pytest.mark.parametrize("t", [5, 12, 16])
def test_general_gap_costs_over_long_windows(engine, oracle, t):
	"""Windows of up to five sentences (up to 320 tokens) with general gap costs: the tensor-memory WSB kernel takes
	windows of up to 160 tokens (fewer warps per SM for the long classes), the scratch kernel the rest -- bit-identical
	to the oracle for every locality, with and without tag weighting."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=700, t=t, seed=60 + t, with_pos=True)
	s = oracle.sim_static(emb, q_id, q_vec)
	rng = np.random.default_rng(t)
	t_pos = rng.integers(0, 17, size=t).astype(np.int8)
	w = (rng.integers(1, 9, size=t) / 4.0).astype(np.float32)
	for window_size, window_step in ((3, 2), (5, 3)):
		dc, oc = make_both(engine, oracle, corpus, window_size=window_size, window_step=window_step, with_pos=True)
		assert dc.max_len > 64
		for locality in LOCALITIES:
			run_case(engine, oracle, dc, oc, s, locality, GAPS["smooth5"], t)
			run_case(engine, oracle, dc, oc, s, locality, GAPS["mixed"], t)
		run_case(engine, oracle, dc, oc, s, _cabi.VX_LOCAL, GAPS["log"], t,
			t_weights=w, t_pos=t_pos, pos_mismatch_penalty=0.4, similarity_threshold=0.1)
		dc.close()


def test_more_ties_than_the_candidate_buffer_holds(engine):
	"""12 000 identical windows: every score ties, the pruned top-k overflows its 8192-key candidate buffer and the search
	(single, setting-fused batch, query-fused batch) must fall back to the exact radix select: the order is still
	(doc desc, slice desc)."""
	n, per_doc = 12000, 100
	token_id = np.zeros(n * 4, np.int32)
	starts = (np.arange(n) % per_doc) * 4
	doc_span_off = np.arange(0, n + 1, per_doc)
	dc = engine.DeviceCorpus(token_id, doc_span_off * 4, starts, starts + 4, doc_span_off)
	sims = [engine.DeviceSim.upload(np.full((1, 3), 0.5, np.float32)), engine.DeviceSim.upload(np.full((1, 2), 0.25, np.float32))]
	qp, _ = _cabi.make_query_params(max_matches=130, min_score=0.0)
	want = [(n // per_doc - 1 - i // per_doc, per_doc - 1 - i % per_doc) for i in range(130)]
	aps, keeps = [], []
	for locality, gap_name in ((_cabi.VX_LOCAL, "none"), (_cabi.VX_LOCAL, "linear0.2"), (_cabi.VX_GLOBAL, "none"), (_cabi.VX_GLOBAL, "linear0.2")):
		ap, keep = align_params(locality, GAPS[gap_name], 4, 3)
		aps.append(ap)
		keeps.append(keep)
	for sim in sims:
		hits, _, _ = dc.search(sim, aps[0], qp)
		assert len(hits) == 130 and np.all(hits["score"] == hits["score"][0])
		assert [(int(h["doc"]), int(h["slice"])) for h in hits] == want
	batch = dc.search_batch(sims, aps, qp)             # settings of a query fused
	mq = dc.search_batch(sims, aps[:1], qp)            # queries of a setting fused
	for q, sim in enumerate(sims):
		for a, ap in enumerate(aps):
			single = dc.search(sim, ap, qp)
			for got in ([batch[q][a]] + ([mq[q][0]] if a == 0 else [])):
				assert len(got[0]) == len(single[0])
				for f in ("score", "doc", "slice"):
					np.testing.assert_array_equal(got[0][f], single[0][f])
				np.testing.assert_array_equal(got[1], single[1])
			if len(single[0]):
				assert [(int(h["doc"]), int(h["slice"])) for h in single[0]] == want[:len(single[0])]
	for sim in sims:
		sim.close()
	dc.close()


@pytest.mark.parametrize("gap_name", ["linear0.2", "affine", "smooth5"])
def test_contextual_fast_path(engine, oracle, gap_name):
	"""Per-token contextual vectors: S_doc formed on the device (ctx_sim_kernel), then the same DP kernels.
	fp32 scores within 1e-5 of the oracle (the dot products sum in a different order)."""
	rng = np.random.default_rng(77)
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=1200, t=7, seed=31)
	n_tok = len(corpus["token_id"])
	for dim in (32, 30):   # 16-byte vector path and the scalar path
		centroids = rng.standard_normal((40, dim)).astype(np.float32)
		x = centroids[corpus["token_id"] % 40] + 0.4 * rng.standard_normal((n_tok, dim)).astype(np.float32)
		q = centroids[rng.integers(0, 40, size=7)] + 0.4 * rng.standard_normal((7, dim)).astype(np.float32)
		dc, oc = make_both(engine, oracle, corpus)
		dc.set_contextual(x)
		sim = engine.DeviceSim.contextual(q)
		s = oracle.sim_contextual(x, q)
		for locality in LOCALITIES:
			ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, 7)
			qp, _ = _cabi.make_query_params(max_matches=30, min_score=0.0)
			g_all = dc.score_all(sim, ap, qp)
			o_all = oc.score_all(s, ap, qp, n_threads=4, contextual=True)
			np.testing.assert_allclose(g_all, o_all, rtol=1e-5, atol=2e-6)
			gpu = dc.search(sim, ap, qp)
			ref = oc.search(s, ap, qp, n_threads=4, contextual=True)
			wd, wsl, _, _ = oc.windows()
			helpers.assert_hits_match(gpu, ref, o_all, wd, wsl, rel_tol=1e-5)
			assert dc.last_launches() <= 8   # the fast kernels, not the exact per-window scorer


def test_search_batch_equals_single_searches(engine, oracle):
	"""vx_search_batch (parameter sweep, BASELINE config 5): queries x settings in one call give exactly the rows of
	the individual vx_search calls, for queries of different lengths and for every gap kind."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=2500, t=8, seed=11)
	dc, oc = make_both(engine, oracle, corpus)
	de = engine.DeviceEmbedding(emb)
	rng = np.random.default_rng(3)
	sims = []
	for t in (8, 5, 12, 3):
		ids = rng.choice(emb.shape[0], size=t, replace=False).astype(np.int32)
		sims.append(engine.DeviceSim.static(de, ids, emb[ids]))
	# an uploaded matrix with negative entries: the fused kernel must keep the explicit Local zero
	sims.append(engine.DeviceSim.upload((rng.random((emb.shape[0], 6)) * 1.3 - 0.3).astype(np.float32)))
	aps, keeps = [], []
	for gap_name in ("linear0.2", "none", "affine", "smooth5"):
		for locality in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
			ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, 12)
			aps.append(ap)
			keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=20, min_score=0.0)
	batch = dc.search_batch(sims, aps, qp)
	assert len(batch) == len(sims) and len(batch[0]) == len(aps)
	for q, sim in enumerate(sims):
		for a, ap in enumerate(aps):
			hits, target, dist = dc.search(sim, ap, qp)
			b_hits, b_target, b_dist = batch[q][a]
			assert len(b_hits) == len(hits)
			for f in ("score", "score_max", "doc", "slice"):
				np.testing.assert_array_equal(b_hits[f], hits[f])
			np.testing.assert_array_equal(b_target, target)
			np.testing.assert_array_equal(b_dist, dist)
	# empty batch
	assert dc.search_batch([], aps, qp) == []


@pytest.mark.parametrize("n_queries", [2, 3, 8, 13])
def test_search_batch_several_queries_per_pass(engine, oracle, n_queries):
	"""vx_search_batch with several queries and ONE linear-gap setting per locality: groups of 8 / 4 / 2 queries share a
	pass over the corpus (score_linear_mq_kernel: interleaved similarity matrices, a lane per (window, query) pair), the
	rest run alone -- every pair returns exactly the rows of its own vx_search, for queries of different lengths, an
	uploaded matrix with negative entries, a boost, and a setting (affine) that is not fused."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=3100, t=8, seed=12)
	rng = np.random.default_rng(4 + n_queries)
	boost = rng.uniform(0.5, 1.5, size=len(corpus["span_start"])).astype(np.float32)
	dc, oc = make_both(engine, oracle, corpus, boost=boost)
	de = engine.DeviceEmbedding(emb)
	sims = []
	for i in range(n_queries):
		t = int(rng.integers(1, 13))
		if i == 1:
			sims.append(engine.DeviceSim.upload((rng.random((emb.shape[0], t)) * 1.3 - 0.3).astype(np.float32)))
		else:
			ids = rng.choice(emb.shape[0], size=t, replace=False).astype(np.int32)
			sims.append(engine.DeviceSim.static(de, ids, emb[ids]))
	aps, keeps = [], []
	for locality, gap_name in ((_cabi.VX_LOCAL, "linear0.2"), (_cabi.VX_GLOBAL, "linear_st"), (_cabi.VX_SEMIGLOBAL, "none"), (_cabi.VX_LOCAL, "affine")):
		ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, 12)
		aps.append(ap)
		keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=30, min_score=0.0)
	batch = dc.search_batch(sims, aps, qp)
	for q, sim in enumerate(sims):
		for a, ap in enumerate(aps):
			hits, target, dist = dc.search(sim, ap, qp)
			b_hits, b_target, b_dist = batch[q][a]
			assert len(b_hits) == len(hits), (q, a)
			for f in ("score", "score_max", "doc", "slice"):
				np.testing.assert_array_equal(b_hits[f], hits[f], err_msg=f"{q} {a} {f}")
			np.testing.assert_array_equal(b_target, target)
			np.testing.assert_array_equal(b_dist, dist)
	for sim in sims:
		sim.close()
	dc.close()


def corpus_with_lengths(lens, vocab, seed):
	"""A one-sentence-per-span corpus with the given sentence lengths (documents of 53 sentences)."""
	rng = np.random.default_rng(seed)
	lens = np.asarray(lens, dtype=np.int64)
	n = len(lens)
	token_id = rng.integers(0, vocab, size=int(lens.sum())).astype(np.int32)
	n_doc = (n + 52) // 53
	doc_span_off = np.minimum(np.arange(n_doc + 1, dtype=np.int64) * 53, n)
	ends = np.cumsum(lens)
	starts = ends - lens
	doc_tok_off = np.concatenate([starts[doc_span_off[:-1]], [int(lens.sum())]]).astype(np.int64)
	doc_of = np.repeat(np.arange(n_doc), np.diff(doc_span_off))
	return dict(token_id=token_id, doc_tok_off=doc_tok_off, span_start=(starts - doc_tok_off[doc_of]).astype(np.int32),
		span_end=(ends - doc_tok_off[doc_of]).astype(np.int32), doc_span_off=doc_span_off)


@pytest.mark.parametrize("t", [4, 7, 12, 16])
def test_wsb_length_class_planner_on_odd_length_distributions(engine, oracle, t):
	"""The tensor-memory WSB launch plan (groups of length classes, the long queue inside the 8-warp launch, vx_wsb.cu) on
	corpora whose window lengths sit where the plan changes: only just above the 8-warp classes, two far-apart modes, a long
	queue that holds too much of the work to be absorbed, every length once, one single long window.  Every window score and
	the top-k equal the oracle's."""
	rng = np.random.default_rng(100 + t)
	emb = rng.standard_normal((300, 16)).astype(np.float32)
	q_id = rng.choice(300, size=t, replace=False).astype(np.int32)
	s = oracle.sim_static(emb, q_id, emb[q_id])
	cases = [
		rng.integers(33, 41, size=1300),                                                   # nothing at or below 32 tokens
		np.concatenate([rng.integers(3, 12, size=2500), rng.integers(41, 49, size=160)]),  # two modes
		np.concatenate([rng.integers(3, 30, size=600), rng.integers(37, 45, size=900)]),   # long classes hold most of the work
		np.concatenate([np.arange(1, 161), np.arange(1, 161)]),                            # every length up to 160 twice
		np.concatenate([np.full(2000, 5), [150]]),                                         # one long window among short ones
		rng.integers(29, 37, size=1500),                                                   # only the top classes of the 8-warp group
	]
	for i, lens in enumerate(cases):
		corpus = corpus_with_lengths(rng.permutation(lens), 300, seed=7 * i + t)
		dc, oc = make_both(engine, oracle, corpus)
		for locality in LOCALITIES:
			run_case(engine, oracle, dc, oc, s, locality, GAPS["smooth5"], t, k=40)
		run_case(engine, oracle, dc, oc, s, _cabi.VX_LOCAL, GAPS["log"], t, k=40)
		dc.close()


def test_search_batch_with_queries_longer_than_the_fused_kernels_take(engine):
	"""A batch whose queries are 20, 40 and 64 tokens long next to two short ones: the long ones have no fused (multi-query /
	multi-setting) kernel and run pair by pair inside the same batch; every pair equals its own vx_search."""
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=1500, t=8, seed=31)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"], 1, 1)
	rng = np.random.default_rng(77)
	sims = [engine.DeviceSim.upload(rng.random((emb.shape[0], t)).astype(np.float32)) for t in (6, 40, 9, 64, 20)]
	aps, keeps = [], []
	for locality, gap_name in ((_cabi.VX_LOCAL, "linear0.2"), (_cabi.VX_GLOBAL, "linear_st"), (_cabi.VX_LOCAL, "affine"), (_cabi.VX_LOCAL, "linear_st")):
		ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, 64)
		aps.append(ap)
		keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=25, min_score=-1e30)
	batch = dc.search_batch(sims, aps, qp)
	for q, sim in enumerate(sims):
		for a, ap in enumerate(aps):
			hits, target, dist = dc.search(sim, ap, qp)
			b_hits, b_target, b_dist = batch[q][a]
			assert len(b_hits) == len(hits) == 25, (q, a)
			for f in ("score", "score_max", "doc", "slice"):
				np.testing.assert_array_equal(b_hits[f], hits[f], err_msg=f"{q} {a} {f}")
			np.testing.assert_array_equal(b_target, target)
			np.testing.assert_array_equal(b_dist, dist)
	for sim in sims:
		sim.close()
	dc.close()


@pytest.mark.parametrize("gap_name", ["linear0.2", "affine", "smooth5", "mixed"])
def test_gpu_scores_equal_reference_solver(engine, gap_name):
	"""The CUDA path against the REFERENCE'S OWN solver (oracle/_ref: unmodified pyalign solver.h, oracle/ref/):
	with the similarity matrix the device built, every sentence score is bit-identical to raw / T of the
	reference solver, for all three localities (Global + affine: border cells go through xt::linspace in the
	reference, last-bit differences allowed -- see tests/test_oracle_vs_ref.py)."""
	from oracle import ref
	if not ref.available():
		pytest.skip("oracle/_ref/libpyalign_ref.so not built")
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=4000, t=7, seed=23)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], 1, 1)
	de = engine.DeviceEmbedding(emb)
	sim = engine.DeviceSim.static(de, q_id, q_vec)
	s = sim.download()
	lens = corpus["lens"]
	sent_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
	for locality in LOCALITIES:
		ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, 7)
		qp, _ = _cabi.make_query_params(max_matches=10, min_score=0.0)
		g_all = dc.score_all(sim, ap, qp)
		resolved = gaps.resolve_gap_costs(GAPS[gap_name], dc.max_len, 7)
		raw = ref.corpus_scores(locality, resolved, s, corpus["token_id"], sent_off, dc.max_len, n_threads=4, doc_len=97)
		expect = raw / np.float32(7)
		if locality == _cabi.VX_GLOBAL and resolved["gap_kind"] == 1:
			np.testing.assert_allclose(g_all, expect, rtol=1e-6, atol=1e-6)
		else:
			np.testing.assert_array_equal(g_all, expect)
	sim.close()
	dc.close()


@pytest.mark.parametrize("t,dim", [(7, 32), (12, 128), (16, 64), (3, 96)])
def test_contextual_fused_equals_unfused(engine, t, dim):
	"""Contextual + linear gap cost: the kernel that fuses the tcgen05 similarity GEMM with the DP (ctx_mma_kernel<TP, true>)
	gives bit for bit the scores, hits and tracebacks of the two-kernel path (GEMM, then score_linear_kernel)."""
	import os
	rng = np.random.default_rng(100 + t)
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=5000, t=t, seed=41)
	n_tok = len(corpus["token_id"])
	centroids = rng.standard_normal((50, dim)).astype(np.float32)
	x = centroids[corpus["token_id"] % 50] + 0.4 * rng.standard_normal((n_tok, dim)).astype(np.float32)
	q = centroids[rng.integers(0, 50, size=t)] + 0.4 * rng.standard_normal((t, dim)).astype(np.float32)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], 1, 1)
	dc.set_contextual(x)
	boost = (0.5 + rng.random(dc.n_windows)).astype(np.float32)
	for use_boost in (False, True):
		if use_boost:
			dc.set_boost(boost)
		for locality in LOCALITIES:
			for gap_name in ("linear0.2", "linear_st", "none"):
				ap, keep = align_params(locality, GAPS[gap_name], dc.max_len, t)
				qp, _ = _cabi.make_query_params(max_matches=40, min_score=-1e30)
				out = {}
				for mode in ("fused", "unfused"):
					if mode == "unfused":
						os.environ["VX_CTX_NOFUSE"] = "1"
					try:
						sim = engine.DeviceSim.contextual(q)
						scores = dc.score_all(sim, ap, qp)
						launches = dc.last_launches()
						sim2 = engine.DeviceSim.contextual(q)
						hits, target, dist = dc.search(sim2, ap, qp)
						out[mode] = (scores, hits.copy(), target.copy(), dist.copy(), launches)
						sim.close()
						sim2.close()
					finally:
						os.environ.pop("VX_CTX_NOFUSE", None)
				f, u = out["fused"], out["unfused"]
				assert f[4] < u[4]                     # one kernel instead of two
				np.testing.assert_array_equal(f[0], u[0])
				for name in ("score", "score_max", "doc", "slice"):
					np.testing.assert_array_equal(f[1][name], u[1][name])
				np.testing.assert_array_equal(f[2], u[2])
				np.testing.assert_array_equal(f[3], u[3])
	dc.close()


@pytest.mark.parametrize("t,dim", [(16, 64), (8, 128)])
def test_contextual_fused_is_race_free(engine, t, dim):
	"""The warp-specialised contextual kernel (TMA ring -> converters -> tcgen05 -> epilogue -> DP warp) gives the same
	bits on every run.  Regression test for a shared-memory stage that was handed back to the TMA producer before the
	converters' loads had completed (about 3 % of the runs of the (16, 64) case differed in a few windows; found with
	scripts/stress_ctx.py)."""
	rng = np.random.default_rng(100 + t)
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=5000, t=t, seed=41)
	n_tok = len(corpus["token_id"])
	centroids = rng.standard_normal((50, dim)).astype(np.float32)
	x = centroids[corpus["token_id"] % 50] + 0.4 * rng.standard_normal((n_tok, dim)).astype(np.float32)
	q = centroids[rng.integers(0, 50, size=t)] + 0.4 * rng.standard_normal((t, dim)).astype(np.float32)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], 1, 1)
	dc.set_contextual(x)
	ap, keep = align_params(_cabi.VX_LOCAL, GAPS["linear0.2"], dc.max_len, t)
	qp, _ = _cabi.make_query_params(max_matches=40, min_score=-1e30)
	first = None
	for rep in range(250):
		sim = engine.DeviceSim.contextual(q)
		scores = dc.score_all(sim, ap, qp)
		sim.close()
		if first is None:
			first = scores
		else:
			np.testing.assert_array_equal(scores, first, err_msg=f"run {rep}")
	dc.close()


@pytest.mark.parametrize("k,min_score", [(100, 0.0), (7, 0.0), (40, 0.5), (500, 0.0), (2000, 0.9)])
def test_gpu_top_k_equals_reference_result_set(engine, k, min_score):
	"""a12 pinned to the reference itself: the device top-k (pruned selection, or the radix select it falls back to) returns
	the winners of the reference's own, unmodified ResultSet + Match::compare_by_score (oracle/ref/resultset_driver.cpp)
	on an input full of score ties across documents -- same scores, same documents, same slices, same order."""
	from oracle import ref
	from tests.test_oracle_vs_ref import tie_problem
	if not ref.result_set_available():
		pytest.skip("oracle/_ref/libresultset_ref.so not built")
	corpus, s, raw, doc, slice_id = tie_problem()
	want_doc, want_slice, want_score = ref.result_set(raw, np.ones_like(raw), doc, slice_id, k, min_score)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
	sim = engine.DeviceSim.upload(s)
	ap, _ = _cabi.make_align_params(_cabi.VX_LOCAL, _cabi.VX_GAP_LINEAR, lin=(0.0, 0.0))
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=min_score)
	hits, _, _ = dc.search(sim, ap, qp)
	np.testing.assert_array_equal(hits["score"], want_score)
	np.testing.assert_array_equal(hits["doc"], want_doc)
	np.testing.assert_array_equal(hits["slice"], want_slice)
	np.testing.assert_array_equal(dc.search_topk_noalloc(sim, ap, qp)["doc"] if k <= _cabi.VX_NOALLOC_MAX_K else want_doc, want_doc)
	sim.close()
	dc.close()
