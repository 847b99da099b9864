"""The oracle (oracle/vx_oracle.c) against the REFERENCE'S OWN solver: the unmodified
vectorian/core/lib/pyalign/pyalign/algorithm/solver.h compiled against an xtensor stand-in into
oracle/_ref/libpyalign_ref.so (oracle/ref/).  Scores must be bit-identical and the single traceback identical,
for every locality x gap-cost kind, including inputs full of ties.  Skipped where the library was not built
(it is built in the development container from /root/reference and travels with the repository snapshot)."""

import numpy as np
import pytest

from vectorian_b200 import _cabi, gaps
from oracle import ref
from tests.helpers import GAPS, LOCALITIES

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libpyalign_ref.so not built")


def equality(s, t, eq, ne):
	return np.array([[eq if a == b else ne for b in t] for a in s], dtype=np.float32)


def both(oracle, sim, locality, gap, max_len_s=None, max_len_t=None):
	len_s, len_t = sim.shape
	r = gaps.resolve_gap_costs(gap, max_len_s or len_s, max_len_t or len_t)
	solver = ref.RefSolver(locality, r, max_len_s or len_s, max_len_t or len_t)
	got_ref = solver.align(sim)
	solver.close()
	r2 = dict(gaps.resolve_gap_costs(gap, len_s, len_t))
	kind = r2.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r2)
	return got_ref, oracle.align(sim, ap)


def test_reference_known_answers_through_ref(oracle):
	"""pyalign/tests/test_nw.py:9-27, test_sw.py:10-29, test_wsb.py:10-30 through the compiled reference solver."""
	(s, t), (so, to) = both(oracle, equality("AATCG", "AACG", 1, -1), _cabi.VX_GLOBAL, gaps.LinearGapCost(2))
	assert s == 2.0 and oracle.edges_from_target(t) == [[0, 0], [1, 1], [3, 2], [4, 3]]
	assert so == s and list(to) == list(t)
	(s, t), (so, to) = both(oracle, equality("AATCG", "AACG", 1, -1), _cabi.VX_LOCAL, gaps.LinearGapCost(2))
	assert s == 2.0 and oracle.edges_from_target(t) == [[3, 2], [4, 3]]
	assert so == s and list(to) == list(t)
	(s, t), (so, to) = both(oracle, equality("CG", "CCGA", 1, -1), _cabi.VX_GLOBAL, gaps.LogarithmicGapCost(3, 1))
	assert s == np.float32(-3.6931472) and oracle.edges_from_target(t) in ([[0, 0], [1, 1]], [[0, 0], [1, 3]])
	assert so == s and list(to) == list(t)


@pytest.mark.parametrize("gap_name", sorted(GAPS.keys()))
@pytest.mark.parametrize("locality", LOCALITIES)
def test_oracle_equals_reference_solver(oracle, locality, gap_name):
	rng = np.random.default_rng(abs(hash((locality, gap_name))) % (2 ** 31))
	gap = GAPS[gap_name]
	max_len_s, max_len_t = 40, 16
	r = gaps.resolve_gap_costs(gap, max_len_s, max_len_t)
	solver = ref.RefSolver(locality, r, max_len_s, max_len_t)   # one instance, many sentences
	n_ties = n_flip = 0
	linspace_border = locality == _cabi.VX_GLOBAL and r["gap_kind"] == 1
	for case in range(120):
		len_s, len_t = int(rng.integers(1, max_len_s + 1)), int(rng.integers(1, max_len_t + 1))
		mode = case % 3
		if mode == 0:
			sim = rng.random((len_s, len_t), dtype=np.float32)
		elif mode == 1:
			sim = (rng.integers(0, 4, size=(len_s, len_t)) / 3.0).astype(np.float32)      # many exact ties
		else:
			sim = np.clip(rng.standard_normal((len_s, len_t)) * 0.5 + 0.2, 0, 1).astype(np.float32)   # many zeros
		s_ref, t_ref = solver.align(sim)
		r2 = dict(gaps.resolve_gap_costs(gap, len_s, len_t))
		kind = r2.pop("gap_kind")
		ap, keep = _cabi.make_align_params(locality, kind, **r2)
		s_or, t_or = oracle.align(sim, ap)
		if linspace_border:
			# Global + affine: the reference fills the border with xt::linspace(0, (n - 1) u, n) + v (solver.h:2868-2872),
			# i.e. through xtensor's own arithmetic (un-vendored; restated in the shim as double(step) * k), the oracle
			# and the kernels with u k + v in float: border cells may differ in the last bit
			assert abs(float(s_or) - float(s_ref)) <= 5e-7 * max(1.0, abs(float(s_ref))), (case, s_or, s_ref)
			n_flip += list(t_or) != list(t_ref)
		else:
			assert np.float32(s_or).tobytes() == np.float32(s_ref).tobytes(), (case, len_s, len_t, s_or, s_ref)
			assert list(t_or) == list(t_ref), (case, len_s, len_t, list(t_or), list(t_ref))
		n_ties += mode == 1
	solver.close()
	assert n_ties > 0 and n_flip <= 6   # a last-bit difference may flip a tie; it must stay rare


@pytest.mark.parametrize("gap_name", ["linear0.2", "linear_st", "affine", "smooth5"])
@pytest.mark.parametrize("locality", LOCALITIES)
def test_oracle_equals_reference_solver_on_long_queries(oracle, locality, gap_name):
	"""Queries of 17 .. 64 tokens (a whole sentence as the query): the GPU score kernels for these lengths are checked against
	the oracle (tests/test_gpu_parity.py), so the oracle is pinned to the reference's own solver at these lengths too."""
	rng = np.random.default_rng(abs(hash((locality, gap_name, "long"))) % (2 ** 31))
	gap = GAPS[gap_name]
	max_len_s, max_len_t = 70, 64
	r = gaps.resolve_gap_costs(gap, max_len_s, max_len_t)
	solver = ref.RefSolver(locality, r, max_len_s, max_len_t)
	linspace_border = locality == _cabi.VX_GLOBAL and r["gap_kind"] == 1
	n_flip = 0
	for case in range(45):
		len_s, len_t = int(rng.integers(1, max_len_s + 1)), int(rng.integers(17, max_len_t + 1))
		if case % 3 == 1:
			sim = (rng.integers(0, 4, size=(len_s, len_t)) / 3.0).astype(np.float32)      # many exact ties
		else:
			sim = np.clip(rng.standard_normal((len_s, len_t)) * 0.5 + 0.2, 0, 1).astype(np.float32)
		s_ref, t_ref = solver.align(sim)
		r2 = dict(gaps.resolve_gap_costs(gap, len_s, len_t))
		kind = r2.pop("gap_kind")
		ap, keep = _cabi.make_align_params(locality, kind, **r2)
		s_or, t_or = oracle.align(sim, ap)
		if linspace_border:
			assert abs(float(s_or) - float(s_ref)) <= 5e-7 * max(1.0, abs(float(s_ref))), (case, s_or, s_ref)
			n_flip += list(t_or) != list(t_ref)
		else:
			assert np.float32(s_or).tobytes() == np.float32(s_ref).tobytes(), (case, len_s, len_t, s_or, s_ref)
			assert list(t_or) == list(t_ref), (case, len_s, len_t, list(t_or), list(t_ref))
	solver.close()
	assert n_flip <= 4


@pytest.mark.parametrize("gap_name", ["smooth5", "mixed", "log"])
@pytest.mark.parametrize("locality", LOCALITIES)
def test_oracle_equals_reference_solver_on_long_windows(oracle, locality, gap_name):
	"""Windows of up to 320 tokens (several sentences) with general gap costs: the lengths of the long WSB length classes and
	of the scratch kernel, which the GPU tests check against the oracle."""
	rng = np.random.default_rng(abs(hash((locality, gap_name, "windows"))) % (2 ** 31))
	gap = GAPS[gap_name]
	max_len_s, max_len_t = 320, 16
	r = gaps.resolve_gap_costs(gap, max_len_s, max_len_t)
	solver = ref.RefSolver(locality, r, max_len_s, max_len_t)
	for case in range(16):
		len_s, len_t = int(rng.integers(65, max_len_s + 1)), int(rng.integers(1, max_len_t + 1))
		if case % 3 == 1:
			sim = (rng.integers(0, 4, size=(len_s, len_t)) / 3.0).astype(np.float32)
		else:
			sim = np.clip(rng.standard_normal((len_s, len_t)) * 0.5 + 0.2, 0, 1).astype(np.float32)
		s_ref, t_ref = solver.align(sim)
		r2 = dict(gaps.resolve_gap_costs(gap, len_s, len_t))
		kind = r2.pop("gap_kind")
		ap, keep = _cabi.make_align_params(locality, kind, **r2)
		s_or, t_or = oracle.align(sim, ap)
		assert np.float32(s_or).tobytes() == np.float32(s_ref).tobytes(), (case, len_s, len_t, s_or, s_ref)
		assert list(t_or) == list(t_ref), (case, len_s, len_t, list(t_or), list(t_ref))
	solver.close()


# ---- heap order pinned to the reference's own ResultSet (oracle/ref/resultset_driver.cpp) ----

def tie_problem(n_doc=40, per_doc=37, seed=0):
	"""One-token windows and a one-token query: the score of a window is S[its token id, 0], i.e. any value we like.
	Every document carries the SAME per_doc distinct scores (in a shuffled slice order): every score is tied n_doc times
	ACROSS documents and never inside one -- the case Match::compare_by_score orders (score, then document id); ties
	inside a document fall through to a comparison of flow POINTERS in the reference (match_impl.h:20-24) and are exempt."""
	rng = np.random.default_rng(seed)
	base = (rng.permutation(per_doc).astype(np.float32) + 1) / np.float32(per_doc + 3)
	n = n_doc * per_doc
	token_id = np.concatenate([rng.permutation(per_doc) for _ in range(n_doc)]).astype(np.int32)
	doc_span_off = np.arange(0, n + 1, per_doc, dtype=np.int64)
	span_start = (np.arange(n) % per_doc).astype(np.int32)
	corpus = dict(token_id=token_id, doc_tok_off=doc_span_off.copy(), span_start=span_start, span_end=span_start + 1,
		doc_span_off=doc_span_off)
	s = base.reshape(per_doc, 1).copy()
	raw = base[token_id]
	doc = np.repeat(np.arange(n_doc, dtype=np.int32), per_doc)
	return corpus, s, raw, doc, span_start.copy()


@pytest.mark.skipif(not ref.result_set_available(), reason="oracle/_ref/libresultset_ref.so not built")
@pytest.mark.parametrize("k,min_score", [(100, 0.0), (7, 0.0), (40, 0.5), (500, 0.0), (2000, 0.9), (1, 0.0)])
def test_oracle_result_set_equals_reference_result_set(oracle, k, min_score):
	from vectorian_b200 import _cabi
	corpus, s, raw, doc, slice_id = tie_problem()
	want_doc, want_slice, want_score = ref.result_set(raw, np.ones_like(raw), doc, slice_id, k, min_score)
	oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
	ap, _ = _cabi.make_align_params(_cabi.VX_LOCAL, _cabi.VX_GAP_LINEAR, lin=(0.0, 0.0))
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=min_score)
	hits, _, _ = oc.search(s, ap, qp, n_threads=3)
	assert len(hits) == len(want_doc) > 0
	np.testing.assert_array_equal(hits["score"], want_score)
	np.testing.assert_array_equal(hits["doc"], want_doc)          # equal scores: the HIGHER document id first
	np.testing.assert_array_equal(hits["slice"], want_slice)
	assert k == 1 or np.sum(np.diff(want_score) == 0) > 1       # the winners really are full of ties


# ---- slices pinned to the reference's own slice/static.h (oracle/ref/slice_driver.cpp) ----

needs_slices = pytest.mark.skipif(not ref.slices_available(), reason="oracle/_ref/libslice_ref.so not built")


@needs_slices
@pytest.mark.parametrize("weights", ["dyadic", "fractions", "none"])
def test_oracle_window_similarity_equals_reference_slices(oracle, weights):
	"""The similarity matrix the oracle's DP sees for a window (fill_sim) against the reference's StaticEmbeddingSlice /
	TagWeightedSlice::similarity (cpp/slice/static.h:71-75, 236-263): query-token weight, POS-mismatch penalty, threshold --
	bit for bit; `unmodified_similarity` stays the plain S rows; max_sum_of_similarities is the float sum of the weights."""
	from tests import helpers
	corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=200, t=9, seed=5, with_pos=True)
	s = oracle.sim_static(emb, q_id, q_vec)
	t = s.shape[1]
	rng = np.random.default_rng(11)
	t_pos = rng.integers(0, 17, size=t).astype(np.int8)
	w = {"dyadic": (rng.integers(1, 9, size=t) / 4.0).astype(np.float32),
		"fractions": (0.3 + 0.1 * np.arange(t)).astype(np.float32), "none": None}[weights]
	oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
		corpus["doc_span_off"], token_pos=corpus["token_pos"])
	if w is None:
		qp, keep = _cabi.make_query_params(max_matches=10, min_score=0.0)
	else:
		qp, keep = _cabi.make_query_params(max_matches=10, min_score=0.0, t_weights=w, t_pos=t_pos,
			pos_mismatch_penalty=0.35, similarity_threshold=0.15)
	doc, sl, tok, ln = oc.windows()
	n_zeroed = 0
	for i in rng.choice(len(tok), size=40, replace=False):
		ids = corpus["token_id"][tok[i]:tok[i] + ln[i]]
		pos = corpus["token_pos"][tok[i]:tok[i] + ln[i]]
		o_sim = oc.window_similarity(s, qp, tok[i], ln[i])
		r_sim, r_unmod, r_max_sum = ref.slice_similarity(s, ids, pos, t_pos, w, 0.35, 0.15)
		assert o_sim.tobytes() == r_sim.tobytes()
		assert r_unmod.tobytes() == s[ids].tobytes()
		if w is not None:
			acc = np.float32(0.0)
			for x in w:
				acc = np.float32(acc + x)
			assert np.float32(r_max_sum) == acc
			n_zeroed += int(np.sum((r_sim == 0.0) & (s[ids] > 0.0)))
		else:
			assert r_max_sum == float(t)
	assert w is None or n_zeroed > 0   # the threshold does cut something in this problem


@needs_slices
def test_filter_pass_equals_reference_filtered_slice(oracle):
	"""Session.filter_pass + the compaction of the token columns (vectorian_b200/session.py) against the reference's
	FilteredSliceFactory::create_slice (cpp/slice/static.h:365-416, TokenFilter::pass cpp/query.h:22-28): the rows of the
	FilteredSlice are the S rows of exactly the tokens the product keeps, in order -- also when nothing or everything passes."""
	from vectorian_b200.session import Session
	rng = np.random.default_rng(3)
	s = rng.random((50, 6)).astype(np.float32)
	for case in range(30):
		n = int(rng.integers(1, 40))
		ids = rng.integers(0, 50, size=n).astype(np.int32)
		pos = rng.integers(0, 20, size=n).astype(np.int8)
		tag = rng.integers(0, 45, size=n).astype(np.int8)
		pos_mask = int(rng.integers(0, 2 ** 20)) if case % 5 else 0
		tag_mask = int(rng.integers(0, 2 ** 45)) if case % 3 else 0
		if case == 7:
			pos_mask = 2 ** 20 - 1    # nothing passes
		keep = Session.filter_pass(pos, tag, (pos_mask, tag_mask))
		r_sim = ref.filtered_similarity(s, ids, pos, tag, pos_mask, tag_mask)
		assert r_sim.shape[0] == int(keep.sum())
		assert r_sim.tobytes() == s[ids[keep]].tobytes()


# ---- window enumeration and booster pinned to the reference's own document.h (oracle/ref/document_driver.cpp) ----

needs_document = pytest.mark.skipif(not ref.document_available(), reason="oracle/_ref/libdocument_ref.so not built")


@needs_document
def test_oracle_window_enumeration_equals_reference_spans_iterate(oracle):
	"""vxo_enumerate_windows against the reference's Spans::iterate over VariableSpans (cpp/document.h:58-104, 147-169):
	slice ids, token offsets and lengths of every window for a range of window sizes / steps, with empty spans (skipped
	when the whole window is empty), a single span, steps larger than the window; max_len as the reference computes it."""
	rng = np.random.default_rng(21)
	for case in range(40):
		n = int(rng.integers(1, 60))
		lens = rng.integers(0 if case % 2 else 1, 12, size=n)
		ends = np.cumsum(lens)
		starts = ends - lens
		n_tok = int(ends[-1])
		for window_size, window_step in ((1, 1), (2, 1), (3, 2), (5, 3), (2, 5), (7, 7)):
			r_slice, r_at, r_len = ref.iterate_windows(starts, ends, window_size, window_step)
			oc = oracle.OracleCorpus(np.zeros(max(n_tok, 1), np.int32), [0, n_tok], starts, ends, [0, n], window_size, window_step)
			o_doc, o_slice, o_tok, o_len = oc.windows()
			np.testing.assert_array_equal(o_slice, r_slice)
			np.testing.assert_array_equal(o_tok, r_at)
			np.testing.assert_array_equal(o_len, r_len)
			assert np.all(o_doc == 0)
			if len(r_len):
				assert int(r_len.max()) <= ref.spans_max_len(starts, ends, window_size)


@needs_document
def test_boost_lookup_equals_reference_booster():
	"""Booster::get_boost (cpp/document.h:175-187) is a plain lookup by slice id -- what vx_corpus_set_boost stores."""
	w = np.random.default_rng(2).uniform(0.25, 2.0, size=37).astype(np.float32)
	for i in (0, 1, 17, 36):
		assert np.float32(ref.boost(w, i)) == w[i]


@needs_document
def test_fixed_spans_iterate_like_unit_variable_spans():
	"""The reference's token-level partition (FixedSpans, cpp/document.h:26-55) enumerates the windows that one-token
	VariableSpans do -- which is how the host API hands a token-level partition to vx_corpus_create."""
	for n in (1, 2, 9, 40):
		unit = np.arange(n, dtype=np.int32)
		for window_size, window_step in ((1, 1), (3, 1), (4, 2), (2, 3)):
			a = ref.iterate_windows(None, None, window_size, window_step, n_fixed=n)
			b = ref.iterate_windows(unit, unit + 1, window_size, window_step)
			for x, y in zip(a, b):
				np.testing.assert_array_equal(x, y)
