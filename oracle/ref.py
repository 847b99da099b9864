"""ctypes binding of oracle/_ref/libpyalign_ref.so: the reference's own, unmodified pyalign solver.h compiled
against the xtensor stand-in of oracle/ref/shim (oracle/ref/ref_driver.cpp).  TEST INFRASTRUCTURE ONLY: a second,
independent check of the oracle's DP and traceback (tests/test_oracle_vs_ref.py).  The library is built in the
development container from /root/reference (oracle/ref/Makefile) and is not part of the product path."""

import ctypes as C
import os

import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libpyalign_ref.so")

_lib = None


def available():
	return os.path.exists(LIB_PATH)


def lib():
	global _lib
	if _lib is None:
		l = C.CDLL(LIB_PATH)
		l.ref_last_error.restype = C.c_char_p
		l.ref_solver_create.restype = C.c_void_p
		l.ref_solver_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
		l.ref_solver_run.restype = C.c_int
		l.ref_solver_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
		l.ref_solver_gap_cost.restype = C.c_float
		l.ref_solver_gap_cost.argtypes = [C.c_void_p, C.c_int, C.c_int]
		l.ref_solver_destroy.argtypes = [C.c_void_p]
		l.ref_corpus_scores.restype = C.c_int
		l.ref_corpus_scores.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
			C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p]
		_lib = l
	return _lib


class RefSolver:
	"""One reference solver instance (pyalign LinearGapCostSolver / AffineGapCostSolver / GeneralGapCostSolver for
	Local / Global / Semiglobal), reused over many sentences like Vectorian reuses it
	(vectorian/core/cpp/metric/alignment.h:241-245)."""

	def __init__(self, locality, resolved, max_len_s, max_len_t):
		"""`resolved` = vectorian_b200.gaps.resolve_gap_costs(gap, max_len_s, max_len_t)."""
		kind = resolved["gap_kind"]
		p = np.zeros((4,), dtype=np.float32)
		w_s = w_t = None
		if kind == 0:
			p[:2] = resolved["lin"]
		elif kind == 1:
			# pyalign.gaps: w(k) = u + v k; AffineCost(u, v): w(k) = u k + v -- flipped like factory.h:41-52
			(u_s, v_s), (u_t, v_t) = resolved["affine_s"], resolved["affine_t"]
			p[:] = (v_s, u_s, v_t, u_t)
		else:
			w_s = np.ascontiguousarray(resolved["w_s"], dtype=np.float32)
			w_t = np.ascontiguousarray(resolved["w_t"], dtype=np.float32)
			assert len(w_s) >= max_len_s + 1 and len(w_t) >= max_len_t + 1
		self._keep = (p, w_s, w_t)
		self._lib = lib()
		self._h = self._lib.ref_solver_create(int(locality), int(kind), int(max_len_s), int(max_len_t), p.ctypes.data,
			w_s.ctypes.data if w_s is not None else None, w_t.ctypes.data if w_t is not None else None)
		if not self._h:
			raise RuntimeError(self._lib.ref_last_error().decode())

	def align(self, sim):
		"""sim[len_s x len_t] -> (score, target[len_t] int16)."""
		sim = np.ascontiguousarray(sim, dtype=np.float32)
		len_s, len_t = sim.shape
		score = C.c_float(0)
		target = np.full((len_t,), -1, dtype=np.int16)
		if self._lib.ref_solver_run(self._h, sim.ctypes.data, len_s, len_t, C.addressof(score), target.ctypes.data) != 0:
			raise RuntimeError(self._lib.ref_last_error().decode())
		return np.float32(score.value), target

	def gap_cost(self, which, n):
		return np.float32(self._lib.ref_solver_gap_cost(self._h, 0 if which == "s" else 1, int(n)))

	def close(self):
		if self._h:
			self._lib.ref_solver_destroy(self._h)
			self._h = None

	def __del__(self):
		self.close()


def _pack(resolved):
	kind = resolved["gap_kind"]
	p = np.zeros((4,), dtype=np.float32)
	w_s = w_t = None
	if kind == 0:
		p[:2] = resolved["lin"]
	elif kind == 1:
		(u_s, v_s), (u_t, v_t) = resolved["affine_s"], resolved["affine_t"]
		p[:] = (v_s, u_s, v_t, u_t)
	else:
		w_s = np.ascontiguousarray(resolved["w_s"], dtype=np.float32)
		w_t = np.ascontiguousarray(resolved["w_t"], dtype=np.float32)
	return kind, p, w_s, w_t


def corpus_scores(locality, resolved, s, token_id, sent_off, max_len_s, n_threads=1, doc_len=1000):
	"""Raw alignment score of every sentence through the reference solver, `n_threads` workers over documents of
	`doc_len` sentences.  s[V x T] float32, token_id int32, sent_off int64[n_sent + 1]."""
	kind, p, w_s, w_t = _pack(resolved)
	s = np.ascontiguousarray(s, dtype=np.float32)
	token_id = np.ascontiguousarray(token_id, dtype=np.int32)
	sent_off = np.ascontiguousarray(sent_off, dtype=np.int64)
	n_sent = len(sent_off) - 1
	out = np.zeros((n_sent,), dtype=np.float32)
	rc = lib().ref_corpus_scores(int(locality), int(kind), int(max_len_s), int(s.shape[1]), p.ctypes.data,
		w_s.ctypes.data if w_s is not None else None, w_t.ctypes.data if w_t is not None else None,
		s.ctypes.data, token_id.ctypes.data, sent_off.ctypes.data, n_sent, int(doc_len), int(n_threads), out.ctypes.data)
	if rc != 0:
		raise RuntimeError("reference solver failed: " + lib().ref_last_error().decode())
	return out


# ---- the reference's own ResultSet (oracle/ref/resultset_driver.cpp) ----

RS_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libresultset_ref.so")
_rs_lib = None


def result_set_available():
	return os.path.exists(RS_LIB_PATH)


def result_set(raw, score_max, doc, slice_id, max_matches, min_score, boost=None):
	"""Windows (document-major, slice ascending) through the reference's unmodified result_set.h / result_set.cpp /
	match_impl.h the way _find drives them: a ResultSet per document (`score > worst_score()` admission), extend() in
	document order, best_n.  Returns (doc, slice, score) arrays of the winners, best first."""
	global _rs_lib
	if _rs_lib is None:
		_rs_lib = C.CDLL(RS_LIB_PATH)
		_rs_lib.ref_result_set.restype = C.c_int
		_rs_lib.ref_result_set.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
			C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
	raw = np.ascontiguousarray(raw, dtype=np.float32)
	score_max = np.ascontiguousarray(score_max, dtype=np.float32)
	doc = np.ascontiguousarray(doc, dtype=np.int32)
	slice_id = np.ascontiguousarray(slice_id, dtype=np.int32)
	b = None if boost is None else np.ascontiguousarray(boost, dtype=np.float32)
	k = int(max_matches)
	o_doc = np.zeros((k,), np.int32)
	o_slice = np.zeros((k,), np.int32)
	o_score = np.zeros((k,), np.float32)
	n = C.c_int32(0)
	_rs_lib.ref_result_set(raw.ctypes.data, score_max.ctypes.data, None if b is None else b.ctypes.data, doc.ctypes.data,
		slice_id.ctypes.data, len(raw), k, float(min_score), o_doc.ctypes.data, o_slice.ctypes.data, o_score.ctypes.data, C.byref(n))
	return o_doc[:n.value], o_slice[:n.value], o_score[:n.value]


# ---- the reference's own slices (oracle/ref/slice_driver.cpp) ----

SLICE_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libslice_ref.so")
_slice_lib = None


def slices_available():
	return os.path.exists(SLICE_LIB_PATH)


def _slices():
	global _slice_lib
	if _slice_lib is None:
		_slice_lib = C.CDLL(SLICE_LIB_PATH)
		_slice_lib.ref_slice_similarity.restype = C.c_int
		_slice_lib.ref_slice_similarity.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
			C.c_int32, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
		_slice_lib.ref_filtered_similarity.restype = C.c_int
		_slice_lib.ref_filtered_similarity.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
			C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
	return _slice_lib


def slice_similarity(s, s_id, s_pos, t_pos, t_weights=None, pos_mismatch_penalty=0.0, similarity_threshold=0.0):
	"""similarity(i, j), unmodified_similarity(i, j) and max_sum_of_similarities() of the reference's StaticEmbeddingSlice
	(t_weights None) or TagWeightedSlice<StaticEmbeddingSlice> (cpp/slice/static.h:9-100, 193-288) over a window with token ids
	s_id / POS ids s_pos and a query with POS ids t_pos."""
	s = np.ascontiguousarray(s, dtype=np.float32)
	s_id = np.ascontiguousarray(s_id, dtype=np.int32)
	s_pos = np.ascontiguousarray(s_pos, dtype=np.int8)
	t_pos = np.ascontiguousarray(t_pos, dtype=np.int8)
	t = s.shape[1]
	sim = np.zeros((len(s_id), t), np.float32)
	unmod = np.zeros((len(s_id), t), np.float32)
	max_sum = C.c_float(0.0)
	w = None if t_weights is None else np.ascontiguousarray(t_weights, dtype=np.float32)
	_slices().ref_slice_similarity(s.ctypes.data, s.shape[0], t, s_id.ctypes.data, s_pos.ctypes.data, len(s_id), t_pos.ctypes.data,
		0 if w is None else 1, None if w is None else w.ctypes.data, float(pos_mismatch_penalty), float(similarity_threshold),
		sim.ctypes.data, unmod.ctypes.data, C.byref(max_sum))
	return sim, unmod, float(max_sum.value)


def filtered_similarity(s, s_id, s_pos, s_tag, pos_mask, tag_mask):
	"""similarity(i, j) of the FilteredSlice the reference's FilteredSliceFactory::create_slice builds (cpp/slice/static.h:365-416;
	TokenFilter::pass, cpp/query.h:22-28): rows = the window's tokens that pass the filter, in order."""
	s = np.ascontiguousarray(s, dtype=np.float32)
	s_id = np.ascontiguousarray(s_id, dtype=np.int32)
	s_pos = np.ascontiguousarray(s_pos, dtype=np.int8)
	s_tag = np.ascontiguousarray(s_tag, dtype=np.int8)
	t = s.shape[1]
	sim = np.zeros((len(s_id), t), np.float32)
	n = C.c_int32(0)
	_slices().ref_filtered_similarity(s.ctypes.data, s.shape[0], t, s_id.ctypes.data, s_pos.ctypes.data, s_tag.ctypes.data, len(s_id),
		int(pos_mask), int(tag_mask), sim.ctypes.data, C.byref(n))
	return sim[:n.value]


# ---- the reference's own span iteration and booster (oracle/ref/document_driver.cpp) ----

DOC_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libdocument_ref.so")
_doc_lib = None


def document_available():
	return os.path.exists(DOC_LIB_PATH)


def _document():
	global _doc_lib
	if _doc_lib is None:
		_doc_lib = C.CDLL(DOC_LIB_PATH)
		_doc_lib.ref_iterate_windows.restype = C.c_int64
		_doc_lib.ref_iterate_windows.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
		_doc_lib.ref_max_len.restype = C.c_int32
		_doc_lib.ref_max_len.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32]
		_doc_lib.ref_boost.restype = C.c_float
		_doc_lib.ref_boost.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
	return _doc_lib


def iterate_windows(span_start, span_end, window_size, window_step, n_fixed=None):
	"""(slice_id, token_at, len_s) of every callback of the reference's Spans::iterate (cpp/document.h:147-169) over
	VariableSpans built from (span_start, span_end), or over FixedSpans(n_fixed) when the spans are None."""
	if span_start is None:
		n, ps, pe = int(n_fixed), None, None
	else:
		span_start = np.ascontiguousarray(span_start, dtype=np.int32)
		span_end = np.ascontiguousarray(span_end, dtype=np.int32)
		n, ps, pe = len(span_start), span_start.ctypes.data, span_end.ctypes.data
	cap = n + 1
	sl = np.zeros((cap,), np.int32)
	at = np.zeros((cap,), np.int64)
	ln = np.zeros((cap,), np.int32)
	k = _document().ref_iterate_windows(ps, pe, n, int(window_size), int(window_step), sl.ctypes.data, at.ctypes.data, ln.ctypes.data, cap)
	assert k <= cap
	return sl[:k], at[:k], ln[:k]


def spans_max_len(span_start, span_end, window_size):
	span_start = np.ascontiguousarray(span_start, dtype=np.int32)
	span_end = np.ascontiguousarray(span_end, dtype=np.int32)
	return int(_document().ref_max_len(span_start.ctypes.data, span_end.ctypes.data, len(span_start), int(window_size)))


def boost(weights, slice_id):
	w = np.ascontiguousarray(weights, dtype=np.float32)
	return float(_document().ref_boost(w.ctypes.data, len(w), int(slice_id)))


# ---- the reference's own result rendering (oracle/ref/flow_driver.cpp) ----

FLOW_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libflow_ref.so")
_flow_lib = None


def flow_available():
	return os.path.exists(FLOW_LIB_PATH)


def regions(s_idx, s_len, s_pos, s_tag, t_idx, t_len, t_pos, token_at, slice_len, target, dist, window_size,
		token_filter=None, gap_s=None, gap_t=None):
	"""Flow::py_regions / py_omitted of the reference (cpp/match/flow.cpp:8-188 with match.h, region.h) for an injective
	alignment target[j] (window position of query token j or -1) with distances dist[j], over document tokens given by
	character offset / length / POS id / tag id and query tokens likewise.  gap_s / gap_t: gap cost by gap length
	(Matcher::gap_cost_s / gap_cost_t).  Returns (regions, omitted) in the shape of oracle/render.py: regions = dicts with
	's' (start, end), 'gap_penalty', 'matched' and, if matched, 'pos_s' (POS id) and 'edges' (dicts with 't_slice',
	't_index', 't_pos' (POS id), 'flow', 'distance'); omitted = [(start, end)]."""
	global _flow_lib
	if _flow_lib is None:
		_flow_lib = C.CDLL(FLOW_LIB_PATH)
		_flow_lib.ref_regions.restype = C.c_int64
		_flow_lib.ref_regions.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
			C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
			C.c_int32, C.c_int32, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]
	a = lambda x, dt: np.ascontiguousarray(x, dtype=dt)
	s_idx, s_len, s_pos, s_tag = a(s_idx, np.int32), a(s_len, np.int8), a(s_pos, np.int8), a(s_tag, np.int8)
	t_idx, t_len, t_pos = a(t_idx, np.int32), a(t_len, np.int8), a(t_pos, np.int8)
	target, dist = a(target, np.int16), a(dist, np.float32)
	gap_s = a(gap_s if gap_s is not None else [0.0], np.float32)
	gap_t = a(gap_t if gap_t is not None else [0.0], np.float32)
	cap = 64 + 16 * (len(s_idx) + len(t_idx))
	out = np.zeros((cap,), np.float64)
	pm, tm = (0, 0) if token_filter is None else token_filter
	n = _flow_lib.ref_regions(s_idx.ctypes.data, s_len.ctypes.data, s_pos.ctypes.data, s_tag.ctypes.data, len(s_idx),
		t_idx.ctypes.data, t_len.ctypes.data, t_pos.ctypes.data, len(t_idx), int(token_at), int(slice_len),
		target.ctypes.data, dist.ctypes.data, int(window_size), 0 if token_filter is None else 1, int(pm), int(tm),
		gap_s.ctypes.data, len(gap_s), gap_t.ctypes.data, len(gap_t), out.ctypes.data, cap)
	assert n >= 0
	it = iter(out[:n])
	nxt = lambda: next(it)
	result = []
	for _ in range(int(nxt())):
		kind, s0, s1, pen = int(nxt()), int(nxt()), int(nxt()), float(np.float32(nxt()))
		r = {'s': (s0, s1), 'gap_penalty': pen, 'matched': kind == 1}
		if kind == 1:
			r['pos_s'] = int(nxt())
			edges = []
			for _ in range(int(nxt())):
				ti, a0, a1, tp, fl, di = int(nxt()), int(nxt()), int(nxt()), int(nxt()), float(np.float32(nxt())), float(np.float32(nxt()))
				edges.append({'t_slice': (a0, a1), 't_index': ti, 't_pos': tp, 'flow': fl, 'distance': di})
			r['edges'] = edges
		result.append(r)
	omitted = [(int(nxt()), int(nxt())) for _ in range(int(nxt()))]
	return result, omitted
