"""ctypes binding of oracle/libvx_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It shares the parameter structs of the product's C ABI
(vectorian_b200/_cabi.py) so that a test drives both with the same arguments.
"""

import ctypes as C
import os
import subprocess

import numpy as np

from vectorian_b200 import _cabi
from vectorian_b200._cabi import AlignParams, QueryParams, Hit, HIT_DTYPE, ptr

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvx_oracle.so")


class Problem(C.Structure):
	"""vxo_problem"""
	_fields_ = [
		("token_id", _cabi.c_int32_p),
		("token_pos", _cabi.c_int8_p),
		("n_tok", C.c_int64),
		("doc_tok_off", _cabi.c_int64_p),
		("n_doc", C.c_int32),
		("span_start", _cabi.c_int32_p),
		("span_end", _cabi.c_int32_p),
		("doc_span_off", _cabi.c_int64_p),
		("window_size", C.c_int32),
		("window_step", C.c_int32),
		("boost", _cabi.c_float_p),
		("s", _cabi.c_float_p),
		("s_rows", C.c_int64),
		("contextual", C.c_int32),
		("t", C.c_int32),
	]


_lib = None


def build():
	subprocess.check_call(["make", "-s", "-C", _HERE, "libvx_oracle.so"])


def lib():
	global _lib
	if _lib is None:
		if not os.path.exists(LIB_PATH):
			build()
		l = C.CDLL(LIB_PATH)
		l.vxo_normalize_rows.restype = None
		l.vxo_normalize_rows.argtypes = [_cabi.c_float_p, C.c_int64, C.c_int32, _cabi.c_float_p]
		l.vxo_sim_static.restype = None
		l.vxo_sim_static.argtypes = [
			_cabi.c_float_p, C.c_int64, C.c_int32, _cabi.c_int32_p, _cabi.c_float_p, C.c_int32, _cabi.c_float_p]
		l.vxo_sim_contextual.restype = None
		l.vxo_sim_contextual.argtypes = [
			_cabi.c_float_p, C.c_int64, C.c_int32, _cabi.c_float_p, C.c_int32, _cabi.c_float_p]
		l.vxo_align.restype = C.c_float
		l.vxo_align.argtypes = [_cabi.c_float_p, C.c_int32, C.c_int32, C.POINTER(AlignParams), _cabi.c_int16_p]
		l.vxo_enumerate_windows.restype = C.c_int64
		l.vxo_enumerate_windows.argtypes = [
			_cabi.c_int64_p, C.c_int32, _cabi.c_int32_p, _cabi.c_int32_p, _cabi.c_int64_p,
			C.c_int32, C.c_int32, _cabi.c_int32_p, _cabi.c_int32_p, _cabi.c_int64_p, _cabi.c_int32_p]
		l.vxo_search.restype = C.c_int
		l.vxo_search.argtypes = [
			C.POINTER(Problem), C.POINTER(AlignParams), C.POINTER(QueryParams), C.c_int32,
			C.POINTER(Hit), _cabi.c_int16_p, _cabi.c_float_p, _cabi.c_int32_p]
		l.vxo_score_all.restype = C.c_int
		l.vxo_score_all.argtypes = [
			C.POINTER(Problem), C.POINTER(AlignParams), C.POINTER(QueryParams), C.c_int32,
			_cabi.c_float_p, C.c_int64]
		_lib = l
	return _lib


def normalize_rows(e):
	e = np.ascontiguousarray(e, dtype=np.float32)
	out = np.empty_like(e)
	lib().vxo_normalize_rows(ptr(e, C.c_float), e.shape[0], e.shape[1], ptr(out, C.c_float))
	return out


def sim_static(e, q_id, q_vec):
	"""S[V x T] from UNnormalised E[V x d] and query vectors [T x d]."""
	ehat = normalize_rows(e)
	q_id = np.ascontiguousarray(q_id, dtype=np.int32)
	q_vec = np.ascontiguousarray(q_vec, dtype=np.float32)
	t = len(q_id)
	s = np.empty((ehat.shape[0], t), dtype=np.float32)
	lib().vxo_sim_static(ptr(ehat, C.c_float), ehat.shape[0], ehat.shape[1],
		ptr(q_id, C.c_int32), ptr(q_vec, C.c_float), t, ptr(s, C.c_float))
	return s


def sim_contextual(x, q_vec):
	x = np.ascontiguousarray(x, dtype=np.float32)
	q_vec = np.ascontiguousarray(q_vec, dtype=np.float32)
	t = q_vec.shape[0]
	s = np.empty((x.shape[0], t), dtype=np.float32)
	lib().vxo_sim_contextual(ptr(x, C.c_float), x.shape[0], x.shape[1], ptr(q_vec, C.c_float), t, ptr(s, C.c_float))
	return s


def align(sim, ap):
	"""One solve + traceback.  Returns (raw score, target[len_t])."""
	sim = np.ascontiguousarray(sim, dtype=np.float32)
	len_s, len_t = sim.shape
	target = np.full((len_t,), -1, dtype=np.int16)
	score = lib().vxo_align(ptr(sim, C.c_float), len_s, len_t, C.byref(ap), ptr(target, C.c_int16))
	return float(np.float32(score)), target


def edges_from_target(target):
	"""[[u, v], ...] sorted by v -- the form pyalign's tests compare (tests/__init__.py:13-24)."""
	return [[int(u), int(v)] for v, u in enumerate(target) if u >= 0]


class OracleCorpus:
	"""Holds the reference-shaped corpus arrays (numpy) for the oracle."""

	def __init__(self, token_id, doc_tok_off, span_start, span_end, doc_span_off,
			window_size=1, window_step=1, token_pos=None, boost=None):
		self.token_id = np.ascontiguousarray(token_id, dtype=np.int32)
		self.token_pos = None if token_pos is None else np.ascontiguousarray(token_pos, dtype=np.int8)
		self.doc_tok_off = np.ascontiguousarray(doc_tok_off, dtype=np.int64)
		self.span_start = np.ascontiguousarray(span_start, dtype=np.int32)
		self.span_end = np.ascontiguousarray(span_end, dtype=np.int32)
		self.doc_span_off = np.ascontiguousarray(doc_span_off, dtype=np.int64)
		self.window_size = int(window_size)
		self.window_step = int(window_step)
		self.boost = None if boost is None else np.ascontiguousarray(boost, dtype=np.float32)
		self.n_doc = len(self.doc_tok_off) - 1

	def windows(self):
		args = (ptr(self.doc_tok_off, C.c_int64), self.n_doc, ptr(self.span_start, C.c_int32),
			ptr(self.span_end, C.c_int32), ptr(self.doc_span_off, C.c_int64), self.window_size, self.window_step)
		null32 = C.cast(None, _cabi.c_int32_p)
		n = lib().vxo_enumerate_windows(*args, null32, null32, C.cast(None, _cabi.c_int64_p), null32)
		doc = np.empty((n,), dtype=np.int32)
		sl = np.empty((n,), dtype=np.int32)
		tok = np.empty((n,), dtype=np.int64)
		ln = np.empty((n,), dtype=np.int32)
		lib().vxo_enumerate_windows(*args, ptr(doc, C.c_int32), ptr(sl, C.c_int32), ptr(tok, C.c_int64), ptr(ln, C.c_int32))
		return doc, sl, tok, ln

	def problem(self, s, contextual=False):
		s = np.ascontiguousarray(s, dtype=np.float32)
		p = Problem()
		p.token_id = ptr(self.token_id, C.c_int32)
		p.token_pos = ptr(self.token_pos, C.c_int8)
		p.n_tok = len(self.token_id)
		p.doc_tok_off = ptr(self.doc_tok_off, C.c_int64)
		p.n_doc = self.n_doc
		p.span_start = ptr(self.span_start, C.c_int32)
		p.span_end = ptr(self.span_end, C.c_int32)
		p.doc_span_off = ptr(self.doc_span_off, C.c_int64)
		p.window_size = self.window_size
		p.window_step = self.window_step
		p.boost = ptr(self.boost, C.c_float)
		p.s = ptr(s, C.c_float)
		p.s_rows = s.shape[0]
		p.contextual = 1 if contextual else 0
		p.t = s.shape[1]
		return p, s

	def search(self, s, ap, qp, n_threads=1, contextual=False):
		"""Returns (hits structured array, target[n x T], dist[n x T])."""
		p, s_keep = self.problem(s, contextual)
		k = max(int(qp.max_matches), 1)
		t = p.t
		hits = np.zeros((k,), dtype=HIT_DTYPE)
		target = np.full((k, max(t, 1)), -1, dtype=np.int16)
		dist = np.ones((k, max(t, 1)), dtype=np.float32)
		n = C.c_int32(0)
		rc = lib().vxo_search(C.byref(p), C.byref(ap), C.byref(qp), n_threads,
			hits.ctypes.data_as(C.POINTER(Hit)), ptr(target, C.c_int16), ptr(dist, C.c_float), C.byref(n))
		assert rc == 0
		return hits[:n.value], target[:n.value, :t], dist[:n.value, :t]

	def window_similarity(self, s, qp, tok, len_s, contextual=False):
		"""The [len_s x T] similarity matrix the DP sees for the window starting at token `tok` (test hook)."""
		p, s_keep = self.problem(s, contextual)
		out = np.zeros((int(len_s), p.t), dtype=np.float32)
		f = lib().vxo_debug_window_similarity
		f.restype = None
		f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
		f(C.byref(p), C.byref(qp), int(tok), int(len_s), out.ctypes.data)
		return out

	def score_all(self, s, ap, qp, n_threads=1, contextual=False):
		p, s_keep = self.problem(s, contextual)
		n = len(self.windows()[0])
		out = np.zeros((n,), dtype=np.float32)
		rc = lib().vxo_score_all(C.byref(p), C.byref(ap), C.byref(qp), n_threads, ptr(out, C.c_float), n)
		assert rc == 0
		return out
