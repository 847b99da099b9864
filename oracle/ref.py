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
