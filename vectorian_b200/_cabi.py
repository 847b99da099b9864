"""ctypes binding of include/vectorian_b200.h (the C-ABI shared library).

This is the only way Python reaches the CUDA path.  There is no CPU fallback:
if ``libvectorian_b200.so`` is missing or no B200 is visible, calls raise.
Replaces the pybind11 module ``vectorian_core`` (reference:
vectorian/core/cpp/module.cpp:30-152).
"""

import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvectorian_b200.so")

VX_LOCAL, VX_GLOBAL, VX_SEMIGLOBAL = 0, 1, 2
VX_GAP_LINEAR, VX_GAP_AFFINE, VX_GAP_GENERAL = 0, 1, 2
VX_MAX_LEN = 16383
VX_NOALLOC_MAX_K = 1024

c_float_p = C.POINTER(C.c_float)
c_int8_p = C.POINTER(C.c_int8)
c_int16_p = C.POINTER(C.c_int16)
c_int32_p = C.POINTER(C.c_int32)
c_int64_p = C.POINTER(C.c_int64)


class AlignParams(C.Structure):
	"""vx_align_params"""
	_fields_ = [
		("locality", C.c_int32),
		("gap_kind", C.c_int32),
		("lin_s", C.c_float), ("lin_t", C.c_float),
		("aff_open_s", C.c_float), ("aff_ext_s", C.c_float),
		("aff_open_t", C.c_float), ("aff_ext_t", C.c_float),
		("w_s", c_float_p), ("n_w_s", C.c_int32),
		("w_t", c_float_p), ("n_w_t", C.c_int32),
	]


class QueryParams(C.Structure):
	"""vx_query_params"""
	_fields_ = [
		("max_matches", C.c_int32),
		("min_score", C.c_float),
		("submatch_weight", C.c_float),
		("tag_weighted", C.c_int32),
		("t_weights", c_float_p),
		("t_pos", c_int8_p),
		("pos_mismatch_penalty", C.c_float),
		("similarity_threshold", C.c_float),
	]


class Hit(C.Structure):
	"""vx_hit"""
	_fields_ = [
		("score", C.c_float),
		("score_max", C.c_float),
		("doc", C.c_int32),
		("slice", C.c_int32),
	]


HIT_DTYPE = np.dtype([
	("score", np.float32), ("score_max", np.float32),
	("doc", np.int32), ("slice", np.int32)])


def ptr(a, ctype):
	"""Pointer to a C-contiguous numpy array (or NULL for None)."""
	if a is None:
		return C.cast(None, C.POINTER(ctype))
	assert a.flags["C_CONTIGUOUS"]
	return a.ctypes.data_as(C.POINTER(ctype))


def make_align_params(locality, gap_kind, lin=(0.0, 0.0), affine_s=(0.0, 0.0), affine_t=(0.0, 0.0),
		w_s=None, w_t=None):
	"""Returns (AlignParams, keepalive)."""
	ap = AlignParams()
	ap.locality = int(locality)
	ap.gap_kind = int(gap_kind)
	ap.lin_s, ap.lin_t = float(lin[0]), float(lin[1])
	ap.aff_open_s, ap.aff_ext_s = float(affine_s[0]), float(affine_s[1])
	ap.aff_open_t, ap.aff_ext_t = float(affine_t[0]), float(affine_t[1])
	keep = []
	if w_s is not None:
		w_s = np.ascontiguousarray(w_s, dtype=np.float32)
		w_t = np.ascontiguousarray(w_t, dtype=np.float32)
		keep = [w_s, w_t]
		ap.w_s, ap.n_w_s = ptr(w_s, C.c_float), len(w_s)
		ap.w_t, ap.n_w_t = ptr(w_t, C.c_float), len(w_t)
	return ap, keep


def make_query_params(max_matches=100, min_score=0.2, submatch_weight=0.0,
		t_weights=None, t_pos=None, pos_mismatch_penalty=0.0, similarity_threshold=0.0):
	"""Defaults follow core.Query.initialize (reference: vectorian/core/cpp/query.cpp:74-93)."""
	qp = QueryParams()
	qp.max_matches = int(max_matches)
	qp.min_score = float(min_score)
	qp.submatch_weight = float(submatch_weight)
	keep = []
	if t_weights is not None:
		t_weights = np.ascontiguousarray(t_weights, dtype=np.float32)
		t_pos = np.ascontiguousarray(t_pos, dtype=np.int8)
		keep = [t_weights, t_pos]
		qp.tag_weighted = 1
		qp.t_weights = ptr(t_weights, C.c_float)
		qp.t_pos = ptr(t_pos, C.c_int8)
		qp.pos_mismatch_penalty = float(pos_mismatch_penalty)
		qp.similarity_threshold = float(similarity_threshold)
	return qp, keep


class VxError(RuntimeError):
	"""Raised for any nonzero vx_status (the reference raises RuntimeError / ValueError from C++)."""


_lib = None
_lock = threading.Lock()

_SIGNATURES = {
	"vx_init": (C.c_int, [C.c_int]),
	"vx_last_error": (C.c_char_p, []),
	"vx_build_info": (C.c_char_p, []),
	"vx_device_sm_count": (C.c_int, []),
	"vx_corpus_create": (C.c_int, [
		c_int32_p, c_int8_p, C.c_int64, c_int64_p, C.c_int32,
		c_int32_p, c_int32_p, c_int64_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_corpus_create_dev": (C.c_int, [
		C.c_void_p, C.c_void_p, C.c_int64, c_int64_p, C.c_int32,
		c_int32_p, c_int32_p, c_int64_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_corpus_destroy": (C.c_int, [C.c_void_p]),
	"vx_corpus_n_windows": (C.c_int64, [C.c_void_p]),
	"vx_corpus_n_tokens": (C.c_int64, [C.c_void_p]),
	"vx_corpus_n_cells": (C.c_int64, [C.c_void_p]),
	"vx_corpus_max_len": (C.c_int32, [C.c_void_p]),
	"vx_corpus_set_boost": (C.c_int, [C.c_void_p, c_float_p, C.c_int64]),
	"vx_corpus_set_contextual": (C.c_int, [C.c_void_p, c_float_p, C.c_int64, C.c_int32]),
	"vx_embedding_create": (C.c_int, [c_float_p, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_embedding_destroy": (C.c_int, [C.c_void_p]),
	"vx_sim_static": (C.c_int, [C.c_void_p, c_int32_p, c_float_p, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_sim_upload": (C.c_int, [c_float_p, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_sim_upload_tokens": (C.c_int, [C.c_void_p, c_float_p, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_sim_contextual": (C.c_int, [c_float_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_sim_download": (C.c_int, [C.c_void_p, c_float_p, C.c_int64, C.c_int32]),
	"vx_sim_build_ms": (C.c_float, [C.c_void_p]),
	"vx_sim_launches": (C.c_int, [C.c_void_p]),
	"vx_sim_destroy": (C.c_int, [C.c_void_p]),
	"vx_search": (C.c_int, [
		C.c_void_p, C.c_void_p, C.POINTER(AlignParams), C.POINTER(QueryParams),
		C.POINTER(Hit), c_int16_p, c_float_p, c_int32_p]),
	"vx_search_topk_noalloc": (C.c_int, [
		C.c_void_p, C.c_void_p, C.POINTER(AlignParams), C.POINTER(QueryParams), C.POINTER(Hit), c_int32_p]),
	"vx_search_batch": (C.c_int, [
		C.c_void_p, C.POINTER(C.c_void_p), C.c_int32, C.POINTER(AlignParams), C.c_int32, C.POINTER(QueryParams), C.c_int32,
		C.POINTER(Hit), c_int16_p, c_float_p, c_int32_p]),
	"vx_score_all": (C.c_int, [
		C.c_void_p, C.c_void_p, C.POINTER(AlignParams), C.POINTER(QueryParams),
		c_float_p, C.c_int64]),
	"vx_abort": (C.c_int, [C.c_void_p]),
	"vx_comm_unique_id": (C.c_int, [C.c_void_p]),
	"vx_comm_create": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
	"vx_comm_destroy": (C.c_int, [C.c_void_p]),
	"vx_corpus_set_shard": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
	"vx_last_gather_ms": (C.c_float, [C.c_void_p]),
	"vx_res_block_bytes": (C.c_int64, [C.c_int32, C.c_int32, c_int64_p]),
	"vx_debug_merge": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
	"vx_last_timing": (C.c_int, [C.c_void_p, c_float_p]),
	"vx_last_launches": (C.c_int, [C.c_void_p]),
	"vx_debug_ctx_sim": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
		C.c_int32, C.c_void_p, C.c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES.keys())


def load_library(path=LIB_PATH):
	"""dlopen the C-ABI library and declare signatures.  Does not touch the GPU."""
	global _lib
	with _lock:
		if _lib is None:
			if not os.path.exists(path):
				raise VxError(
					f"{path} is missing: the CUDA extension has not been built "
					f"(run `python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU fallback")
			lib = C.CDLL(path)
			for name, (res, args) in _SIGNATURES.items():
				fn = getattr(lib, name)
				fn.restype = res
				fn.argtypes = args
			_lib = lib
	return _lib


_initialised = False


def lib():
	"""The library, with the CUDA device of this process selected (LOCAL_RANK or 0)."""
	global _initialised
	l = load_library()
	if not _initialised:
		device = int(os.environ.get("LOCAL_RANK", "0"))
		check(l.vx_init(device), l)
		_initialised = True
	return l


def check(status, l=None):
	if status != 0:
		l = l or load_library()
		msg = l.vx_last_error()
		raise VxError(f"vx_status {status}: {msg.decode() if msg else '?'}")
