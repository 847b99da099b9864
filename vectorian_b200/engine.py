"""Thin object layer over the C ABI: device corpus, embedding, similarity matrix, search.

These are the objects the reference keeps behind pybind11 (``core.Document`` /
``core.Vocabulary`` / ``core.Query`` / ``core.ResultSet``, reference:
vectorian/core/cpp/module.cpp:30-152); everything here forwards to
libvectorian_b200.so and raises if the CUDA library or device is missing.
"""

import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import ptr, check, HIT_DTYPE, Hit


MAX_SIM_COLUMNS = 16   # query columns per vx_sim_static launch (VX_MAX_T_FAST)


class Comm:
	"""The ranks of a sentence-sharded corpus (vx_comm: an NCCL communicator bound inside the C ABI).  The reference's
	counterpart is the document thread pool sharing one ResultSet (vectorian/index.py:544-555)."""

	def __init__(self, unique_id, world, rank):
		lib = _cabi.lib()
		self._lib = lib
		self._h = C.c_void_p()
		uid = np.ascontiguousarray(unique_id, dtype=np.uint8)
		assert uid.shape == (128,)
		self.world, self.rank = int(world), int(rank)
		check(lib.vx_comm_create(uid.ctypes.data_as(C.c_void_p), self.world, self.rank, C.byref(self._h)), lib)

	@staticmethod
	def unique_id():
		lib = _cabi.lib()
		uid = np.zeros((128,), dtype=np.uint8)
		check(lib.vx_comm_unique_id(uid.ctypes.data_as(C.c_void_p)), lib)
		return uid

	@property
	def handle(self):
		return self._h

	def close(self):
		if self._h:
			self._lib.vx_comm_destroy(self._h)
			self._h = C.c_void_p()


def merge_blocks(blocks, t, k):
	"""vx_debug_merge: merge per-rank (hits, target, dist) triples with the DEVICE merge kernel of the sharded search
	(test hook; document ids must already be global).  Returns the merged triple."""
	lib = _cabi.lib()
	off = np.zeros((8,), dtype=np.int64)
	nbytes = int(lib.vx_res_block_bytes(t, k, ptr(off, C.c_int64)))
	o_score, o_max, o_win, o_doc, o_slice, o_target, o_dist, o_n = [int(x) for x in off]
	buf = np.zeros((len(blocks), nbytes), dtype=np.uint8)
	for r, (h, tg, ds) in enumerate(blocks):
		n = len(h)
		assert n <= k
		row = buf[r]
		row[o_score:o_score + 4 * n].view(np.float32)[:] = h["score"]
		row[o_max:o_max + 4 * n].view(np.float32)[:] = h["score_max"]
		row[o_doc:o_doc + 4 * n].view(np.int32)[:] = h["doc"]
		row[o_slice:o_slice + 4 * n].view(np.int32)[:] = h["slice"]
		if t:
			row[o_target:o_target + 2 * n * t].view(np.int16)[:] = np.asarray(tg, dtype=np.int16).reshape(-1)
			row[o_dist:o_dist + 4 * n * t].view(np.float32)[:] = np.asarray(ds, dtype=np.float32).reshape(-1)
		row[o_n:o_n + 4].view(np.int32)[0] = n
	out = np.zeros((nbytes,), dtype=np.uint8)
	check(lib.vx_debug_merge(buf.ctypes.data_as(C.c_void_p), len(blocks), t, k, out.ctypes.data_as(C.c_void_p)), lib)
	n = int(out[o_n:o_n + 4].view(np.int32)[0])
	hits = np.zeros((n,), dtype=HIT_DTYPE)
	hits["score"] = out[o_score:o_score + 4 * n].view(np.float32)
	hits["score_max"] = out[o_max:o_max + 4 * n].view(np.float32)
	hits["doc"] = out[o_doc:o_doc + 4 * n].view(np.int32)
	hits["slice"] = out[o_slice:o_slice + 4 * n].view(np.int32)
	target = out[o_target:o_target + 2 * n * t].view(np.int16).reshape(n, t).copy()
	dist = out[o_dist:o_dist + 4 * n * t].view(np.float32).reshape(n, t).copy()
	return hits, target, dist


class DeviceCorpus:
	"""Documents packed once into device-resident token tiles for one partition.

	Replaces the list of ``core.Document`` + ``Spans::iterate``
	(reference: vectorian/core/cpp/document.h:147-169, cpp/document.cpp:30-58).
	"""

	def __init__(self, token_id, doc_tok_off, span_start, span_end, doc_span_off,
			window_size=1, window_step=1, token_pos=None, on_device=False):
		lib = _cabi.lib()
		self._lib = lib
		self._h = C.c_void_p()
		doc_tok_off = np.ascontiguousarray(doc_tok_off, dtype=np.int64)
		span_start = np.ascontiguousarray(span_start, dtype=np.int32)
		span_end = np.ascontiguousarray(span_end, dtype=np.int32)
		doc_span_off = np.ascontiguousarray(doc_span_off, dtype=np.int64)
		n_doc = len(doc_tok_off) - 1
		if on_device:
			# token_id / token_pos are torch CUDA tensors (int32 / int8)
			n_tok = int(token_id.numel())
			check(lib.vx_corpus_create_dev(
				C.c_void_p(token_id.data_ptr()),
				C.c_void_p(token_pos.data_ptr() if token_pos is not None else None),
				n_tok, ptr(doc_tok_off, C.c_int64), n_doc,
				ptr(span_start, C.c_int32), ptr(span_end, C.c_int32), ptr(doc_span_off, C.c_int64),
				int(window_size), int(window_step), C.byref(self._h)), lib)
		else:
			token_id = np.ascontiguousarray(token_id, dtype=np.int32)
			if token_pos is not None:
				token_pos = np.ascontiguousarray(token_pos, dtype=np.int8)
			check(lib.vx_corpus_create(
				ptr(token_id, C.c_int32), ptr(token_pos, C.c_int8), len(token_id),
				ptr(doc_tok_off, C.c_int64), n_doc,
				ptr(span_start, C.c_int32), ptr(span_end, C.c_int32), ptr(doc_span_off, C.c_int64),
				int(window_size), int(window_step), C.byref(self._h)), lib)
		self.n_doc = n_doc
		self.n_spans = int(doc_span_off[-1])
		self.sharded = False     # set_shard: search / search_batch return the GLOBAL result on every rank

	def set_shard(self, comm, doc_offset):
		"""vx_corpus_set_shard: this corpus is one rank's shard; doc_offset = global id of its first document."""
		check(self._lib.vx_corpus_set_shard(self._h, comm.handle if comm is not None else None, int(doc_offset)), self._lib)
		self.sharded = comm is not None
		self._comm = comm

	def close(self):
		if self._h:
			self._lib.vx_corpus_destroy(self._h)
			self._h = C.c_void_p()

	def __del__(self):
		try:
			self.close()
		except Exception:
			pass

	@property
	def handle(self):
		return self._h

	@property
	def n_windows(self):
		return int(self._lib.vx_corpus_n_windows(self._h))

	@property
	def n_tokens(self):
		return int(self._lib.vx_corpus_n_tokens(self._h))

	@property
	def n_cells(self):
		return int(self._lib.vx_corpus_n_cells(self._h))

	@property
	def max_len(self):
		return int(self._lib.vx_corpus_max_len(self._h))

	def set_boost(self, boost):
		if boost is None:
			check(self._lib.vx_corpus_set_boost(self._h, ptr(None, C.c_float), 0), self._lib)
			return
		boost = np.ascontiguousarray(boost, dtype=np.float32)
		check(self._lib.vx_corpus_set_boost(self._h, ptr(boost, C.c_float), len(boost)), self._lib)

	def set_contextual(self, x):
		"""x: [n_tokens x d] float32, a numpy array or a contiguous torch CUDA tensor (copied device to device)."""
		if hasattr(x, "data_ptr"):
			if not x.is_contiguous() or x.element_size() != 4:
				raise ValueError("contextual vectors must be a contiguous float32 tensor")
			check(self._lib.vx_corpus_set_contextual(self._h, C.cast(C.c_void_p(x.data_ptr()), C.POINTER(C.c_float)),
				x.shape[0], x.shape[1]), self._lib)
			return
		x = np.ascontiguousarray(x, dtype=np.float32)
		check(self._lib.vx_corpus_set_contextual(self._h, ptr(x, C.c_float), x.shape[0], x.shape[1]), self._lib)

	def abort(self):
		self._lib.vx_abort(self._h)

	def search(self, sim, ap, qp):
		"""Returns (hits structured array, target[n x T] int16, dist[n x T] float32)."""
		k = max(int(qp.max_matches), 1)
		t = max(sim.t, 1)
		hits = np.zeros((k,), dtype=HIT_DTYPE)
		target = np.full((k, t), -1, dtype=np.int16)
		dist = np.ones((k, t), dtype=np.float32)
		n = C.c_int32(0)
		check(self._lib.vx_search(self._h, sim.handle, C.byref(ap), C.byref(qp),
			hits.ctypes.data_as(C.POINTER(Hit)), ptr(target, C.c_int16), ptr(dist, C.c_float), C.byref(n)), self._lib)
		return hits[:n.value], target[:n.value, :sim.t], dist[:n.value, :sim.t]

	def search_topk_noalloc(self, sim, ap, qp):
		"""vx_search_topk_noalloc: the hits of `search` (no tracebacks) through the entry point that allocates nothing
		(scratch in a static device array; max_matches <= VX_NOALLOC_MAX_K, fast paths only).  No counterpart in the
		reference."""
		k = max(int(qp.max_matches), 1)
		hits = np.zeros((k,), dtype=HIT_DTYPE)
		n = C.c_int32(0)
		check(self._lib.vx_search_topk_noalloc(self._h, sim.handle, C.byref(ap), C.byref(qp),
			hits.ctypes.data_as(C.POINTER(Hit)), C.byref(n)), self._lib)
		return hits[:n.value]

	def search_batch(self, sims, aps, qp, raw=False):
		"""Parameter sweep (vx_search_batch): every query in `sims` against every alignment setting in `aps` in one
		call.  Returns a list over queries of lists over settings of (hits, target, dist) as `search` returns them."""
		n_q, n_a = len(sims), len(aps)
		k = max(int(qp.max_matches), 1)
		t_stride = max([s.t for s in sims] + [1])
		n_pairs = n_q * n_a
		hits = np.zeros((n_pairs, k), dtype=HIT_DTYPE)
		target = np.full((n_pairs, k, t_stride), -1, dtype=np.int16)
		dist = np.ones((n_pairs, k, t_stride), dtype=np.float32)
		n = np.zeros((max(n_pairs, 1),), dtype=np.int32)
		handles = (C.c_void_p * max(n_q, 1))(*[s.handle for s in sims])
		ap_arr = (_cabi.AlignParams * max(n_a, 1))()
		for i, a in enumerate(aps):
			C.memmove(C.byref(ap_arr[i]), C.byref(a), C.sizeof(_cabi.AlignParams))   # keeps pointers into the caller's tables
		check(self._lib.vx_search_batch(self._h, handles, n_q, ap_arr, n_a, C.byref(qp), t_stride,
			hits.ctypes.data_as(C.POINTER(Hit)), ptr(target, C.c_int16), ptr(dist, C.c_float), ptr(n, C.c_int32)), self._lib)
		if raw:
			# unsplit arrays for distributed.gather_topk_many (pair p = q * len(aps) + a)
			from .distributed import BatchRows
			return BatchRows(hits, target, dist, n[:n_pairs], [sims[p // max(n_a, 1)].t for p in range(n_pairs)])
		out = []
		for q in range(n_q):
			row = []
			for a in range(n_a):
				p = q * n_a + a
				m = int(n[p])
				row.append((hits[p, :m], target[p, :m, :sims[q].t], dist[p, :m, :sims[q].t]))
			out.append(row)
		return out

	def score_all(self, sim, ap, qp):
		n = self.n_windows
		out = np.zeros((n,), dtype=np.float32)
		check(self._lib.vx_score_all(self._h, sim.handle, C.byref(ap), C.byref(qp), ptr(out, C.c_float), n), self._lib)
		return out

	def last_timing(self):
		ms = np.zeros((3,), dtype=np.float32)
		check(self._lib.vx_last_timing(self._h, ptr(ms, C.c_float)), self._lib)
		return dict(dp_ms=float(ms[0]), topk_ms=float(ms[1]), traceback_ms=float(ms[2]),
			gather_merge_ms=float(self._lib.vx_last_gather_ms(self._h)))

	def last_launches(self):
		return int(self._lib.vx_last_launches(self._h))


class DeviceEmbedding:
	"""A static embedding table, L2-normalised on the device (reference: embedding/vectors.py:71-86)."""

	def __init__(self, vectors):
		lib = _cabi.lib()
		self._lib = lib
		self._h = C.c_void_p()
		vectors = np.ascontiguousarray(vectors, dtype=np.float32)
		self.shape = vectors.shape
		check(lib.vx_embedding_create(ptr(vectors, C.c_float), vectors.shape[0], vectors.shape[1], C.byref(self._h)), lib)

	def close(self):
		if self._h:
			self._lib.vx_embedding_destroy(self._h)
			self._h = C.c_void_p()

	def __del__(self):
		try:
			self.close()
		except Exception:
			pass

	@property
	def handle(self):
		return self._h


class DeviceSim:
	"""The query's similarity matrix on the device (reference: cpp/metric/metric.h:9-52)."""

	def __init__(self, handle, rows, t, lib):
		self._h = handle
		self._lib = lib
		self.rows = rows
		self.t = t

	@classmethod
	def static(cls, embedding, q_id, q_vec):
		"""S = clip01(cos(E, Q)) with the identity fix-up (reference: cpp/metric/static.cpp:9-78)."""
		lib = _cabi.lib()
		q_id = np.ascontiguousarray(q_id, dtype=np.int32)
		q_vec = np.ascontiguousarray(q_vec, dtype=np.float32).reshape(len(q_id), -1)
		h = C.c_void_p()
		check(lib.vx_sim_static(embedding.handle, ptr(q_id, C.c_int32), ptr(q_vec, C.c_float), len(q_id), C.byref(h)), lib)
		return cls(h, embedding.shape[0], len(q_id), lib)

	@classmethod
	def upload(cls, s):
		lib = _cabi.lib()
		s = np.ascontiguousarray(s, dtype=np.float32)
		h = C.c_void_p()
		check(lib.vx_sim_upload(ptr(s, C.c_float), s.shape[0], s.shape[1], C.byref(h)), lib)
		return cls(h, s.shape[0], s.shape[1], lib)

	@classmethod
	def upload_tokens(cls, corpus, s_tok):
		"""Per-token similarity rows S_tok[n_tokens x T] (token order of the corpus), valid for `corpus` only
		(reference: ContextualSimilarityMatrix, cpp/metric/modifier.cpp:18-74, cpp/metric/static.cpp:154-189)."""
		lib = _cabi.lib()
		s_tok = np.ascontiguousarray(s_tok, dtype=np.float32)
		h = C.c_void_p()
		check(lib.vx_sim_upload_tokens(corpus.handle, ptr(s_tok, C.c_float), s_tok.shape[0], s_tok.shape[1], C.byref(h)), lib)
		return cls(h, s_tok.shape[0], s_tok.shape[1], lib)

	@classmethod
	def contextual(cls, q_vec):
		lib = _cabi.lib()
		q_vec = np.ascontiguousarray(q_vec, dtype=np.float32)
		h = C.c_void_p()
		check(lib.vx_sim_contextual(ptr(q_vec, C.c_float), q_vec.shape[0], q_vec.shape[1], C.byref(h)), lib)
		return cls(h, 0, q_vec.shape[0], lib)

	@property
	def build_ms(self):
		return float(self._lib.vx_sim_build_ms(self._h))

	@property
	def launches(self):
		return int(self._lib.vx_sim_launches(self._h))

	def download(self):
		out = np.zeros((self.rows, self.t), dtype=np.float32)
		check(self._lib.vx_sim_download(self._h, ptr(out, C.c_float), self.rows, self.t), self._lib)
		return out

	def close(self):
		if self._h:
			self._lib.vx_sim_destroy(self._h)
			self._h = C.c_void_p()

	def __del__(self):
		try:
			self.close()
		except Exception:
			pass

	@property
	def handle(self):
		return self._h
