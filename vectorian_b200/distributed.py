"""Sentence-sharded multi-GPU search: one process per GPU, one small gather of per-rank top-k.

The reference has no distributed path; its only fan-out is a thread pool over documents whose
per-document ResultSets are merged with ResultSet::extend (reference: vectorian/index.py:544-555,
vectorian/core/cpp/result_set.h:70-93).  Here each rank owns a contiguous range of documents and runs the
whole device search on its shard.

On GPUs (NCCL process group) the exchange happens INSIDE the C ABI: `device_comm()` builds a vx_comm (the unique
id travels over torch.distributed once), `DeviceCorpus.set_shard` attaches it, and from then on vx_search /
vx_search_batch all-gather the k result rows of every rank on the corpus stream, merge them with one kernel under
(score desc, doc desc, slice desc) and copy k rows to the host (csrc/vx_shard.cu) -- no host bounce.

`gather_topk[_many]` below is the host path for process groups without NCCL (gloo: the CPU tests of the shard /
merge logic) with ONE merge implementation, `merge_hits`.
"""

import numpy as np
import torch
import torch.distributed as dist

from ._cabi import HIT_DTYPE

_comm = None


def device_comm(group=None):
	"""The vx_comm of this process group (NCCL backends only), created once: rank 0 draws the NCCL unique id,
	torch.distributed broadcasts its 128 bytes, every rank joins.  None without a process group or without NCCL."""
	global _comm
	if _comm is not None:
		return _comm
	if not (dist.is_available() and dist.is_initialized()) or dist.get_backend(group) != "nccl":
		return None
	from . import engine
	rank, world = dist.get_rank(group), dist.get_world_size(group)
	uid = engine.Comm.unique_id() if rank == 0 else np.zeros((128,), dtype=np.uint8)
	t = torch.from_numpy(uid).to(torch.device("cuda", torch.cuda.current_device()))
	dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
	_comm = engine.Comm(t.cpu().numpy(), world, rank)
	return _comm


def shard_documents(n_doc, world_size, rank, weights=None):
	"""Contiguous document range [lo, hi) of `rank`, balanced by `weights` (e.g. DP cells per document)."""
	if weights is None:
		weights = np.ones((n_doc,), dtype=np.float64)
	cum = np.concatenate([[0.0], np.cumsum(np.asarray(weights, dtype=np.float64))])
	total = cum[-1]
	bounds = [int(np.searchsorted(cum, total * r / world_size, side="left")) for r in range(world_size + 1)]
	bounds[0], bounds[-1] = 0, n_doc
	for r in range(1, world_size + 1):
		bounds[r] = max(bounds[r], bounds[r - 1])
	return bounds[rank], bounds[rank + 1]


def merge_hits(parts, k):
	"""Merge per-shard (hits, target, dist) triples (document ids already global) into the best k."""
	hits = np.concatenate([p[0] for p in parts]) if parts else np.zeros((0,), dtype=HIT_DTYPE)
	if len(hits) == 0:
		t = parts[0][1].shape[1] if parts else 0
		return hits, np.zeros((0, t), np.int16), np.zeros((0, t), np.float32)
	target = np.concatenate([p[1] for p in parts])
	distance = np.concatenate([p[2] for p in parts])
	# score desc, doc desc, slice desc  (Match::compare_by_score made total)
	order = np.lexsort((-hits["slice"].astype(np.int64), -hits["doc"].astype(np.int64), -hits["score"].astype(np.float64)))
	order = order[:k]
	return hits[order], target[order], distance[order]


class BatchRows:
	"""The result arrays of one vx_search_batch call, unsplit: hits [P, k] (HIT_DTYPE), target [P, k, T] int16,
	dist [P, k, T] float32, n [P] valid rows per pair, widths [P] query length of every pair (engine.search_batch(raw=True))."""

	def __init__(self, hits, target, dist, n, widths):
		self.hits, self.target, self.dist, self.n, self.widths = hits, target, dist, np.asarray(n), list(widths)

	def triples(self):
		return [(self.hits[i, :m], self.target[i, :m, :w], self.dist[i, :m, :w]) for i, (m, w) in enumerate(zip(self.n, self.widths))]


def gather_topk_many(results, k, doc_offset=0, group=None, device=None):
	"""Host path (no NCCL): all-gather the local top-k of SEVERAL searches (e.g. the queries x settings pairs of a
	parameter sweep) in ONE collective and merge each with merge_hits; `results` is a list of (hits, target, dist)
	triples or a BatchRows, every rank returns the list of global results."""
	triples = results.triples() if isinstance(results, BatchRows) else list(results)
	if not (dist.is_available() and dist.is_initialized()):
		out = []
		for hits, target, distance in triples:
			h = hits.copy()
			h["doc"] += doc_offset
			out.append(merge_hits([(h, target, distance)], k))
		return out
	if not triples:
		return []
	world = dist.get_world_size(group)
	widths = [r[1].shape[1] if r[1].ndim == 2 else 0 for r in triples]
	t = max(widths)
	n_pairs = len(triples)
	# one float32 / int32 row per hit: score, score_max, doc, slice, target[t], dist[t]; row k of a pair holds its count
	buf = np.zeros((n_pairs, k + 1, 4 + 2 * t), dtype=np.float32)
	ibuf = buf.view(np.int32)
	for i, (h, tg, ds) in enumerate(triples):
		n, ti = len(h), widths[i]
		buf[i, :n, 0] = h["score"]
		buf[i, :n, 1] = h["score_max"]
		ibuf[i, :n, 2] = h["doc"] + doc_offset
		ibuf[i, :n, 3] = h["slice"]
		if ti:
			ibuf[i, :n, 4:4 + ti] = tg
			buf[i, :n, 4 + t:4 + t + ti] = ds
		ibuf[i, k, 0] = n
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
	send = torch.from_numpy(buf).to(device)
	recv = torch.empty((world,) + buf.shape, dtype=torch.float32, device=device)
	dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
	out = recv.cpu().numpy()
	iout = out.view(np.int32)
	merged = []
	for i in range(n_pairs):
		parts = []
		for r in range(world):
			m = int(iout[r, i, k, 0])
			h = np.zeros((m,), dtype=HIT_DTYPE)
			h["score"] = out[r, i, :m, 0]
			h["score_max"] = out[r, i, :m, 1]
			h["doc"] = iout[r, i, :m, 2]
			h["slice"] = iout[r, i, :m, 3]
			parts.append((h, iout[r, i, :m, 4:4 + widths[i]].astype(np.int16), out[r, i, :m, 4 + t:4 + t + widths[i]].copy()))
		merged.append(merge_hits(parts, k))
	return merged


def gather_topk(hits, target, distance, k, doc_offset=0, group=None, device=None):
	"""Host path (no NCCL): all-gather every rank's local top-k and merge; every rank returns the global result."""
	return gather_topk_many([(hits, target, distance)], k, doc_offset=doc_offset, group=group, device=device)[0]
