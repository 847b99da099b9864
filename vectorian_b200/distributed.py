"""Sentence-sharded multi-GPU search: one process per GPU, one small gather of per-rank top-k.

The reference has no distributed path; its only fan-out is a thread pool over documents whose
per-document ResultSets are merged with ResultSet::extend (reference: vectorian/index.py:544-555,
vectorian/core/cpp/result_set.h:70-93).  Here each rank owns a contiguous range of documents,
runs the whole device search on its shard, and the k best (score, doc, slice, flow) rows of every
rank are all-gathered (k x (16 + 6T) bytes per rank, NCCL over NVLink on GPUs, gloo on CPU) and merged
with the same total order (score desc, doc desc, slice desc).
"""

import numpy as np
import torch
import torch.distributed as dist

from ._cabi import HIT_DTYPE


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


def gather_topk(hits, target, distance, k, doc_offset=0, group=None, device=None):
	"""All-gather every rank's local top-k and merge; every rank returns the global result.

	One collective: rows are packed as float32 [k, 4 + 2T] = (score, score_max, doc, slice,
	target[T], dist[T]) plus the row count in row k.  doc/slice/target are exact in float32 only up
	to 2^24, so they travel bit-cast, not converted.
	"""
	if not (dist.is_available() and dist.is_initialized()):
		h = hits.copy()
		h["doc"] += doc_offset
		return merge_hits([(h, target, distance)], k)
	world = dist.get_world_size(group)
	t = target.shape[1] if target.ndim == 2 else 0
	n = len(hits)
	width = 4 + 2 * t
	buf = np.zeros((k + 1, width), dtype=np.float32)
	ibuf = buf.view(np.int32)
	buf[:n, 0] = hits["score"]
	buf[:n, 1] = hits["score_max"]
	ibuf[:n, 2] = hits["doc"] + doc_offset
	ibuf[:n, 3] = hits["slice"]
	ibuf[:n, 4:4 + t] = target.astype(np.int32)
	buf[:n, 4 + t:] = distance
	ibuf[k, 0] = n
	backend = dist.get_backend(group)
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
	send = torch.from_numpy(buf).to(device)
	recv = torch.empty((world, k + 1, width), dtype=torch.float32, device=device)
	dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
	out = recv.cpu().numpy()
	iout = out.view(np.int32)
	parts = []
	for r in range(world):
		m = int(iout[r, k, 0])
		h = np.zeros((m,), dtype=HIT_DTYPE)
		h["score"] = out[r, :m, 0]
		h["score_max"] = out[r, :m, 1]
		h["doc"] = iout[r, :m, 2]
		h["slice"] = iout[r, :m, 3]
		parts.append((h, iout[r, :m, 4:4 + t].astype(np.int16), out[r, :m, 4 + t:].copy()))
	return merge_hits(parts, k)
