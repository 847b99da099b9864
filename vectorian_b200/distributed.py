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


def _pack_rows(hits, target, distance, k, t, doc_offset):
	"""float32 [k + 1, 4 + 2T] = (score, score_max, doc, slice, target[T], dist[T]) rows + the row count in row k.
	doc / slice / target travel bit-cast (exact in float32 only up to 2^24 if converted)."""
	n = len(hits)
	buf = np.zeros((k + 1, 4 + 2 * t), dtype=np.float32)
	ibuf = buf.view(np.int32)
	buf[:n, 0] = hits["score"]
	buf[:n, 1] = hits["score_max"]
	ibuf[:n, 2] = hits["doc"] + doc_offset
	ibuf[:n, 3] = hits["slice"]
	if t:
		ibuf[:n, 4:4 + t] = target[:, :t].astype(np.int32)
		buf[:n, 4 + t:] = distance[:, :t]
	ibuf[k, 0] = n
	return buf


def _unpack_rows(out, k, t):
	iout = out.view(np.int32)
	m = int(iout[k, 0])
	h = np.zeros((m,), dtype=HIT_DTYPE)
	h["score"] = out[:m, 0]
	h["score_max"] = out[:m, 1]
	h["doc"] = iout[:m, 2]
	h["slice"] = iout[:m, 3]
	return h, iout[:m, 4:4 + t].astype(np.int16), out[:m, 4 + t:].copy()


def gather_topk_many(results, k, doc_offset=0, group=None, device=None):
	"""All-gather the local top-k of SEVERAL searches (e.g. the queries x settings pairs of a parameter sweep) in ONE
	collective and merge each; `results` is a list of (hits, target, dist) triples, every rank returns the list of
	global results."""
	if not (dist.is_available() and dist.is_initialized()):
		out = []
		for hits, target, distance in results:
			h = hits.copy()
			h["doc"] += doc_offset
			out.append(merge_hits([(h, target, distance)], k))
		return out
	if not results:
		return []
	world = dist.get_world_size(group)
	t = max(r[1].shape[1] if r[1].ndim == 2 else 0 for r in results)
	buf = np.stack([_pack_rows(h, tg, ds, k, tg.shape[1] if tg.ndim == 2 else 0, doc_offset) if (tg.ndim == 2 and tg.shape[1] == t)
		else _pack_rows(h, np.pad(tg, ((0, 0), (0, t - tg.shape[1])), constant_values=-1), np.pad(ds, ((0, 0), (0, t - ds.shape[1])), constant_values=1.0), k, t, doc_offset)
		for h, tg, ds in results])
	backend = dist.get_backend(group)
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
	send = torch.from_numpy(buf).to(device)
	recv = torch.empty((world,) + buf.shape, dtype=torch.float32, device=device)
	dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
	out = recv.cpu().numpy()
	merged = []
	for i, (h, tg, ds) in enumerate(results):
		ti = tg.shape[1] if tg.ndim == 2 else 0
		parts = []
		for r in range(world):
			hh, tt, dd = _unpack_rows(out[r, i], k, t)
			parts.append((hh, tt[:, :ti], dd[:, :ti]))
		merged.append(merge_hits(parts, k))
	return merged


def gather_topk(hits, target, distance, k, doc_offset=0, group=None, device=None):
	"""All-gather every rank's local top-k and merge; every rank returns the global result (one collective of
	k x (16 + 8T) bytes per rank)."""
	return gather_topk_many([(hits, target, distance)], k, doc_offset=doc_offset, group=group, device=device)[0]
