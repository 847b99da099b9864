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

_staging = {}   # (buffer shape, world, device) -> pinned / device staging tensors of gather_topk_many
HOST_MERGE_ROWS = 16384   # gathered rows up to which gather_topk_many merges on the host instead of the device


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
	"""All-gather the local top-k of SEVERAL searches (e.g. the queries x settings pairs of a parameter sweep) in ONE
	collective and merge each; `results` is a list of (hits, target, dist) triples, every rank returns the list of
	global results.  The merge runs where the gathered rows are (torch, on the GPU under NCCL): two sorts over
	[pairs, world * k] keys instead of a Python loop per pair, then one copy of the pairs x k winning rows to the host."""
	if not (dist.is_available() and dist.is_initialized()):
		out = []
		for hits, target, distance in (results.triples() if isinstance(results, BatchRows) else results):
			h = hits.copy()
			h["doc"] += doc_offset
			out.append(merge_hits([(h, target, distance)], k))
		return out
	if not isinstance(results, BatchRows) and not results:
		return []
	world = dist.get_world_size(group)
	if isinstance(results, BatchRows):
		# the arrays of vx_search_batch as they are: packed with a handful of whole-array copies
		n_pairs, t = results.hits.shape[0], results.target.shape[2]
		widths = list(results.widths)
		buf = np.zeros((n_pairs, k + 1, 4 + 2 * t), dtype=np.float32)
		ibuf = buf.view(np.int32)
		live = np.arange(k)[None, :] < results.n[:, None]
		buf[:, :k, 0] = np.where(live, results.hits["score"], 0.0)
		buf[:, :k, 1] = np.where(live, results.hits["score_max"], 0.0)
		ibuf[:, :k, 2] = np.where(live, results.hits["doc"] + doc_offset, 0)
		ibuf[:, :k, 3] = np.where(live, results.hits["slice"], 0)
		ibuf[:, :k, 4:4 + t] = results.target
		buf[:, :k, 4 + t:] = results.dist
		ibuf[:, k, 0] = results.n
	else:
		widths = [r[1].shape[1] if r[1].ndim == 2 else 0 for r in results]
		t = max(widths)
		n_pairs = len(results)
		buf = np.zeros((n_pairs, k + 1, 4 + 2 * t), dtype=np.float32)
		ibuf = buf.view(np.int32)
		buf[:, :, 4 + t:] = 1.0
		ibuf[:, :, 4:4 + t] = -1
		for i, (h, tg, ds) in enumerate(results):
			n, ti = len(h), widths[i]
			buf[i, :n, 0] = h["score"]
			buf[i, :n, 1] = h["score_max"]
			ibuf[i, :n, 2] = h["doc"] + doc_offset
			ibuf[i, :n, 3] = h["slice"]
			if ti:
				ibuf[i, :n, 4:4 + ti] = tg
				buf[i, :n, 4 + t:4 + t + ti] = ds
			ibuf[i, k, 0] = n
	backend = dist.get_backend(group)
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
	host_merge = n_pairs * world * k <= HOST_MERGE_ROWS
	if host_merge and device.type == "cuda":
		# the latency path of a single search: cached pinned staging + device buffers, asynchronous copies, one wait
		key = (buf.shape, world, device.index)
		st = _staging.get(key)
		if st is None:
			st = (torch.empty(buf.shape, dtype=torch.float32).pin_memory(), torch.empty(buf.shape, dtype=torch.float32, device=device),
				torch.empty((world,) + buf.shape, dtype=torch.float32, device=device), torch.empty((world,) + buf.shape, dtype=torch.float32).pin_memory())
			if len(_staging) > 32:
				_staging.clear()
			_staging[key] = st
		pin_send, send, recv, pin_recv = st
		pin_send.numpy()[...] = buf
		send.copy_(pin_send, non_blocking=True)
		dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
		pin_recv.copy_(recv, non_blocking=True)
		torch.cuda.current_stream(device).synchronize()
		recv_host = pin_recv.numpy()
		pin_top = None
	elif device.type == "cuda":
		# sweeps: the same staging (pinned send buffer, cached device buffers); the merge below stays on the device
		key = ("sweep", buf.shape, world, device.index)
		st = _staging.get(key)
		if st is None:
			st = (torch.empty(buf.shape, dtype=torch.float32).pin_memory(), torch.empty(buf.shape, dtype=torch.float32, device=device),
				torch.empty((world,) + buf.shape, dtype=torch.float32, device=device),
				torch.empty((n_pairs, k, buf.shape[2]), dtype=torch.float32).pin_memory())
			if len(_staging) > 32:
				_staging.clear()
			_staging[key] = st
		pin_send, send, recv, pin_top = st
		pin_send.numpy()[...] = buf
		send.copy_(pin_send, non_blocking=True)
		dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
		recv_host = None
	else:
		send = torch.from_numpy(buf).to(device)
		recv = torch.empty((world,) + buf.shape, dtype=torch.float32, device=device)
		dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
		recv_host = None
		pin_top = None

	if host_merge:
		# a single search (or a handful): one copy to the host and a lexsort per pair cost less than a dozen small kernels
		out = recv_host if recv_host is not None else recv.cpu().numpy()
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

	# merge on the device: total order score desc, doc desc, slice desc (Match::compare_by_score made total)
	irecv = recv.view(torch.int32)
	count = irecv[:, :, k, 0].permute(1, 0)                                   # [pairs, world]
	rows = recv[:, :, :k, :].permute(1, 0, 2, 3).reshape(n_pairs, world * k, -1)
	irows = irecv[:, :, :k, :].permute(1, 0, 2, 3).reshape(n_pairs, world * k, -1)
	valid = (torch.arange(k, device=device)[None, None, :] < count[:, :, None]).reshape(n_pairs, world * k)
	score = torch.where(valid, rows[:, :, 0], torch.full_like(rows[:, :, 0], float("-inf")))
	tie = irows[:, :, 2].to(torch.int64) * (1 << 32) + irows[:, :, 3].to(torch.int64)
	tie = torch.where(valid, tie, torch.full_like(tie, -1))
	o1 = torch.argsort(tie, dim=1, descending=True)
	o2 = torch.sort(torch.gather(score, 1, o1), dim=1, descending=True, stable=True).indices
	order = torch.gather(o1, 1, o2)[:, :k]
	top = torch.gather(rows, 1, order[:, :, None].expand(-1, -1, rows.shape[2])).contiguous()
	n_out = torch.clamp(valid.sum(dim=1), max=k)
	if device.type == "cuda" and pin_top is not None:
		pin_top.copy_(top, non_blocking=True)
		n_h = n_out.cpu().numpy()             # synchronises the stream: the pinned copy above has landed
		top_h = pin_top.numpy()
	else:
		top_h = top.cpu().numpy()
		n_h = n_out.cpu().numpy()
	itop = top_h.view(np.int32)
	all_hits = np.zeros((n_pairs, k), dtype=HIT_DTYPE)
	all_hits["score"] = top_h[:, :, 0]
	all_hits["score_max"] = top_h[:, :, 1]
	all_hits["doc"] = itop[:, :, 2]
	all_hits["slice"] = itop[:, :, 3]
	all_target = itop[:, :, 4:4 + t].astype(np.int16)
	all_dist = np.ascontiguousarray(top_h[:, :, 4 + t:])
	return [(all_hits[i, :n_h[i]], all_target[i, :n_h[i], :widths[i]], all_dist[i, :n_h[i], :widths[i]]) for i in range(n_pairs)]


def gather_topk(hits, target, distance, k, doc_offset=0, group=None, device=None):
	"""All-gather every rank's local top-k and merge; every rank returns the global result (one collective of
	k x (16 + 8T) bytes per rank)."""
	return gather_topk_many([(hits, target, distance)], k, doc_offset=doc_offset, group=group, device=device)[0]
