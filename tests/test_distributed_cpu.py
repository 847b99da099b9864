"""Sentence-sharded search, host logic on CPU: world_size-2 gloo all-gather + merge of per-rank top-k."""

import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vectorian_b200._cabi import HIT_DTYPE
from vectorian_b200.distributed import gather_topk, gather_topk_many, merge_hits, shard_documents


def _free_port():
	with socket.socket() as s:
		s.bind(("127.0.0.1", 0))
		return s.getsockname()[1]


def _rank_rows(rank, n, t, k_seed=0):
	rng = np.random.default_rng(100 + rank + k_seed)
	hits = np.zeros((n,), dtype=HIT_DTYPE)
	# scores on a coarse grid so that ties across ranks occur
	hits["score"] = np.sort(rng.integers(0, 12, size=n).astype(np.float32) / 16)[::-1]
	hits["score_max"] = t
	hits["doc"] = rng.integers(0, 5, size=n)          # local document ids
	hits["slice"] = rng.permutation(1000)[:n]
	target = rng.integers(-1, 40000, size=(n, t)).astype(np.int16)
	distance = rng.random((n, t)).astype(np.float32)
	return hits, target, distance


def _worker(rank, world, port, k, t, out_dir):
	os.environ["MASTER_ADDR"] = "127.0.0.1"
	os.environ["MASTER_PORT"] = str(port)
	dist.init_process_group("gloo", rank=rank, world_size=world)
	n_local = [7, 3][rank]
	hits, target, distance = _rank_rows(rank, n_local, t)
	g = gather_topk(hits, target, distance, k, doc_offset=rank * 5)
	np.savez(os.path.join(out_dir, f"r{rank}.npz"), hits=g[0], target=g[1], distance=g[2])
	# several searches (a sweep's pairs, queries of different lengths) in one collective
	import vectorian_b200.distributed as vd
	rows = [_rank_rows(rank, n_local, tt, k_seed=10 * i) for i, tt in enumerate((t, 3, t))]
	many = gather_topk_many(rows, k, doc_offset=rank * 5)
	# the unsplit arrays of a batched search (engine.search_batch(raw=True)) give the same answer
	rows3 = [_rank_rows(rank, n_local, t, k_seed=10 * i) for i in range(3)]
	kk = 8
	hits = np.zeros((3, kk), dtype=HIT_DTYPE)
	target = np.full((3, kk, t), -1, dtype=np.int16)
	distance = np.ones((3, kk, t), dtype=np.float32)
	for i, (h, tg, ds) in enumerate(rows3):
		hits[i, :len(h)], target[i, :len(h)], distance[i, :len(h)] = h, tg, ds
	want = gather_topk_many(rows3, kk, doc_offset=rank * 5)
	got = gather_topk_many(vd.BatchRows(hits, target, distance, [len(r[0]) for r in rows3], [t] * 3), kk, doc_offset=rank * 5)
	for g0, g1 in zip(want, got):
		for x0, x1 in zip(g0, g1):
			np.testing.assert_array_equal(x0, x1)
	np.savez(os.path.join(out_dir, f"m{rank}.npz"), **{f"{name}{i}": g[j] for i, g in enumerate(many) for j, name in enumerate(("hits", "target", "distance"))})
	dist.destroy_process_group()


def test_gather_topk_world2(tmp_path):
	world, k, t = 2, 6, 5
	mp.spawn(_worker, args=(world, _free_port(), k, t, str(tmp_path)), nprocs=world, join=True)
	parts = []
	for rank, n_local in enumerate([7, 3]):
		h, tg, ds = _rank_rows(rank, n_local, t)
		h = h.copy()
		h["doc"] += rank * 5
		parts.append((h, tg, ds))
	want = merge_hits(parts, k)
	for rank in range(world):
		got = np.load(os.path.join(str(tmp_path), f"r{rank}.npz"))
		np.testing.assert_array_equal(got["hits"], want[0])
		np.testing.assert_array_equal(got["target"], want[1])      # int16 targets survive the float32 transport bit-exactly
		np.testing.assert_array_equal(got["distance"], want[2])
	for i, tt in enumerate((t, 3, t)):
		parts = []
		for rank, n_local in enumerate([7, 3]):
			h, tg, ds = _rank_rows(rank, n_local, tt, k_seed=10 * i)
			h = h.copy()
			h["doc"] += rank * 5
			parts.append((h, tg, ds))
		want_i = merge_hits(parts, k)
		for rank in range(world):
			got = np.load(os.path.join(str(tmp_path), f"m{rank}.npz"))
			np.testing.assert_array_equal(got[f"hits{i}"], want_i[0])
			np.testing.assert_array_equal(got[f"target{i}"], want_i[1])
			np.testing.assert_array_equal(got[f"distance{i}"], want_i[2])
	# total order: score desc, doc desc, slice desc
	h = want[0]
	for a, b in zip(h[:-1], h[1:]):
		assert (a["score"], a["doc"], a["slice"]) > (b["score"], b["doc"], b["slice"])


def test_merge_empty_and_shards():
	e = merge_hits([], 5)
	assert len(e[0]) == 0
	h, tg, ds = _rank_rows(0, 4, 3)
	m = merge_hits([(h[:0], tg[:0], ds[:0]), (h, tg, ds)], 10)
	assert len(m[0]) == 4
	# contiguous, covering, balanced by weight
	w = np.array([10, 1, 1, 1, 1, 10, 1, 1, 1, 1], dtype=np.float64)
	bounds = [shard_documents(len(w), 2, r, w) for r in range(2)]
	assert bounds[0][0] == 0 and bounds[0][1] == bounds[1][0] and bounds[1][1] == len(w)
	assert abs(w[:bounds[0][1]].sum() - w.sum() / 2) <= 10
	for world in (1, 3, 8, 16):
		b = [shard_documents(5, world, r) for r in range(world)]
		assert b[0][0] == 0 and b[-1][1] == 5
		assert all(x[1] == y[0] for x, y in zip(b[:-1], b[1:]))


def test_gather_without_process_group():
	"""No process group (one GPU): gather_topk / gather_topk_many only make the document ids global and keep the order."""
	import vectorian_b200.distributed as vd
	assert not dist.is_initialized()
	h, tg, ds = _rank_rows(0, 6, 4)
	g = gather_topk(h, tg, ds, 4, doc_offset=7)
	want = merge_hits([(np.array([(s, m, d + 7, sl) for s, m, d, sl in h.tolist()], dtype=HIT_DTYPE), tg, ds)], 4)
	np.testing.assert_array_equal(g[0], want[0])
	np.testing.assert_array_equal(g[1], want[1])
	k = 8
	hits = np.zeros((2, k), dtype=HIT_DTYPE)
	target = np.full((2, k, 4), -1, dtype=np.int16)
	distance = np.ones((2, k, 4), dtype=np.float32)
	rows = [_rank_rows(0, 6, 4), _rank_rows(1, 3, 4)]
	for i, (hh, tt, dd) in enumerate(rows):
		hits[i, :len(hh)], target[i, :len(hh)], distance[i, :len(hh)] = hh, tt, dd
	a = gather_topk_many(vd.BatchRows(hits, target, distance, [6, 3], [4, 4]), k, doc_offset=2)
	b = gather_topk_many(rows, k, doc_offset=2)
	for x, y in zip(a, b):
		for u, v in zip(x, y):
			np.testing.assert_array_equal(u, v)
