"""Sentence shards on the hardware: the device merge kernel, the communicator inside the C ABI, and -- where two GPUs
are visible -- N shards + NCCL gather + merge against one GPU holding the whole corpus (ResultSet::extend + best_n,
reference: vectorian/core/cpp/result_set.h:70-93, cpp/result_set.cpp:3-21, vectorian/index.py:544-555)."""

import os
import subprocess
import sys

import numpy as np
import pytest

from vectorian_b200 import _cabi, engine
from vectorian_b200._cabi import HIT_DTYPE
from vectorian_b200.distributed import merge_hits

from .helpers import GAPS, align_params, small_problem

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rank_rows(rng, rank, n, t, grid):
	"""n rows of one rank in result order; scores on a coarse grid so that ties within and across ranks occur."""
	hits = np.zeros((n,), dtype=HIT_DTYPE)
	hits["score"] = rng.integers(0, grid, size=n).astype(np.float32) / 16
	hits["score_max"] = t
	hits["doc"] = rank * 5 + rng.integers(0, 5, size=n)       # the documents of different ranks are disjoint
	hits["slice"] = rng.permutation(4000)[:n]
	order = np.lexsort((-hits["slice"].astype(np.int64), -hits["doc"].astype(np.int64), -hits["score"].astype(np.float64)))
	hits = hits[order]
	target = rng.integers(-1, 30000, size=(n, t)).astype(np.int16)
	dist = rng.random((n, t)).astype(np.float32)
	return hits, target, dist


@pytest.mark.parametrize("world,k,t,grid", [(1, 10, 4, 12), (2, 6, 5, 12), (3, 100, 12, 3), (8, 100, 12, 40), (8, 37, 1, 2), (64, 20, 3, 6)])
def test_device_merge_equals_host_merge(world, k, t, grid):
	rng = np.random.default_rng(world * 1000 + k)
	counts = [int(rng.integers(0, k + 1)) for _ in range(world)]
	counts[0] = k                      # one full list
	if world > 1:
		counts[1] = 0                  # one empty shard
	parts = [_rank_rows(rng, r, n, t, grid) for r, n in enumerate(counts)]
	want = merge_hits(parts, k)
	got = engine.merge_blocks(parts, t, k)
	for name in ("score", "score_max", "doc", "slice"):
		np.testing.assert_array_equal(got[0][name], want[0][name], err_msg=name)
	np.testing.assert_array_equal(got[1], want[1])
	np.testing.assert_array_equal(got[2], want[2])


def test_single_rank_communicator_is_transparent():
	"""A world of one: vx_search / vx_search_batch through all-gather + merge return the plain result with the shard's
	document offset added (exercises the NCCL binding and the gather path on a 1-GPU box)."""
	corpus, emb, q_id, q_vec = small_problem(n_sentences=5000, t=12, seed=11)
	dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
	de = engine.DeviceEmbedding(emb)
	sim = engine.DeviceSim.static(de, q_id, q_vec)
	comm = engine.Comm(engine.Comm.unique_id(), 1, 0)
	try:
		for locality, gap, k in ((_cabi.VX_LOCAL, "smooth5", 100), (_cabi.VX_GLOBAL, "linear0.2", 30), (_cabi.VX_SEMIGLOBAL, "affine", 64)):
			ap, keep = align_params(locality, GAPS[gap], dc.max_len, sim.t)
			qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
			want = dc.search(sim, ap, qp)
			dc.set_shard(comm, 1000)
			got = dc.search(sim, ap, qp)
			assert dc.last_timing()["gather_merge_ms"] > 0.0
			dc.set_shard(None, 0)
			assert len(got[0]) == len(want[0]) == k
			np.testing.assert_array_equal(got[0]["doc"], want[0]["doc"] + 1000)
			for name in ("score", "score_max", "slice"):
				np.testing.assert_array_equal(got[0][name], want[0][name])
			np.testing.assert_array_equal(got[1], want[1])
			np.testing.assert_array_equal(got[2], want[2])
		# batched sweep
		aps, keeps = [], []
		for locality in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
			for gap in ("none", "linear0.2"):
				ap, keep = align_params(locality, GAPS[gap], dc.max_len, sim.t)
				aps.append(ap)
				keeps.append(keep)
		qp, _ = _cabi.make_query_params(max_matches=25, min_score=0.0)
		want = dc.search_batch([sim], aps, qp)
		dc.set_shard(comm, 7)
		got = dc.search_batch([sim], aps, qp)
		dc.set_shard(None, 0)
		for w, g in zip(want[0], got[0]):
			np.testing.assert_array_equal(g[0]["doc"], w[0]["doc"] + 7)
			np.testing.assert_array_equal(g[0]["score"], w[0]["score"])
			np.testing.assert_array_equal(g[1], w[1])
			np.testing.assert_array_equal(g[2], w[2])
	finally:
		dc.set_shard(None, 0)
		comm.close()
		sim.close()
		de.close()
		dc.close()


def test_device_resident_corpus_gets_the_bounds_check():
	"""vx_corpus_create_dev scans the ids on the device: a similarity matrix with too few rows is rejected instead of
	being gathered out of bounds, negative ids are refused."""
	import torch
	corpus, emb, q_id, q_vec = small_problem(n_sentences=500, vocab=300, t=4, seed=3)
	tok = torch.from_numpy(corpus["token_id"]).cuda()
	dc = engine.DeviceCorpus(tok, corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"], on_device=True)
	small = engine.DeviceSim.upload(np.zeros((int(corpus["token_id"].max()), 4), dtype=np.float32))   # one row short
	ap, keep = align_params(_cabi.VX_LOCAL, GAPS["linear0.2"], dc.max_len, 4)
	qp, _ = _cabi.make_query_params(max_matches=5, min_score=0.0)
	with pytest.raises(_cabi.VxError, match="fewer rows"):
		dc.search(small, ap, qp)
	bad = tok.clone()
	bad[17] = -3
	with pytest.raises(_cabi.VxError, match="negative token id"):
		engine.DeviceCorpus(bad, corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"], on_device=True)
	small.close()
	dc.close()


def test_two_ranks_equal_one_gpu():
	"""N shards + NCCL merge == one GPU, whole corpus -- on the hardware, when at least two GPUs are visible."""
	import torch
	n = torch.cuda.device_count()
	if n < 2:
		pytest.skip("needs at least two GPUs")
	world = 4 if n >= 4 else 2
	env = dict(os.environ, MASTER_ADDR="127.0.0.1")
	r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
		"--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "scripts", "check_sharded.py")],
		capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
	assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
	assert "sharded == unsharded on every rank: True" in r.stdout
