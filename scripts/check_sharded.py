#!/usr/bin/env python
"""N shards + NCCL all-gather + device merge == ONE GPU holding the whole corpus (ResultSet::extend + best_n semantics,
reference: vectorian/index.py:544-555, vectorian/core/cpp/result_set.h:70-93, cpp/result_set.cpp:3-21).

Every rank packs (a) its contiguous range of documents, attached to the communicator, and (b) the WHOLE corpus,
unattached, on its own GPU; vx_search / vx_search_batch over (a) must return exactly what they return over (b):
scores, document ids, slice ids, tracebacks and distances, in order -- including the order inside score ties.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/check_sharded.py [sentences]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

from vectorian_b200 import _cabi, engine, gaps, synthetic  # noqa: E402
from vectorian_b200.distributed import device_comm, shard_documents  # noqa: E402


def params(locality, gap, max_len, t):
	r = gaps.resolve_gap_costs(gap, max_len, t)
	kind = r.pop("gap_kind")
	return _cabi.make_align_params(locality, kind, **r)


def tiled_corpus(base_sentences, reps, vocab, sentences_per_doc=53):
	"""`reps` copies of one small corpus back to back: every sentence -- hence every score -- occurs `reps` times, in
	different documents and (sharded) on different ranks, so the top-k is full of exact ties."""
	base = synthetic.make_corpus(base_sentences, vocab, seed=5, sentences_per_doc=sentences_per_doc)
	lens = np.tile(base["lens"], reps)
	token_id = np.tile(base["token_id"], reps)
	n = len(lens)
	n_doc = (n + sentences_per_doc - 1) // sentences_per_doc
	doc_span_off = np.minimum(np.arange(n_doc + 1, dtype=np.int64) * sentences_per_doc, n)
	ends = np.cumsum(lens.astype(np.int64))
	starts = ends - lens
	doc_tok_off = np.concatenate([starts[doc_span_off[:-1]], [len(token_id)]]).astype(np.int64)
	doc_of = np.repeat(np.arange(n_doc), np.diff(doc_span_off))
	return dict(token_id=token_id, doc_tok_off=doc_tok_off, span_start=(starts - doc_tok_off[doc_of]).astype(np.int32),
		span_end=(ends - doc_tok_off[doc_of]).astype(np.int32), doc_span_off=doc_span_off, lens=lens, n_doc=n_doc)


def check_corpus(vocab, dim, cent, n_sent, rank, world, tiled=False):
	corpus = tiled_corpus(n_sent // 30, 30, vocab) if tiled else synthetic.make_corpus(n_sent, vocab, seed=5, sentences_per_doc=53)
	emb = synthetic.make_embedding(vocab, dim, seed=6, n_centroids=cent)   # few centroids: many exact score ties
	whole = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
	n_doc = corpus["n_doc"]
	lo, hi = shard_documents(n_doc, world, rank, np.diff(corpus["doc_tok_off"]))
	if world > 2 and rank == world - 1:
		lo = hi            # an EMPTY shard must still take part in the collectives
	elif world > 2 and rank == world - 2:
		hi = n_doc
	t0, t1 = corpus["doc_tok_off"][lo], corpus["doc_tok_off"][hi]
	s0, s1 = corpus["doc_span_off"][lo], corpus["doc_span_off"][hi]
	shard = engine.DeviceCorpus(corpus["token_id"][t0:t1], corpus["doc_tok_off"][lo:hi + 1] - t0, corpus["span_start"][s0:s1],
		corpus["span_end"][s0:s1], corpus["doc_span_off"][lo:hi + 1] - s0)
	shard.set_shard(device_comm(), lo)
	de = engine.DeviceEmbedding(emb)
	ok = True
	n_checked = 0
	max_len = whole.max_len
	settings = [(_cabi.VX_LOCAL, gaps.smooth_gap_cost(5)), (_cabi.VX_GLOBAL, gaps.LinearGapCost(0.2)),
		(_cabi.VX_SEMIGLOBAL, gaps.AffineGapCost(0.3, 0.1)), (_cabi.VX_LOCAL, gaps.LinearGapCost(0.0))]
	sims = []
	for qi, t in enumerate((12, 5, 8)):
		q = synthetic.make_query(t, vocab, seed=40 + qi)
		sims.append(engine.DeviceSim.static(de, q, emb[q].copy()))
	for k, min_score in ((100, 0.0), (7, 0.0), (100, 0.55), (50, -5.0)):
		qp, _ = _cabi.make_query_params(max_matches=k, min_score=min_score)
		for sim in sims:
			for loc, gap in settings:
				ap, keep = params(loc, gap, max_len, sim.t)
				want = whole.search(sim, ap, qp)
				got = shard.search(sim, ap, qp)
				same = len(want[0]) == len(got[0]) and all(np.array_equal(w, g) for w, g in zip(want, got))
				if not same and rank == 0:
					print("MISMATCH k=%d min=%g t=%d loc=%d: %d vs %d rows" % (k, min_score, sim.t, loc, len(want[0]), len(got[0])), flush=True)
				ok = ok and same
				n_checked += 1
	# the batched sweep: ONE gather + ONE merge launch for all pairs
	qp, _ = _cabi.make_query_params(max_matches=20, min_score=0.0)
	aps, keeps = [], []
	for loc in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
		for u in (0.0, 0.1, 0.2, 0.5):
			ap, keep = params(loc, gaps.LinearGapCost(u), max_len, 12)
			aps.append(ap)
			keeps.append(keep)
	ap, keep = params(_cabi.VX_LOCAL, gaps.smooth_gap_cost(5), max_len, 12)
	aps.append(ap)
	want = whole.search_batch(sims, aps, qp)
	got = shard.search_batch(sims, aps, qp)
	for wq, gq in zip(want, got):
		for w, g in zip(wq, gq):
			same = len(w[0]) == len(g[0]) and all(np.array_equal(x, y) for x, y in zip(w, g))
			ok = ok and same
			n_checked += 1
	ties = int(np.sum(np.diff(whole.search(sims[0], params(_cabi.VX_LOCAL, gaps.LinearGapCost(0.0), max_len, 12)[0],
		_cabi.make_query_params(max_matches=100, min_score=0.0)[0])[0]["score"]) == 0))
	ms = shard.last_timing()["gather_merge_ms"]
	for sim in sims:
		sim.close()
	de.close()
	shard.set_shard(None, 0)
	shard.close()
	whole.close()
	return ok, n_checked, ties, ms


def main():
	n_sent = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
	local_rank = int(os.environ.get("LOCAL_RANK", "0"))
	torch.cuda.set_device(local_rank)
	dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
	rank, world = dist.get_rank(), dist.get_world_size()
	ok = True
	n_checked = 0
	ties = 0
	last_ms = 0.0
	# a varied corpus, then 30 copies of a small one: every score occurs 30 times, spread over the ranks, so the top-k
	# and its order are decided by (doc desc, slice desc) across ranks
	for vocab, dim, cent, n_here, tiled in ((3000, 32, 20, n_sent, False), (3000, 32, 20, min(n_sent, 60000), True)):
		ok_c, n_c, ties_c, last_ms = check_corpus(vocab, dim, cent, n_here, rank, world, tiled)
		ok = ok and ok_c
		n_checked += n_c
		ties += ties_c
	flag = torch.tensor([1 if ok else 0], device="cuda")
	dist.all_reduce(flag, op=dist.ReduceOp.MIN)
	if rank == 0:
		print("sharded == unsharded on every rank: %s (%d searches, world %d, %d sentences, %d equal neighbours in two top-100 lists, gather+merge %.3f ms)" % (
			bool(flag.item()), n_checked, world, n_sent, ties, last_ms), flush=True)
	dist.destroy_process_group()
	sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
	main()
