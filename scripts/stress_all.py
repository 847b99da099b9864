#!/usr/bin/env python
"""Run-to-run determinism of every search path: the scores of all windows, the hits and the tracebacks must be bit-identical
on every repetition (race hunting; usage: python scripts/stress_all.py [reps] [sentences])."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vectorian_b200 import _cabi, engine, gaps, synthetic  # noqa: E402


def main():
	reps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
	n_sent = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
	vocab, dim = 50_000, 300
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, device="cuda:0")
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	emb = synthetic.make_embedding(vocab, dim)
	de = engine.DeviceEmbedding(emb)
	cases = [
		("wsb local T=12", _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12),
		("wsb global T=7", _cabi.VX_GLOBAL, gaps.smooth_gap_cost(3), 7),
		("wsb semiglobal T=16", _cabi.VX_SEMIGLOBAL, gaps.LogarithmicGapCost(0.2, 0.1), 16),
		("linear local T=8", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8),
		("linear global T=12", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), 12),
		("affine local T=8", _cabi.VX_LOCAL, gaps.AffineGapCost(u=0.3, v=0.1), 8),
		("affine semiglobal T=5", _cabi.VX_SEMIGLOBAL, gaps.AffineGapCost(u=0.2, v=0.2), 5),
	]
	total_bad = 0
	for name, loc, gap, t in cases:
		r = gaps.resolve_gap_costs(gap, dc.max_len, t)
		kind = r.pop("gap_kind")
		ap, keep = _cabi.make_align_params(loc, kind, **r)
		qp, _ = _cabi.make_query_params(max_matches=100, min_score=-1e30 if loc != _cabi.VX_LOCAL else 0.0)
		q = synthetic.make_query(t, vocab, seed=7 + t)
		first = None
		bad = 0
		for rep in range(reps):
			sim = engine.DeviceSim.static(de, q, emb[q])
			s = sim.download() if rep % 10 == 0 else None
			hits, target, dist = dc.search(sim, ap, qp)
			cur = [hits.copy(), target.copy(), dist.copy()]
			if rep % 10 == 0:
				cur.append(dc.score_all(sim, ap, qp))
				cur.append(s)
			sim.close()
			if first is None:
				first = cur
				continue
			ok = all(np.array_equal(first[0][f], cur[0][f]) for f in ("score", "score_max", "doc", "slice"))
			ok = ok and np.array_equal(first[1], cur[1]) and np.array_equal(first[2], cur[2])
			if len(cur) > 3:
				ok = ok and np.array_equal(first[3], cur[3]) and np.array_equal(first[4], cur[4])
			if not ok:
				bad += 1
				if bad <= 3:
					d = np.flatnonzero(first[0]["score"] != cur[0]["score"])
					print(f"  {name}: rep {rep} differs; hit scores differing: {len(d)}", flush=True)
					if len(cur) > 3:
						da = np.flatnonzero(first[3] != cur[3])
						print(f"    all-window scores differing: {len(da)} {da[:6]}; S differing: {int((first[4] != cur[4]).sum())}", flush=True)
		print(f"{name}: {bad} of {reps - 1} repetitions differ", flush=True)
		total_bad += bad

	# batched sweep (fused multi-setting kernel)
	aps, keeps = [], []
	for loc in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
		for u in (0.0, 0.1, 0.2, 0.5):
			r = gaps.resolve_gap_costs(gaps.LinearGapCost(u), dc.max_len, 12)
			kind = r.pop("gap_kind")
			ap, keep = _cabi.make_align_params(loc, kind, **r)
			aps.append(ap)
			keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
	queries = [synthetic.make_query(12, vocab, seed=300 + i) for i in range(4)]
	first, bad = None, 0
	for rep in range(max(reps // 4, 5)):
		sims = [engine.DeviceSim.static(de, q, emb[q]) for q in queries]
		out = dc.search_batch(sims, aps, qp)
		for sim in sims:
			sim.close()
		flat = [np.concatenate([o[0]["score"] for row in out for o in row]), np.concatenate([o[1].ravel() for row in out for o in row]),
			np.concatenate([o[0]["slice"] for row in out for o in row])]
		if first is None:
			first = flat
		elif not all(np.array_equal(a, b) for a, b in zip(first, flat)):
			bad += 1
	print(f"batched sweep: {bad} repetitions differ", flush=True)
	total_bad += bad

	# round 2: several queries per pass (score_linear_mq_kernel + multi-array pruned top-k)
	first, bad = None, 0
	queries = [synthetic.make_query(5 + i, vocab, seed=400 + i) for i in range(8)]
	for rep in range(max(reps // 4, 5)):
		sims = [engine.DeviceSim.static(de, q, emb[q]) for q in queries]
		out = dc.search_batch(sims, aps[:1] + aps[4:5], qp)
		for sim in sims:
			sim.close()
		flat = [np.concatenate([o[0]["score"] for row in out for o in row]), np.concatenate([o[1].ravel() for row in out for o in row]),
			np.concatenate([o[0]["slice"] for row in out for o in row])]
		if first is None:
			first = flat
		elif not all(np.array_equal(a, b) for a, b in zip(first, flat)):
			bad += 1
	print(f"multi-query batch: {bad} repetitions differ", flush=True)
	total_bad += bad

	# round 2: tag weighting on the fast kernels, a 24-token query, windows of three sentences (WSB classes up to 160 rows)
	import torch
	g = torch.Generator(device="cuda:0")
	g.manual_seed(3)
	n_small = min(n_sent, 300_000)
	tok2, tables2 = synthetic.make_corpus_device(n_small, vocab, seed=77, device="cuda:0")
	pos2 = torch.randint(0, 17, (int(tok2.numel()),), generator=g, device="cuda:0", dtype=torch.int8)
	extra = [
		("tag-weighted wsb local T=12", dict(), _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, True),
		("tag-weighted linear global T=8", dict(), _cabi.VX_GLOBAL, gaps.LinearGapCost(0.2), 8, True),
		("linear local T=24", dict(), _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 24, False),
		("wsb local T=24 (scratch kernel)", dict(), _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 24, False),
		("linear local T=40 (vx_align_long)", dict(), _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 40, False),
		("linear semiglobal T=64", dict(), _cabi.VX_SEMIGLOBAL, gaps.LinearGapCost(0.1), 64, False),
		("affine global T=48", dict(), _cabi.VX_GLOBAL, gaps.AffineGapCost(u=0.3, v=0.1), 48, False),
		("wsb local T=12, windows of 3 sentences", dict(window_size=3, window_step=1), _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, False),
		("wsb semiglobal T=16, windows of 3 sentences", dict(window_size=3, window_step=2), _cabi.VX_SEMIGLOBAL, gaps.smooth_gap_cost(3), 16, False),
	]
	ehat = emb / np.linalg.norm(emb, axis=1, keepdims=True)
	for name, wkw, loc, gap, t, tagged in extra:
		dc2 = engine.DeviceCorpus(tok2, tables2["doc_tok_off"], tables2["span_start"], tables2["span_end"], tables2["doc_span_off"],
			on_device=True, token_pos=pos2, **wkw)
		r = gaps.resolve_gap_costs(gap, dc2.max_len, t)
		kind = r.pop("gap_kind")
		ap, keep = _cabi.make_align_params(loc, kind, **r)
		if tagged:
			qp2, keep_q = _cabi.make_query_params(max_matches=100, min_score=0.0, t_weights=(np.arange(t) % 4 + 1).astype(np.float32) / 2,
				t_pos=(np.arange(t) % 17).astype(np.int8), pos_mismatch_penalty=0.3, similarity_threshold=0.1)
		else:
			qp2, keep_q = _cabi.make_query_params(max_matches=100, min_score=0.0 if loc == _cabi.VX_LOCAL else -1e30)
		q = synthetic.make_query(t, vocab, seed=500 + t)
		s_host = np.clip(ehat @ ehat[q].T, 0.0, 1.0).astype(np.float32)
		first, bad = None, 0
		for rep in range(max(reps // 2, 5)):
			sim = engine.DeviceSim.upload(s_host)
			hits, target, dist = dc2.search(sim, ap, qp2)
			cur = [hits.copy(), target.copy(), dist.copy()]
			if rep % 10 == 0:
				cur.append(dc2.score_all(sim, ap, qp2))
			sim.close()
			if first is None:
				first = cur
				continue
			ok = all(np.array_equal(first[0][f], cur[0][f]) for f in ("score", "doc", "slice")) and np.array_equal(first[1], cur[1])
			if len(cur) > 3:
				ok = ok and np.array_equal(first[3], cur[3])
			bad += 0 if ok else 1
		print(f"{name}: {bad} repetitions differ", flush=True)
		total_bad += bad
		dc2.close()
	print("TOTAL differing repetitions:", total_bad)


if __name__ == "__main__":
	main()
