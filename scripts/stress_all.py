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
	print("TOTAL differing repetitions:", total_bad)


if __name__ == "__main__":
	main()
