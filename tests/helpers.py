"""Shared helpers of the parity tests (oracle vs. CUDA path through the C ABI)."""

import numpy as np

from vectorian_b200 import _cabi, gaps, synthetic

LOCALITIES = [_cabi.VX_LOCAL, _cabi.VX_GLOBAL, _cabi.VX_SEMIGLOBAL]
LOCALITY_NAMES = {0: "local", 1: "global", 2: "semiglobal"}

GAPS = {
	"none": None,
	"linear0.2": gaps.LinearGapCost(0.2),
	"linear_st": {"s": gaps.LinearGapCost(0.1), "t": gaps.LinearGapCost(0.35)},
	"affine": gaps.AffineGapCost(u=0.3, v=0.1),
	"affine_st": {"s": gaps.AffineGapCost(0.5, 0.05), "t": gaps.AffineGapCost(u=0.2, v=0.2)},
	"smooth5": gaps.smooth_gap_cost(5),
	"log": gaps.LogarithmicGapCost(0.2, 0.1),
	"const": gaps.ConstantGapCost(0.25),
	"mixed": {"s": gaps.LinearGapCost(0.2), "t": gaps.smooth_gap_cost(3)},
}


def align_params(locality, gap, max_len_s, max_len_t):
	r = gaps.resolve_gap_costs(gap, max_len_s, max_len_t)
	kind = r.pop("gap_kind")
	return _cabi.make_align_params(locality, kind, **r)


def small_problem(n_sentences=3000, vocab=1500, dim=24, t=8, seed=7, with_pos=False, clustered=True):
	"""A corpus + embedding + query small enough for the oracle to finish in well under a second."""
	corpus = synthetic.make_corpus(n_sentences, vocab, seed=seed, sentences_per_doc=97, with_pos=with_pos)
	emb = synthetic.make_embedding(vocab, dim, seed=seed + 1, clustered=clustered, n_centroids=60)
	q_id = synthetic.make_query(t, vocab, seed=seed + 2)
	q_vec = emb[q_id].copy()
	return corpus, emb, q_id, q_vec


def assert_hits_match(gpu, ref, all_scores, win_doc, win_slice, rel_tol=0.0):
	"""Top-k parity modulo ties.

	gpu / ref = (hits, target, dist).  `all_scores` are the oracle's scores of every window
	(window order) with win_doc / win_slice giving each window's (doc, slice).  Rules
	(BASELINE.json north_star): the sorted score lists are identical; every
	(doc, slice) whose score is strictly above the k-th score must appear in both with an
	identical traceback; entries tied with the k-th score only have to be windows that really
	have that score.
	"""
	g_hits, g_target, g_dist = gpu
	r_hits, r_target, r_dist = ref
	assert len(g_hits) == len(r_hits), (len(g_hits), len(r_hits))
	if len(g_hits) == 0:
		return
	if rel_tol == 0.0:
		np.testing.assert_array_equal(g_hits["score"], r_hits["score"])
	else:
		np.testing.assert_allclose(g_hits["score"], r_hits["score"], rtol=rel_tol, atol=0)
	# descending order
	assert np.all(np.diff(g_hits["score"]) <= 0)
	kth = r_hits["score"][-1]
	score_of = {(int(d), int(s)): float(x) for d, s, x in zip(win_doc, win_slice, all_scores)}
	r_index = {(int(h["doc"]), int(h["slice"])): i for i, h in enumerate(r_hits)}
	for i, h in enumerate(g_hits):
		key = (int(h["doc"]), int(h["slice"]))
		assert key in score_of
		if rel_tol == 0.0:
			assert np.float32(score_of[key]) == h["score"], (key, score_of[key], h["score"])
		if h["score"] > kth + abs(kth) * rel_tol:
			assert key in r_index, f"{key} (score {h['score']}) missing from the oracle's top-k"
			j = r_index[key]
			np.testing.assert_array_equal(g_target[i], r_target[j])
			if rel_tol == 0.0:
				np.testing.assert_array_equal(g_dist[i], r_dist[j])
				assert h["score_max"] == r_hits["score_max"][j]
			else:
				np.testing.assert_allclose(g_dist[i], r_dist[j], rtol=1e-4, atol=1e-5)
	# the total order inside the GPU result: score desc, doc desc, slice desc
	for a, b in zip(g_hits[:-1], g_hits[1:]):
		if a["score"] == b["score"]:
			assert (a["doc"], a["slice"]) > (b["doc"], b["slice"])
