"""BASELINE.json's configurations at FULL size (per GPU), checked through size-independent properties.

The oracle cannot score 50 M sentences in seconds, so each test
  * runs the CUDA path (C ABI) over the whole corpus,
  * checks the top-k against the score of EVERY window (`vx_score_all`): the hits are exactly the windows above
    the k-th score, ordered by (score desc, doc desc, slice desc)  -- ResultSet semantics, cpp/result_set.h:17-153,
  * reads the tokens of the k winners and of a few thousand random windows back from the device, runs the ORACLE
    on that sample with the same similarity matrix and demands bit-identical scores (token / index / DP work is
    bit-exact) and identical tracebacks for the winners.
For the contextual config the similarity is formed on the device, so the bar is the north star's 1e-5.
"""

import numpy as np
import pytest

from vectorian_b200 import _cabi, gaps, synthetic
from tests.helpers import align_params

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
	from vectorian_b200 import engine as e
	_cabi.lib()
	return e


def sample_corpus(token_dev, lens, idx):
	"""Tokens of the sentences `idx` (window_size 1: window i is sentence i) as a one-document host corpus."""
	import torch
	ends = np.cumsum(lens.astype(np.int64))
	starts = ends - lens
	ln = lens[idx].astype(np.int64)
	st = torch.from_numpy(starts[idx]).to(token_dev.device)
	lt = torch.from_numpy(ln).to(token_dev.device)
	first = torch.cumsum(lt, 0) - lt
	pos = torch.arange(int(ln.sum()), device=token_dev.device) - torch.repeat_interleave(first, lt) + torch.repeat_interleave(st, lt)
	sub_end = np.cumsum(ln)
	return dict(token_pos_index=pos, token_id=token_dev[pos].cpu().numpy(),
		doc_tok_off=np.array([0, int(ln.sum())], dtype=np.int64),
		span_start=(sub_end - ln).astype(np.int32), span_end=sub_end.astype(np.int32),
		doc_span_off=np.array([0, len(idx)], dtype=np.int64))


def check_topk_against_all(hits, g_all, win_doc, win_slice, k):
	"""The hits are the k best windows of g_all under (score desc, doc desc, slice desc); ties at the k-th place
	may be resolved either way only if the score is equal."""
	sc = hits["score"]
	assert np.all(np.diff(sc) <= 0)
	for a, b in zip(hits[:-1], hits[1:]):
		if a["score"] == b["score"]:
			assert (a["doc"], a["slice"]) > (b["doc"], b["slice"])
	kth = sc[-1]
	above = np.flatnonzero(g_all > kth)
	key = {(int(win_doc[i]), int(win_slice[i])) for i in above}
	got = {(int(h["doc"]), int(h["slice"])) for h in hits if h["score"] > kth}
	assert key == got
	# the k-th score is the k-th largest of all scores
	part = np.partition(g_all, len(g_all) - len(hits))[len(g_all) - len(hits):]
	np.testing.assert_array_equal(np.sort(part)[::-1], sc)


def run_static_full(engine, oracle, n_sent, vocab, dim, t, locality, gap, k=100, n_sample=3000):
	import torch
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, device="cuda:0")
	lens = tables["lens"]
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	assert dc.n_windows == n_sent
	emb = synthetic.make_embedding(vocab, dim, clustered=True)
	de = engine.DeviceEmbedding(emb)
	q = synthetic.make_query(t, vocab, seed=99)
	ap, keep = align_params(locality, gap, dc.max_len, t)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	sim = engine.DeviceSim.static(de, q, emb[q])
	s = sim.download()
	# the similarity matrix itself: 1e-5 of the oracle's (north star tolerance; BLAS-free summation order differs)
	np.testing.assert_allclose(s, oracle.sim_static(emb, q, emb[q]), rtol=0, atol=1e-5)
	hits, target, dist = dc.search(sim, ap, qp)
	sim2 = engine.DeviceSim.static(de, q, emb[q])
	g_all = dc.score_all(sim2, ap, qp)
	win_doc = np.repeat(np.arange(tables["n_doc"], dtype=np.int32), np.diff(tables["doc_span_off"]))
	win_slice = (np.arange(n_sent, dtype=np.int64) - tables["doc_span_off"][win_doc]).astype(np.int32)
	assert len(hits) == k
	check_topk_against_all(hits, g_all, win_doc, win_slice, k)

	# oracle on the winners + a random sample, same S: bit-exact scores, identical tracebacks
	rng = np.random.default_rng(3)
	win_idx = (tables["doc_span_off"][hits["doc"]] + hits["slice"]).astype(np.int64)
	idx = np.concatenate([win_idx, rng.integers(0, n_sent, size=n_sample), [0, n_sent - 1]])
	sub = sample_corpus(tok, lens, idx)
	oc = oracle.OracleCorpus(sub["token_id"], sub["doc_tok_off"], sub["span_start"], sub["span_end"], sub["doc_span_off"])
	o_all = oc.score_all(s, ap, qp, n_threads=4)
	np.testing.assert_array_equal(o_all, g_all[idx])
	qp_all, _ = _cabi.make_query_params(max_matches=len(idx), min_score=-1e30)
	o_hits, o_target, o_dist = oc.search(s, ap, qp_all, n_threads=4)
	by_slice = {int(h["slice"]): i for i, h in enumerate(o_hits)}
	for i in range(k):
		j = by_slice[i]   # winner i is sentence i of the sample document
		assert o_hits["score"][j] == hits["score"][i]
		assert o_hits["score_max"][j] == hits["score_max"][i]
		np.testing.assert_array_equal(o_target[j], target[i])
		np.testing.assert_array_equal(o_dist[j], dist[i])
	sim.close()
	sim2.close()
	dc.close()
	de.close()
	del tok
	torch.cuda.empty_cache()


def test_c3_wsb_50m_sentences(engine, oracle):
	"""BASELINE config 3 at its per-GPU size: WSB Local, smooth_gap_cost(5), T = 12, 50 M sentences, top-100."""
	run_static_full(engine, oracle, 50_000_000, 50_000, 300, 12, _cabi.VX_LOCAL, gaps.smooth_gap_cost(5))


def test_c2_sw_1m_sentences(engine, oracle):
	"""BASELINE config 2: Smith-Waterman Local, LinearGapCost(0.2), T = 8, 1 M sentences, V = 200 k x 300, top-100."""
	run_static_full(engine, oracle, 1_000_000, 200_000, 300, 8, _cabi.VX_LOCAL, gaps.LinearGapCost(0.2))


def test_c5_settings_10m_sentences(engine, oracle):
	"""BASELINE config 5 at its size (10 M sentences, T = 12): the batched sweep returns for every (query, setting)
	pair what a single search returns, and single searches are pinned to the oracle by the sample check."""
	import torch
	n_sent, vocab, t = 10_000_000, 50_000, 12
	run_static_full(engine, oracle, n_sent, vocab, 300, t, _cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), n_sample=1000)
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, device="cuda:0")
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	emb = synthetic.make_embedding(vocab, 300, clustered=True)
	de = engine.DeviceEmbedding(emb)
	aps, keeps = [], []
	for loc in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
		for u in (0.0, 0.1, 0.2, 0.5):
			ap, keep = align_params(loc, gaps.LinearGapCost(u), dc.max_len, t)
			aps.append(ap)
			keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
	queries = [synthetic.make_query(t, vocab, seed=200 + i) for i in range(3)]
	sims = [engine.DeviceSim.static(de, q, emb[q]) for q in queries]
	out = dc.search_batch(sims, aps, qp)
	for qi, sim in enumerate(sims):
		for ai, ap in enumerate(aps):
			hits, target, dist = dc.search(sim, ap, qp)
			b_hits, b_target, b_dist = out[qi][ai]
			for name in ("score", "score_max", "doc", "slice"):
				np.testing.assert_array_equal(hits[name], b_hits[name])
			np.testing.assert_array_equal(target, b_target)
			np.testing.assert_array_equal(dist, b_dist)
	for sim in sims:
		sim.close()
	dc.close()
	de.close()
	del tok
	torch.cuda.empty_cache()


def test_c4_contextual_full_shard(engine, oracle):
	"""BASELINE config 4, one GPU's shard of the 8-GPU run: 625 k sentences, 15.6 M tokens x 128 fp32 (8 GB of
	vectors: byte offsets beyond 32 bits), NW Global LinearGapCost(0.1), T = 8.  Scores within 1e-5 of the oracle on
	the winners + a random sample (vectors read back from the device); top-k consistent with the scores of all windows."""
	import torch
	n_sent, dim, t, k = 625_000, 128, 8, 100
	tok, tables = synthetic.make_corpus_device(n_sent, 50_000, device="cuda:0")
	lens = tables["lens"]
	n_tok = int(tok.numel())
	g = torch.Generator(device="cuda:0")
	g.manual_seed(5)
	# clustered vectors so that alignments are non-trivial: 4096 centroids + noise, generated on the device
	centroids = torch.randn((4096, dim), generator=g, device="cuda:0")
	x = torch.empty((n_tok, dim), dtype=torch.float32, device="cuda:0")
	chunk = 1 << 22
	for i in range(0, n_tok, chunk):
		n = min(chunk, n_tok - i)
		x[i:i + n] = centroids[(tok[i:i + n] % 4096).long()] + 0.5 * torch.randn((n, dim), generator=g, device="cuda:0")
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	dc.set_contextual(x)
	q_tok = torch.randint(0, n_tok, (t,), generator=g, device="cuda:0")
	q = (x[q_tok] + 0.1 * torch.randn((t, dim), generator=g, device="cuda:0")).cpu().numpy()
	ap, keep = align_params(_cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), dc.max_len, t)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=-1e30)
	sim = engine.DeviceSim.contextual(q)
	hits, target, dist = dc.search(sim, ap, qp)
	assert dc.last_launches() <= 8
	sim2 = engine.DeviceSim.contextual(q)
	g_all = dc.score_all(sim2, ap, qp)
	win_doc = np.repeat(np.arange(tables["n_doc"], dtype=np.int32), np.diff(tables["doc_span_off"]))
	win_slice = (np.arange(n_sent, dtype=np.int64) - tables["doc_span_off"][win_doc]).astype(np.int32)
	assert len(hits) == k
	check_topk_against_all(hits, g_all, win_doc, win_slice, k)

	rng = np.random.default_rng(4)
	win_idx = (tables["doc_span_off"][hits["doc"]] + hits["slice"]).astype(np.int64)
	idx = np.concatenate([win_idx, rng.integers(0, n_sent, size=2000), [0, n_sent - 1]])
	sub = sample_corpus(tok, lens, idx)
	x_sub = x[sub["token_pos_index"]].cpu().numpy()
	s = oracle.sim_contextual(x_sub, q)
	oc = oracle.OracleCorpus(sub["token_id"], sub["doc_tok_off"], sub["span_start"], sub["span_end"], sub["doc_span_off"])
	o_all = oc.score_all(s, ap, qp, n_threads=4, contextual=True)
	np.testing.assert_allclose(g_all[idx], o_all, rtol=1e-5, atol=2e-6)
	qp_all, _ = _cabi.make_query_params(max_matches=len(idx), min_score=-1e30)
	o_hits, o_target, o_dist = oc.search(s, ap, qp_all, n_threads=4, contextual=True)
	by_slice = {int(h["slice"]): i for i, h in enumerate(o_hits)}
	same_tb = 0
	for i in range(k):
		j = by_slice[i]
		np.testing.assert_allclose(o_hits["score"][j], hits["score"][i], rtol=1e-5, atol=2e-6)
		if np.array_equal(o_target[j], target[i]):
			same_tb += 1
			continue
		# A differing traceback must be a genuine tie within the tolerance: the GPU's alignment, scored with the ORACLE's
		# similarity matrix (Global, linear gap 0.1: the matched similarities minus 0.1 for every unmatched token of the
		# sentence and of the query), reaches the oracle's optimum within 1e-5.
		lo, n_s = int(sub["span_start"][i]), int(sub["span_end"][i] - sub["span_start"][i])   # rows of s: the sampled tokens, back to back
		matched = [(int(u), v) for v, u in enumerate(target[i]) if u >= 0]
		assert all(0 <= u < n_s for u, _ in matched) and all(a[0] < b[0] for a, b in zip(matched[:-1], matched[1:]))
		raw = sum(float(s[lo + u, v]) for u, v in matched) - 0.1 * (n_s - len(matched)) - 0.1 * (t - len(matched))
		np.testing.assert_allclose(raw / t, o_hits["score"][j], rtol=1e-5, atol=2e-6)
	# ... and they are rare on random vectors
	assert same_tb >= k - 2
	sim.close()
	sim2.close()
	dc.close()
	del x, tok
	torch.cuda.empty_cache()
