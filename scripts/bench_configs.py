#!/usr/bin/env python
"""Secondary measurements: the other BASELINE.json configs on ONE GPU, per-phase device times and
roofline fractions (bench.py stays the C3 headline).  Usage: python scripts/bench_configs.py [c2] [c1] [c4] ...
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vectorian_b200 import _cabi, engine, gaps, synthetic  # noqa: E402

# Multi-GPU: launched with torchrun (one process per GPU) every rank holds 1 / WORLD of the sentences (its own seeded shard),
# runs the whole query on it, and the per-rank top-k are merged by one all-gather; times are the max over ranks.
RANK = int(os.environ.get("RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
DEV = "cuda:%d" % int(os.environ.get("LOCAL_RANK", "0"))


def dist_setup():
	if WORLD > 1:
		import torch
		import torch.distributed as dist
		torch.cuda.set_device(DEV)
		dist.init_process_group("nccl", device_id=torch.device(DEV))


def sync_all():
	import torch
	if WORLD > 1:
		import torch.distributed as dist
		dist.barrier()
	torch.cuda.synchronize()


def reduce_max(values):
	"""max over ranks of a dict of floats"""
	if WORLD == 1:
		return values
	import torch
	import torch.distributed as dist
	keys = sorted(values)
	x = torch.tensor([values[k] for k in keys], dtype=torch.float64, device=DEV)
	dist.all_reduce(x, op=dist.ReduceOp.MAX)
	return dict(zip(keys, x.tolist()))


COLLECT = None   # bench.py sets this to a list: results are collected instead of printed


def emit(out):
	if RANK == 0:
		out["n_gpus"] = WORLD
		if COLLECT is not None:
			COLLECT.append(out)
		else:
			print(json.dumps(out), flush=True)


def attach_shard(dc, n_doc):
	"""Multi-GPU: gather + merge of the ranks' result rows happen inside vx_search / vx_search_batch (csrc/vx_shard.cu)."""
	if WORLD > 1:
		from vectorian_b200.distributed import device_comm
		dc.set_shard(device_comm(), RANK * n_doc)


def peaks():
	try:
		return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
	except Exception:
		return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}


def run_static(name, locality, gap, t, n_sent, vocab, dim, k=100, steps=10, cand_per_cell=None, embedding=None, tag_weights=None):
	n_sent = n_sent // WORLD
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, seed=synthetic.CORPUS_SEED + RANK, device=DEV)
	pos = None
	if tag_weights is not None:
		import torch
		g = torch.Generator(device=DEV)
		g.manual_seed(7 + RANK)
		pos = torch.randint(0, 17, (int(tok.numel()),), generator=g, device=DEV, dtype=torch.int8)   # 17 universal POS tags
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True,
		token_pos=pos)
	attach_shard(dc, tables["n_doc"])
	del tok, pos
	emb = embedding if embedding is not None else synthetic.make_embedding(vocab, dim)
	de = engine.DeviceEmbedding(emb)
	r = gaps.resolve_gap_costs(gap, 64, t)
	kind = r.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r)
	if tag_weights is not None:
		qp, _keep_q = _cabi.make_query_params(max_matches=k, min_score=0.0, t_weights=np.asarray(tag_weights, np.float32),
			t_pos=(np.arange(t) % 17).astype(np.int8), pos_mismatch_penalty=0.3, similarity_threshold=0.1)
	else:
		qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	queries = [synthetic.make_query(t, vocab, seed=99 + i) for i in range(8)]
	tm = {"sim_ms": [], "dp_ms": [], "topk_ms": [], "traceback_ms": [], "wall_ms": []}
	for i in range(3 + steps):
		q = queries[i % 8]
		sync_all()
		t0 = time.perf_counter()
		sim = engine.DeviceSim.static(de, q, emb[q])
		hits, target, dist = dc.search(sim, ap, qp)
		t1 = time.perf_counter()
		if i >= 3:
			lt = dc.last_timing()
			for kk in ("dp_ms", "topk_ms", "traceback_ms"):
				tm[kk].append(lt[kk])
			tm["sim_ms"].append(sim.build_ms)
			tm["wall_ms"].append((t1 - t0) * 1e3)
		sim.close()
	p = peaks()
	med = reduce_max({kk: float(np.median(v)) for kk, v in tm.items()})
	dev_ms = med["sim_ms"] + med["dp_ms"] + med["topk_ms"] + med["traceback_ms"]
	lens = tables["lens"].astype(np.float64)
	cells = float((lens * t).sum()) * WORLD
	n_sent = n_sent * WORLD
	sms = _cabi.lib().vx_device_sm_count()
	issue_peak = sms * 128 * p.get("sm_max_mhz", 1965.0) * 1e6 * WORLD
	hbm_peak = p["hbm_gbs"] * WORLD
	tok_total = float(lens.sum()) * WORLD
	out = {
		"config": name, "sentences": n_sent, "T": t, "V": vocab, "d": dim, "n_hits": int(len(hits)),
		"ms": med, "device_ms": dev_ms, "sentences_per_s_device": n_sent / dev_ms * 1e3, "sentences_per_s_e2e": n_sent / med["wall_ms"] * 1e3,
		"gcups": cells / dev_ms / 1e6,
		"dp_hbm_frac": (4 * tok_total + 4 * n_sent) / (med["dp_ms"] * 1e-3) / 1e9 / hbm_peak,
		"dp_l2_gather_gbs": cells * 4 / (med["dp_ms"] * 1e-3) / 1e9,
		"sim_hbm_frac": (4.0 * vocab * dim + 4.0 * vocab * ((t + 3) // 4 * 4)) / (med["sim_ms"] * 1e-3) / 1e9 / p["hbm_gbs"],
	}
	if cand_per_cell:
		ops = 2 * cand_per_cell * cells
		out["dp_issue_frac"] = ops / (med["dp_ms"] * 1e-3) / issue_peak
	emit(out)
	dc.close()
	return out


def run_contextual(name, locality, gap, t, n_sent, dim, k=100, steps=10):
	import torch
	n_sent = n_sent // WORLD
	tok, tables = synthetic.make_corpus_device(n_sent, 50_000, seed=synthetic.CORPUS_SEED + RANK, device=DEV)
	n_tok = int(tok.numel())
	g = torch.Generator(device=DEV)
	g.manual_seed(5 + RANK)
	x = torch.empty((n_tok, dim), dtype=torch.float32, device=DEV)
	chunk = 1 << 22
	for i in range(0, n_tok, chunk):
		x[i:i + min(chunk, n_tok - i)].normal_(generator=g)
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	dc.set_contextual(x)
	attach_shard(dc, tables["n_doc"])
	rng = np.random.default_rng(5)
	qrows = x[torch.from_numpy(rng.integers(0, n_tok, size=(3 + steps, t))).to(DEV)].cpu().numpy()
	del x, tok
	torch.cuda.empty_cache()
	r = gaps.resolve_gap_costs(gap, 64, t)
	kind = r.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=-1e30)
	tm = {"dp_ms": [], "topk_ms": [], "traceback_ms": [], "wall_ms": []}
	for i in range(3 + steps):
		q = qrows[i] + 0.1 * rng.standard_normal((t, dim), dtype=np.float32)
		sync_all()
		t0 = time.perf_counter()
		sim = engine.DeviceSim.contextual(q)
		hits, target, dist = dc.search(sim, ap, qp)
		t1 = time.perf_counter()
		if i >= 3:
			lt = dc.last_timing()
			for kk in ("dp_ms", "topk_ms", "traceback_ms"):
				tm[kk].append(lt[kk])
			tm["wall_ms"].append((t1 - t0) * 1e3)
		sim.close()
	p = peaks()
	med = reduce_max({kk: float(np.median(v)) for kk, v in tm.items()})
	tp = (t + 3) // 4 * 4
	lens = tables["lens"].astype(np.float64)
	n_pad = float((np.ceil(lens / 4) * 4).sum()) * WORLD   # token positions incl. quad padding (approx.: 32-lane tiles)
	gemm_bytes = n_pad * (4.0 * dim + 4.0 * tp)
	n_all = n_sent * WORLD
	out = {"config": name, "sentences": n_all, "tokens": n_tok * WORLD, "T": t, "d": dim, "ms": med,
		"sentences_per_s_device": n_all / (med["dp_ms"] + med["topk_ms"] + med["traceback_ms"]) * 1e3,
		"sentences_per_s_e2e": n_all / med["wall_ms"] * 1e3,
		"note": "dp_ms = contextual similarity GEMM + DP kernels; vectors generated on the device (random normal)",
		"gemm_dp_hbm_frac_lower_bound": gemm_bytes / (med["dp_ms"] * 1e-3) / 1e9 / (p["hbm_gbs"] * WORLD), "top_score": float(hits["score"][0])}
	emit(out)
	dc.close()


def run_sweep(name, n_sent, t=12, vocab=50_000, dim=300, n_queries=64, k=100, embedding=None):
	"""C5: n_queries x 8 settings ({LinearGapCost(u), u in 0,.1,.2,.5} x {Local, Global}) without re-preprocessing."""
	n_sent = n_sent // WORLD
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, seed=synthetic.CORPUS_SEED + RANK, device=DEV)
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	attach_shard(dc, tables["n_doc"])
	del tok
	emb = embedding if embedding is not None else synthetic.make_embedding(vocab, dim)
	de = engine.DeviceEmbedding(emb)
	aps, keeps = [], []
	for loc in (_cabi.VX_LOCAL, _cabi.VX_GLOBAL):
		for u in (0.0, 0.1, 0.2, 0.5):
			r = gaps.resolve_gap_costs(gaps.LinearGapCost(u), 64, t)
			kind = r.pop("gap_kind")
			ap, keep = _cabi.make_align_params(loc, kind, **r)
			aps.append(ap)
			keeps.append(keep)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	queries = [synthetic.make_query(t, vocab, seed=99 + i) for i in range(n_queries)]
	res = {}
	for mode in ("single", "batch"):
		if mode == "single" and WORLD > 1:
			continue
		for rep in range(2):   # first repetition warms up
			sync_all()
			t0 = time.perf_counter()
			sims = [engine.DeviceSim.static(de, q, emb[q]) for q in queries]
			ta = time.perf_counter()
			if mode == "single":
				top = [[dc.search(sim, ap, qp)[0]["score"][0] for ap in aps] for sim in sims]
			else:
				out = dc.search_batch(sims, aps, qp)   # WORLD > 1: ONE all-gather + ONE merge launch for all pairs, inside the call
				tb = time.perf_counter()
				host = {"sims_ms": (ta - t0) * 1e3, "search_batch_ms": (tb - ta) * 1e3,
					"gather_merge_ms": dc.last_timing()["gather_merge_ms"]}
				top = [[out[q][a][0]["score"][0] for a in range(len(aps))] for q in range(len(sims))]
				dev_ms = dc.last_timing()["dp_ms"]
			t1 = time.perf_counter()
			for sim in sims:
				sim.close()
		res[mode] = {"wall_ms": (t1 - t0) * 1e3, "top": top}
	if WORLD == 1:
		assert res["single"]["top"] == res["batch"]["top"]
	pairs = n_queries * len(aps)
	m = reduce_max({"batch_wall_ms": res["batch"]["wall_ms"], "batch_device_ms": dev_ms})
	out = {"config": name, "sentences": n_sent * WORLD, "T": t, "queries": n_queries, "settings": len(aps),
		"single_calls_wall_ms": res["single"]["wall_ms"] if "single" in res else None,
		"batch_wall_ms": m["batch_wall_ms"], "batch_device_ms": m["batch_device_ms"],
		"ms_per_pair_batch": m["batch_wall_ms"] / pairs,
		"sentence_alignments_per_s_e2e": n_sent * WORLD * pairs / m["batch_wall_ms"] * 1e3,
		"launches": dc.last_launches(), "host_ms_rank0": host}
	emit(out)
	dc.close()


def run_windows(name, locality, gap, t, n_sent, window_size, window_step, vocab=50_000, dim=300, k=100, steps=5):
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, seed=synthetic.CORPUS_SEED, device=DEV)
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"],
		window_size=window_size, window_step=window_step, on_device=True)
	del tok
	emb = synthetic.make_embedding(vocab, dim)
	de = engine.DeviceEmbedding(emb)
	q = synthetic.make_query(t, vocab, seed=123)
	r = gaps.resolve_gap_costs(gap, dc.max_len, t)
	kind = r.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	tm = []
	for i in range(2 + steps):
		t0 = time.perf_counter()
		sim = engine.DeviceSim.static(de, q, emb[q])
		hits = dc.search(sim, ap, qp)[0]
		if i >= 2:
			tm.append((time.perf_counter() - t0) * 1e3)
		sim.close()
	emit({"config": name, "windows": dc.n_windows, "max_len": dc.max_len, "cells": dc.n_cells, "T": t, "wall_ms": float(np.median(tm)), "ms": dc.last_timing(),
		"windows_per_s_e2e": dc.n_windows / float(np.median(tm)) * 1e3, "gcups": dc.n_cells * t / float(np.median(tm)) / 1e6, "launches": dc.last_launches()})
	dc.close()


def run_long_query(name, locality, gap, t, n_sent, vocab=50_000, dim=300, k=100, steps=5):
	"""Queries of more than 16 tokens: S is uploaded (vx_sim_upload), the search takes the TP = 20 .. 32 score kernels."""
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, seed=synthetic.CORPUS_SEED, device=DEV)
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	del tok
	emb = synthetic.make_embedding(vocab, dim)
	ehat = emb / np.linalg.norm(emb, axis=1, keepdims=True)
	q = synthetic.make_query(t, vocab, seed=123)
	s_host = np.clip(ehat @ ehat[q].T, 0.0, 1.0).astype(np.float32)
	s_host[q, np.arange(t)] = 1.0
	sim = engine.DeviceSim.upload(s_host)
	r = gaps.resolve_gap_costs(gap, 64, t)
	kind = r.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	tm = []
	for i in range(2 + steps):
		t0 = time.perf_counter()
		hits = dc.search(sim, ap, qp)[0]
		if i >= 2:
			tm.append((time.perf_counter() - t0) * 1e3)
	lt = dc.last_timing()
	emit({"config": name, "sentences": n_sent, "T": t, "wall_ms": float(np.median(tm)), "ms": lt,
		"sentences_per_s_e2e": n_sent / float(np.median(tm)) * 1e3, "launches": dc.last_launches(), "top_score": float(hits["score"][0])})
	sim.close()
	dc.close()


def run_multi_query(name, locality, gap, t, n_sent, vocab, dim, n_queries=8, k=100, reps=5):
	"""Several queries against ONE linear-gap setting: n_queries single vx_search calls vs one
	vx_search_batch (score_linear_mq_kernel), same results."""
	n_sent = n_sent // WORLD
	tok, tables = synthetic.make_corpus_device(n_sent, vocab, seed=synthetic.CORPUS_SEED + RANK, device=DEV)
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
	attach_shard(dc, tables["n_doc"])
	del tok
	emb = synthetic.make_embedding(vocab, dim)
	de = engine.DeviceEmbedding(emb)
	r = gaps.resolve_gap_costs(gap, 64, t)
	kind = r.pop("gap_kind")
	ap, keep = _cabi.make_align_params(locality, kind, **r)
	qp, _ = _cabi.make_query_params(max_matches=k, min_score=0.0)
	queries = [synthetic.make_query(t, vocab, seed=99 + i) for i in range(n_queries)]
	res = {}
	for mode in ("single", "batch"):
		best = None
		for rep in range(reps):
			sync_all()
			t0 = time.perf_counter()
			sims = [engine.DeviceSim.static(de, q, emb[q]) for q in queries]
			if mode == "single":
				top = [dc.search(sim, ap, qp)[0] for sim in sims]
			else:
				top = [row[0][0] for row in dc.search_batch(sims, [ap], qp)]
			dt = (time.perf_counter() - t0) * 1e3
			for sim in sims:
				sim.close()
			best = dt if best is None or dt < best else best
		res[mode] = (best, top)
	same = all(np.array_equal(a, b) for a, b in zip(res["single"][1], res["batch"][1]))
	m = reduce_max({"single_ms": res["single"][0], "batch_ms": res["batch"][0]})
	emit({"config": name, "sentences": n_sent * WORLD, "T": t, "queries": n_queries, "single_calls_wall_ms": m["single_ms"],
		"batch_wall_ms": m["batch_ms"], "speedup": m["single_ms"] / m["batch_ms"],
		"sentence_queries_per_s_single": n_sent * WORLD * n_queries / m["single_ms"] * 1e3,
		"sentence_queries_per_s_batch": n_sent * WORLD * n_queries / m["batch_ms"] * 1e3, "identical_results": bool(same)})
	dc.close()


def main():
	which = sys.argv[1:] or ["c1", "c2"]
	dist_setup()
	if "c5" in which:
		run_sweep("C5 sweep: 64 queries x {linear 0,.1,.2,.5} x {Local,Global}, T=12, 10M sentences (one GPU holds the whole corpus)", 10_000_000)
	if "c5s" in which:   # one GPU's share of C5 on 8 GPUs (1.25 M sentences), on a single GPU
		run_sweep("C5 shard: 64 queries x 8 settings, T=12, 1.25M sentences (what each of 8 GPUs holds)", 1_250_000)
	if "c1" in which:   # the reference's own CPU-runnable case, on the GPU for comparison
		run_static("C1 NW Global linear(0.2) T=5 10k V=50k", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.2), 5, 10_000, 50_000, 300, cand_per_cell=3)
	if "c2" in which:
		run_static("C2 SW Local linear(0.2) T=8 1M V=200k", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8, 1_000_000, 200_000, 300, cand_per_cell=3.5)
	if "c2big" in which:
		run_static("C2x SW Local linear(0.2) T=8 20M V=200k", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8, 20_000_000, 200_000, 300, cand_per_cell=3.5)
	if "c4full" in which:   # BASELINE config 4 as stated: 5 M sentences (125 M tokens x 128 fp32 = 64 GB of vectors) over all GPUs
		run_contextual("C4 (full config) NW Global linear(0.1) contextual d=128 T=8, 5M sentences over %d GPU(s)" % WORLD,
			_cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), 8, 5_000_000, 128)
	if "c3" in which:       # the headline config with 50 M sentences in TOTAL (strong-scaling view; bench.py holds 50 M per GPU)
		run_static("C3 (50M sentences in total) WSB Local smooth_gap_cost(5) T=12 V=50k over %d GPU(s)" % WORLD,
			_cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, 50_000_000, 50_000, 300)
	if "c4" in which:
		run_contextual("C4 NW Global linear(0.1) contextual d=128 T=8, 200k sentences (1/3 of one GPU's C4 shard)", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), 8, 200_000, 128)
		run_contextual("C4b same, T=16", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), 16, 200_000, 128)
	if "wsbloc" in which:
		for loc, nm in ((_cabi.VX_LOCAL, "Local"), (_cabi.VX_GLOBAL, "Global"), (_cabi.VX_SEMIGLOBAL, "Semiglobal")):
			run_static("WSB %s smooth_gap_cost(5) T=12 10M V=50k" % nm, loc, gaps.smooth_gap_cost(5), 12, 10_000_000, 50_000, 300)
	if "win" in which:    # three-sentence windows (~75 tokens, up to 192): general gap costs over windows longer than 64 tokens
		run_windows("WSB Local smooth_gap_cost(5) T=12, windows of 3 sentences (step 1), 1M sentences", _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, 1_000_000, 3, 1)
	if "long" in which:   # a whole sentence as the query: 24 tokens (S assembled from two 16-column similarity launches by the host API; here uploaded)
		for nm, loc, gap in (("SW Local linear(0.2)", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2)), ("Gotoh Local affine(0.3,0.1)", _cabi.VX_LOCAL, gaps.AffineGapCost(u=0.3, v=0.1)),
				("WSB Local smooth_gap_cost(5)", _cabi.VX_LOCAL, gaps.smooth_gap_cost(5))):
			run_long_query("long query T=24: %s, 2M sentences" % nm, loc, gap, 24, 2_000_000)
		for t in (40, 64):   # longer than 32 tokens: the linear kernel goes up to 64 query tokens, the affine one to 48
			run_long_query("long query T=%d: SW Local linear(0.2), 2M sentences" % t, _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), t, 2_000_000)
		run_long_query("long query T=40: Gotoh Local affine(0.3,0.1), 2M sentences", _cabi.VX_LOCAL, gaps.AffineGapCost(u=0.3, v=0.1), 40, 2_000_000)
	if "mq" in which:
		run_multi_query("C2-mq: 8 SW Local linear(0.2) queries, T=8, 20M sentences, V=200k", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8, 20_000_000, 200_000, 300)
		run_multi_query("C2-mq12: 8 NW Global linear(0.2) queries, T=12, 20M sentences, V=50k", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.2), 12, 20_000_000, 50_000, 300)
		run_multi_query("C2-mq4: 4 SW Local linear(0.2) queries, T=8, 1M sentences, V=200k", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8, 1_000_000, 200_000, 300, n_queries=4)
	if "tag" in which:    # tag-weighted similarity on the fast kernels vs the per-window scorer (weights that are not exact in fp32)
		w = [1.0, 0.5, 1.5, 1.0, 2.0, 0.25, 1.0, 1.0, 0.75, 1.0, 1.25, 1.0]
		run_static("C3-tag: WSB Local smooth_gap_cost(5) T=12 tag-weighted (dyadic weights, fast kernel), 10M", _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, 10_000_000, 50_000, 300, tag_weights=w)
		run_static("C2-tag: SW Local linear(0.2) T=12 tag-weighted (dyadic weights, fast kernel), 10M", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 12, 10_000_000, 50_000, 300, cand_per_cell=3.5, tag_weights=w)
		run_static("C3-tag-exact: same, weights 0.3 .. 1.3 (per-window scorer), 200k", _cabi.VX_LOCAL, gaps.smooth_gap_cost(5), 12, 200_000, 50_000, 300, steps=3, tag_weights=[0.3 + 0.1 * i for i in range(12)])
	if "gotoh" in which:
		run_static("Gotoh Local affine(0.3,0.1) T=8 20M V=50k", _cabi.VX_LOCAL, gaps.AffineGapCost(u=0.3, v=0.1), 8, 20_000_000, 50_000, 300, cand_per_cell=7.5)


if __name__ == "__main__":
	main()
	if WORLD > 1:
		import torch.distributed as dist
		dist.destroy_process_group()
