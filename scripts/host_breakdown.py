#!/usr/bin/env python
"""Where the host time of one query goes (wall clock around each public call), C3 at a given corpus size."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vectorian_b200 import _cabi, engine, gaps, synthetic

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6_250_000
tok, tables = synthetic.make_corpus_device(n, 50_000, device="cuda:0")
dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
emb = synthetic.make_embedding(50_000, 300)
de = engine.DeviceEmbedding(emb)
r = gaps.resolve_gap_costs(gaps.smooth_gap_cost(5), dc.max_len, 12)
kind = r.pop("gap_kind")
ap, keep = _cabi.make_align_params(_cabi.VX_LOCAL, kind, **r)
qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
qs = [synthetic.make_query(12, 50_000, seed=99 + i) for i in range(8)]
q_ids = [torch.from_numpy(q).pin_memory() for q in qs]
q_vecs = [torch.from_numpy(emb[q].copy()).pin_memory() for q in qs]
acc = {"sim_call": 0.0, "search_call": 0.0, "timing_queries": 0.0, "close": 0.0, "device": 0.0}
steps = 200
for i in range(steps + 20):
	j = i % 8
	t0 = time.perf_counter()
	sim = engine.DeviceSim.static(de, q_ids[j].numpy(), q_vecs[j].numpy())
	t1 = time.perf_counter()
	hits = dc.search(sim, ap, qp)
	t2 = time.perf_counter()
	dev = sim.build_ms + sum(dc.last_timing().values())
	t3 = time.perf_counter()
	sim.close()
	t4 = time.perf_counter()
	if i >= 20:
		acc["sim_call"] += t1 - t0; acc["search_call"] += t2 - t1; acc["timing_queries"] += t3 - t2; acc["close"] += t4 - t3; acc["device"] += dev / 1e3
print({k: round(v / steps * 1e3, 4) for k, v in acc.items()}, "ms per query; wall", round(sum(v for k, v in acc.items() if k != "device") / steps * 1e3, 4))
