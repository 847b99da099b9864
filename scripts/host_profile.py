"""Where does the host-side time of one search go? (run on the GPU box)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vectorian_b200 import _cabi, engine, synthetic, gaps
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
tok, tables = synthetic.make_corpus_device(n, bench.VOCAB, device="cuda:0")
dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"], tables["doc_span_off"], on_device=True)
emb = synthetic.make_embedding(bench.VOCAB, bench.DIM)
de = engine.DeviceEmbedding(emb)
q = synthetic.make_query(12, bench.VOCAB)
qv = emb[q].copy()
ap, keep = bench.wsb_params(dc.max_len, 12)
qp, _ = _cabi.make_query_params(max_matches=100, min_score=0.0)
for it in range(6):
	t0 = time.perf_counter(); sim = engine.DeviceSim.static(de, q, qv)
	t1 = time.perf_counter(); r = dc.search(sim, ap, qp)
	t2 = time.perf_counter(); tm = dc.last_timing(); bm = sim.build_ms
	t3 = time.perf_counter(); sim.close()
	t4 = time.perf_counter()
	print(f"sim.static {1e3*(t1-t0):.3f} ms (dev {bm:.3f}) | search {1e3*(t2-t1):.3f} ms (dev {sum(tm.values()):.3f}: {tm}) | timing {1e3*(t3-t2):.3f} | close {1e3*(t4-t3):.3f}")
