import numpy as np, sys
sys.path.insert(0, '.')
from vectorian_b200 import _cabi, engine, gaps
from oracle import oracle
from tests import helpers
corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=400, t=21, seed=5)
s = oracle.sim_static(emb, q_id, q_vec)
dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
for loc in (0,1,2):
  for g in ("smooth5","affine","linear0.2"):
    ap, keep = helpers.align_params(loc, helpers.GAPS[g], dc.max_len, 21)
    qp, _ = _cabi.make_query_params(max_matches=10, min_score=0.0)
    sim = engine.DeviceSim.upload(s)
    a = dc.score_all(sim, ap, qp); b = oc.score_all(s, ap, qp, n_threads=1)
    bad = np.nonzero(a != b)[0]
    print(loc, g, "bad", bad, a[bad], b[bad], corpus["lens"][bad], "maxlen", dc.max_len)
    a2 = dc.score_all(sim, ap, qp)
    print("  rerun same?", np.array_equal(a, a2))
