import numpy as np, sys
sys.path.insert(0, '.')
from vectorian_b200 import _cabi, gaps
from oracle import oracle
from tests import helpers
corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=400, t=21, seed=5)
s = oracle.sim_static(emb, q_id, q_vec)
oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"], corpus["doc_span_off"])
ap, keep = helpers.align_params(0, helpers.GAPS["smooth5"], 47, 21)
qp, _ = _cabi.make_query_params(max_matches=10, min_score=0.0)
b = oc.score_all(s, ap, qp, n_threads=1)
for i in range(20):
    a = oc.score_all(s, ap, qp, n_threads=4)
    bad = np.nonzero(a != b)[0]
    if len(bad): print(i, bad, a[bad], b[bad])
    s2 = oracle.sim_static(emb, q_id, q_vec)
    if not np.array_equal(s, s2): print("sim differs", np.abs(s-s2).max())
print("done")
