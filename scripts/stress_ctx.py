#!/usr/bin/env python
"""Stress the fused contextual GEMM + DP kernel against the two-kernel path and the FP32 SIMT path (race hunting)."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vectorian_b200 import _cabi, engine, gaps  # noqa: E402
from tests import helpers  # noqa: E402
from tests.helpers import GAPS, LOCALITIES, align_params  # noqa: E402


def scores(dc, q, ap, qp, env=None):
	for k in ("VX_CTX_NOFUSE", "VX_CTX_SIMT"):
		os.environ.pop(k, None)
	if env:
		os.environ[env] = "1"
	try:
		sim = engine.DeviceSim.contextual(q)
		out = dc.score_all(sim, ap, qp)
		sim.close()
		return out
	finally:
		if env:
			os.environ.pop(env, None)


def main():
	reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
	shapes = [(16, 64), (12, 128), (7, 32), (3, 96), (8, 128), (16, 256), (8, 768)]
	bad = 0
	for t, dim in shapes:
		rng = np.random.default_rng(100 + t)
		corpus, emb, q_id, q_vec = helpers.small_problem(n_sentences=5000, t=t, seed=41)
		n_tok = len(corpus["token_id"])
		centroids = rng.standard_normal((50, dim)).astype(np.float32)
		x = centroids[corpus["token_id"] % 50] + 0.4 * rng.standard_normal((n_tok, dim)).astype(np.float32)
		q = centroids[rng.integers(0, 50, size=t)] + 0.4 * rng.standard_normal((t, dim)).astype(np.float32)
		dc = engine.DeviceCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"], corpus["span_end"],
			corpus["doc_span_off"], 1, 1)
		dc.set_contextual(x)
		lens = (corpus["span_end"] - corpus["span_start"])
		ap, keep = align_params(_cabi.VX_LOCAL, GAPS["linear0.2"], dc.max_len, t)
		qp, _ = _cabi.make_query_params(max_matches=40, min_score=-1e30)
		base_f = scores(dc, q, ap, qp)
		base_u = scores(dc, q, ap, qp, "VX_CTX_NOFUSE")
		print(f"shape T={t} d={dim}: fused==unfused first run: {np.array_equal(base_f, base_u)}", flush=True)
		for rep in range(reps):
			f = scores(dc, q, ap, qp)
			u = scores(dc, q, ap, qp, "VX_CTX_NOFUSE")
			for name, arr in (("fused", f), ("unfused", u)):
				d = np.flatnonzero(arr != base_u)
				if len(d):
					bad += 1
					for i in d[:4]:
						print(f"  rep {rep} {name}: window {i} len {lens[i]} got {arr[i]!r} want {base_u[i]!r}", flush=True)
		dc.close()
	print("VX_CTX_DBG", os.environ.get("VX_CTX_DBG"), "mismatching runs:", bad)


if __name__ == "__main__":
	main()
