#!/usr/bin/env python
"""Sharded sessions (one process per GPU, torchrun): index.sweep == individual index.find calls, and both equal the
single-process answers.  Usage: python -m torch.distributed.run --nproc-per-node 2 scripts/check_sweep_sharded.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

from vectorian_b200 import gaps  # noqa: E402
from vectorian_b200.alignment import GlobalAlignment, LocalAlignment  # noqa: E402
from vectorian_b200.corpus import whitespace_nlp  # noqa: E402
from vectorian_b200.index import sweep  # noqa: E402
from vectorian_b200.sim.span import OptimizedSpanSim  # noqa: E402
from vectorian_b200.sim.token import EmbeddingTokenSim  # noqa: E402
from vectorian_b200.sim.vector import CosineSim  # noqa: E402
from tests.test_gpu_api import make_session  # noqa: E402


def rows(result):
	return [(m.prepared_doc.id, m.slice_id, np.float32(m.score), tuple(int(x) for x in m.flow["target"])) for m in result]


def main():
	local_rank = int(os.environ.get("LOCAL_RANK", "0"))
	torch.cuda.set_device(local_rank)
	dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
	session, (emb,) = make_session(n_docs=9, sent_per_doc=80)
	assert session.shard[1] == dist.get_world_size()
	ts = EmbeddingTokenSim(emb, CosineSim())
	partition = session.partition("sentence")
	indexes = [partition.index(OptimizedSpanSim(ts, a), nlp=whitespace_nlp) for a in (
		LocalAlignment(gap=gaps.LinearGapCost(0.0)), LocalAlignment(gap=gaps.LinearGapCost(0.2)),
		GlobalAlignment(gap=gaps.LinearGapCost(0.1)), GlobalAlignment(gap=gaps.LinearGapCost(0.5)),
		LocalAlignment(gap=gaps.smooth_gap_cost(5)))]
	texts = ["w3 w17 w5 w120 w1 w9", "w2 w8 w30 w1", "w77 w5 w3 w200 w14 w6 w9 w1"]
	got = sweep(indexes, texts, n=15, min_score=-10.0)
	ok = True
	for qi, text in enumerate(texts):
		for ai, index in enumerate(indexes):
			want = index.find(text, n=15, min_score=-10.0)
			ok = ok and rows(got[qi][ai]) == rows(want)
	# twice more, the second time with the queries reversed: the result buffers (send / gather / merged / pinned) are
	# reused from call to call and must not alias the results of an earlier one
	dev1 = sweep(indexes, texts, n=15, min_score=-10.0)
	dev1_rows = [[rows(r) for r in row] for row in dev1]
	dev2 = sweep(indexes, list(reversed(texts)), n=15, min_score=-10.0)
	for qi in range(len(texts)):
		for ai in range(len(indexes)):
			ok = ok and dev1_rows[qi][ai] == rows(got[qi][ai]) and rows(dev1[qi][ai]) == dev1_rows[qi][ai]
			ok = ok and rows(dev2[len(texts) - 1 - qi][ai]) == rows(got[qi][ai])
	assert indexes[0].session.device_corpus(partition).sharded
	docs_seen = {r[0] for row in got for res in row for r in rows(res)}
	flag = torch.tensor([1 if ok else 0], device="cuda")
	dist.all_reduce(flag, op=dist.ReduceOp.MIN)
	if dist.get_rank() == 0:
		print("sharded sweep == sharded finds on every rank:", bool(flag.item()), "; documents in the results:", sorted(docs_seen)[:12])
	dist.destroy_process_group()
	sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
	main()
