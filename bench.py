#!/usr/bin/env python
"""Headline benchmark: Waterman-Smith-Beyer brute-force search (BASELINE.json configs[2]).

One step = one query through the whole hot path over the resident corpus:
similarity GEMM (query x vocabulary) -> per-sentence WSB alignment -> device top-k ->
traceback of the winners -> (N > 1) all-gather + device merge of the per-rank top-k.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--sentences S] [--weak]

N > 1 is launched by torchrun (one rank per GPU, NCCL).  STRONG scaling, as configs[2] names it: `--sentences`
(default 50M) is the corpus IN TOTAL, sharded by documents over the N ranks (50M / N sentences per GPU); every rank
generates the same global corpus and keeps its own documents.  `--weak` keeps `--sentences` per GPU instead.
After the timed loop the sharded result of one query is checked on the hardware against a merge built from every
rank's vx_score_all scores ("sharded_parity").
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
	sys.path.insert(0, ROOT)

T_QUERY = 12
VOCAB = 50_000
DIM = 300
TOPK = 100
N_QUERIES = 8


def wsb_params(max_len, t):
	from vectorian_b200 import _cabi, gaps
	r = gaps.resolve_gap_costs(gaps.smooth_gap_cost(5), max_len, t)
	kind = r.pop("gap_kind")
	return _cabi.make_align_params(_cabi.VX_LOCAL, kind, **r)


def ops_per_sentence(lens, t):
	"""Algorithmic FP32 lane-ops of Local WSB: 2 per candidate + 1 zero-compare per cell."""
	l = lens.astype(np.float64)
	cand = l * t + t * l * (l + 1) / 2 + l * t * (t + 1) / 2
	return 2 * cand + l * t


class ClockSampler:
	"""nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md: `-lms 200`, started before, killed
	after).  The polling is not free: on 8 GPUs, where the timed region is 50 ms long, a 100 ms period cost 2-4 % of the
	end-to-end rate through longer waits in the collective (profiles/r2_final/README.md); hence the recipe's period and the
	start before the warm-up."""

	def __init__(self, index):
		self.rows = []
		q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
			"clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
			"clocks_event_reasons.sw_power_cap")
		try:
			self.p = subprocess.Popen(
				["nvidia-smi", f"--id={index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200"],
				stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
			self.t = threading.Thread(target=self._read, daemon=True)
			self.t.start()
		except Exception:
			self.p = None

	def _read(self):
		for line in self.p.stdout:
			self.rows.append((time.perf_counter(), line.strip()))

	def wait_first(self, timeout_s):
		t_end = time.perf_counter() + timeout_s
		while self.p is not None and not self.rows and time.perf_counter() < t_end:
			time.sleep(0.02)

	def stop(self, t0, t1):
		if self.p is None:
			return None
		time.sleep(0.15)
		self.p.terminate()
		sm, mx, reasons = [], [], set()
		for ts, line in self.rows:
			f = [x.strip() for x in line.split(",")]
			if len(f) < 7 or not (t0 <= ts <= t1 + 0.2):
				continue
			try:
				sm.append(float(f[0]))
				mx.append(float(f[1]))
			except ValueError:
				continue
			for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
				if val.lower().startswith("active"):
					reasons.add(name)
		if not sm:
			return None
		return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_rate(sample_sentences, threads, seed=1234):
	"""Times the oracle (CPU port of the reference path) on a bounded sample of the same workload."""
	from vectorian_b200 import _cabi, synthetic
	from oracle import oracle
	corpus = synthetic.make_corpus(sample_sentences, VOCAB, seed=seed)
	emb = synthetic.make_embedding(VOCAB, DIM)
	q_id = synthetic.make_query(T_QUERY, VOCAB)
	oc = oracle.OracleCorpus(corpus["token_id"], corpus["doc_tok_off"], corpus["span_start"],
		corpus["span_end"], corpus["doc_span_off"])
	ap, keep = wsb_params(int(corpus["lens"].max()), T_QUERY)
	qp, _ = _cabi.make_query_params(max_matches=TOPK, min_score=0.0)
	t0 = time.perf_counter()
	s = oracle.sim_static(emb, q_id, emb[q_id])
	hits, _, _ = oc.search(s, ap, qp, n_threads=threads)
	dt = time.perf_counter() - t0
	return sample_sentences / dt, dt, len(hits)


def cpu_reference_rate(sample_sentences, threads, seed=1234):
	"""Times the reference's OWN solver (oracle/_ref: the unmodified pyalign solver.h compiled against an xtensor
	stand-in, oracle/ref/) on a bounded sample of the same workload, the way the reference runs the path: S with
	numpy (vectorian/sim/vector.py:66-78), then per sentence solve() + alignment() on a pool of threads over
	documents (vectorian/index.py:544-551, cpp/metric/alignment.h:247-294), then the k best."""
	from vectorian_b200 import _cabi, gaps, synthetic
	from oracle import ref
	corpus = synthetic.make_corpus(sample_sentences, VOCAB, seed=seed)
	emb = synthetic.make_embedding(VOCAB, DIM)
	ehat = emb / np.maximum(np.linalg.norm(emb, axis=1, keepdims=True), 1e-30)   # Vectors.normalized is cached per embedding
	q_id = synthetic.make_query(T_QUERY, VOCAB)
	lens = corpus["lens"]
	sent_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
	max_len = int(lens.max())
	resolved = gaps.resolve_gap_costs(gaps.smooth_gap_cost(5), max_len, T_QUERY)
	t0 = time.perf_counter()
	s = np.ascontiguousarray(ehat @ ehat[q_id].T, dtype=np.float32)
	s[q_id, np.arange(T_QUERY)] = 1.0
	np.clip(s, 0.0, 1.0, out=s)
	raw = ref.corpus_scores(_cabi.VX_LOCAL, resolved, s, corpus["token_id"], sent_off, max_len, n_threads=threads)
	score = raw / np.float32(T_QUERY)
	top = np.argpartition(-score, min(TOPK, len(score) - 1))[:TOPK]
	top = top[np.argsort(-score[top], kind="stable")]
	dt = time.perf_counter() - t0
	return sample_sentences / dt, dt, len(top)


def cpu_baseline(sample_sentences, threads, seed=1234):
	"""(rate, seconds, kind, sample description): the compiled reference solver where oracle/_ref exists, else the port."""
	from oracle import ref
	if ref.available():
		r, dt, _ = cpu_reference_rate(sample_sentences, threads, seed)
		return r, dt, "reference"
	r, dt, _ = cpu_port_rate(sample_sentences, threads, seed)
	return r, dt, "port"


def run_reference(args):
	"""--impl reference: the reference's CPU implementation of the path on the host cores.

	`import vectorian` cannot run here (h5py / spaCy / xtensor absent, no network), but the hot path's arithmetic --
	the vendored pyalign solver.h -- compiles against an xtensor stand-in (oracle/ref/, built into oracle/_ref/ in the
	development container): kind "reference".  Where that library is missing the oracle port is timed (kind "port").
	All host threads; each step is a bounded sample of the workload."""
	rank = int(os.environ.get("RANK", "0"))
	if rank != 0:
		return
	threads = os.cpu_count() or 1
	sample = args.cpu_sample
	rates = []
	kind = "port"
	for i in range(args.warmup + args.steps):
		r, dt, kind = cpu_baseline(sample, threads, seed=1234 + i)
		if i >= args.warmup:
			rates.append((r, dt))
	value = float(np.mean([r for r, _ in rates]))
	line = {
		"impl": "reference", "metric": "sentences/sec (WSB alignment, 12-token query, top-100)",
		"value": value, "unit": "sentences/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
		"ms_per_step": float(np.mean([dt for _, dt in rates]) * 1e3), "higher_is_better": True, "scaling": "weak" if args.weak else "strong",
		"vs_baseline": None, "dtype": "f32", "data": "synthetic",
		"config": {"workload": workload_name(args), "sample_sentences_per_step": sample},
		"cpu_baseline": {"value": value, "unit": "sentences/s", "cores": threads, "kind": kind,
			"sample": f"{sample} sentences per step of the same synthetic workload",
			"what": "pyalign solver.h (unmodified, xtensor stand-in) + numpy S + top-k" if kind == "reference" else "oracle port"},
		"e2e": {"value": value, "unit": "sentences/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
		"gcups": value * 25 * T_QUERY / 1e9,
	}
	emit_line(json.dumps(line))


def workload_name(args):
	world = max(int(getattr(args, "gpus", 1) or 1), 1)
	if getattr(args, "weak", False):
		size = f"{args.sentences} sentences per GPU"
	elif world > 1:
		size = f"{args.sentences} sentences in total, {args.sentences // world} per GPU (sharded by documents over {world} GPUs)"
	else:
		size = f"{args.sentences} sentences in total"
	return (f"C3: Waterman-Smith-Beyer Local, smooth_gap_cost(5), T={T_QUERY}, V={VOCAB}x{DIM} static embedding, "
		f"top-{TOPK}, {size} (~25 tokens each, Zipf ids)")


def shard_tables(tok, tables, lo, hi):
	"""The documents [lo, hi) of a corpus made by synthetic.make_corpus_device: (token view, span tables)."""
	t0, t1 = int(tables["doc_tok_off"][lo]), int(tables["doc_tok_off"][hi])
	s0, s1 = int(tables["doc_span_off"][lo]), int(tables["doc_span_off"][hi])
	sub = dict(doc_tok_off=tables["doc_tok_off"][lo:hi + 1] - t0, span_start=tables["span_start"][s0:s1],
		span_end=tables["span_end"][s0:s1], doc_span_off=tables["doc_span_off"][lo:hi + 1] - s0,
		lens=tables["lens"][s0:s1], n_doc=hi - lo)
	return tok[t0:t1], sub


def sharded_parity(dc, sim, ap, qp, hits, rank, world, doc_lo, sentences_per_doc=1000):
	"""On-hardware check of 'N shards + NCCL gather + device merge' (ResultSet::extend + best_n, reference:
	vectorian/core/cpp/result_set.h:70-93, cpp/result_set.cpp:3-21): every rank scores ALL its windows (vx_score_all,
	independent of the device top-k and of the merge), keeps its best k under (score desc, doc desc, slice desc) with
	numpy, the lists are gathered as Python objects and merged on rank 0; the result must equal the hits the timed
	path returned -- scores, document ids and slice ids, in order."""
	import torch.distributed as dist
	score = dc.score_all(sim, ap, qp)
	k = int(qp.max_matches)
	ok = score > np.float32(qp.min_score)
	idx = np.nonzero(ok)[0]
	if len(idx) > k:
		part = np.argpartition(-score[idx], k - 1)[:k]
		kth = score[idx][part].min()
		idx = idx[score[idx] >= kth]          # everything tied with the k-th score stays in
	doc = idx // sentences_per_doc + doc_lo    # synthetic corpus: every sentence is one window, 1000 per document
	sl = idx % sentences_per_doc
	rows = list(zip(score[idx].tolist(), doc.tolist(), sl.tolist()))
	gathered = [None] * world
	if world > 1:
		dist.all_gather_object(gathered, rows)
	else:
		gathered = [rows]
	if rank != 0:
		return None
	allrows = sorted((r for part in gathered for r in part), key=lambda r: (-r[0], -r[1], -r[2]))[:k]
	got = [(float(h["score"]), int(h["doc"]), int(h["slice"])) for h in hits]
	if got == [(float(np.float32(a)), b, c) for a, b, c in allrows]:
		return "ok"
	return "MISMATCH: %d rows, first difference at %s" % (len(got), next((i for i, (x, y) in enumerate(zip(got, allrows)) if x != (float(np.float32(y[0])), y[1], y[2])), "length"))


def main():
	ap_ = argparse.ArgumentParser()
	ap_.add_argument("--gpus", type=int, default=1)
	ap_.add_argument("--steps", type=int, default=20)
	ap_.add_argument("--warmup", type=int, default=3)
	ap_.add_argument("--impl", default="ours")
	ap_.add_argument("--sentences", type=int, default=50_000_000)
	ap_.add_argument("--cpu-sample", type=int, default=1_000_000)
	ap_.add_argument("--no-cpu-baseline", action="store_true")
	ap_.add_argument("--no-clocks", action="store_true")
	ap_.add_argument("--weak", action="store_true", help="--sentences per GPU instead of in total")
	ap_.add_argument("--no-secondary", action="store_true")
	args = ap_.parse_args()
	args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

	if args.impl == "reference":
		run_reference(args)
		return

	import torch
	import torch.distributed as dist
	from vectorian_b200 import _cabi, engine, synthetic
	from vectorian_b200.distributed import device_comm

	rank = int(os.environ.get("RANK", "0"))
	world = int(os.environ.get("WORLD_SIZE", "1"))
	local_rank = int(os.environ.get("LOCAL_RANK", "0"))
	args.gpus = world
	torch.cuda.set_device(local_rank)
	if world > 1:
		dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

	# ---- resident state: corpus shard + embedding (setup, untimed) ----
	t_setup = time.perf_counter()
	if args.weak or world == 1:
		tok, tables = synthetic.make_corpus_device(args.sentences, VOCAB, seed=synthetic.CORPUS_SEED + (rank if args.weak else 0),
			device=f"cuda:{local_rank}")
		doc_lo = rank * tables["n_doc"]
	else:
		# strong scaling: the SAME global corpus on every rank, this rank keeps a contiguous range of its documents
		tok_all, tables_all = synthetic.make_corpus_device(args.sentences, VOCAB, seed=synthetic.CORPUS_SEED, device=f"cuda:{local_rank}")
		n_doc_all = tables_all["n_doc"]
		doc_lo, doc_hi = n_doc_all * rank // world, n_doc_all * (rank + 1) // world
		tok, tables = shard_tables(tok_all, tables_all, doc_lo, doc_hi)
		del tok_all
	torch.cuda.synchronize()
	dc = engine.DeviceCorpus(tok, tables["doc_tok_off"], tables["span_start"], tables["span_end"],
		tables["doc_span_off"], on_device=True)
	lens = tables["lens"]
	del tok
	torch.cuda.empty_cache()
	if world > 1:
		comm = device_comm()
		dc.set_shard(comm, doc_lo)      # vx_search now returns the global top-k on every rank
	emb = synthetic.make_embedding(VOCAB, DIM)
	de = engine.DeviceEmbedding(emb)
	queries = [synthetic.make_query(T_QUERY, VOCAB, seed=synthetic.QUERY_SEED + i) for i in range(N_QUERIES)]
	# per-step inputs live in pinned host memory: query token ids + their (unnormalised) vectors
	q_ids = [torch.from_numpy(q).pin_memory() for q in queries]
	q_vecs = [torch.from_numpy(emb[q].copy()).pin_memory() for q in queries]
	ap, keep = wsb_params(dc.max_len, T_QUERY)
	qp, _ = _cabi.make_query_params(max_matches=TOPK, min_score=0.0)
	setup_s = time.perf_counter() - t_setup

	def step(i):
		"""The public-API call a user makes; inputs from pinned host memory, results back on the host.  With N > 1 the
		all-gather of the ranks' result rows and their merge happen inside vx_search (device time in gather_merge_ms)."""
		j = i % N_QUERIES
		sim = engine.DeviceSim.static(de, q_ids[j].numpy(), q_vecs[j].numpy())
		hits, target, distance = dc.search(sim, ap, qp)
		timing = dc.last_timing()            # read once per step: every read is a handful of cudaEventElapsedTime calls
		dev_ms = sim.build_ms + sum(timing.values())
		launches = sim.launches + dc.last_launches()
		sim.close()
		return hits, dev_ms, launches, timing

	def barrier():
		if world > 1:
			dist.barrier()
		torch.cuda.synchronize()

	# The sampler is a polling nvidia-smi process: start it BEFORE the warm-up and wait for its first line, so that its
	# start-up (NVML initialisation touches the driver of every GPU of the box) is over when the timed region begins --
	# at N = 8 that region is only 50 ms long.
	sampler = ClockSampler(local_rank) if rank == 0 and not args.no_clocks else None
	if sampler:
		sampler.wait_first(3.0)
	for i in range(args.warmup):
		step(i)

	barrier()
	t0 = time.perf_counter()
	dev_ms_total, dp_ms_total, launches_total = 0.0, 0.0, 0
	last_hits = None
	gather_ms_total = 0.0
	for i in range(args.steps):
		last_hits, dev_ms, launches, timing = step(args.warmup + i)
		dev_ms_total += dev_ms
		dp_ms_total += timing["dp_ms"]
		gather_ms_total += timing["gather_merge_ms"]
		launches_total += launches
	barrier()
	t1 = time.perf_counter()
	clocks = sampler.stop(t0, t1) if sampler else None

	# ---- on-hardware check of the sharded path (untimed): the hits of one more query vs. every rank's vx_score_all ----
	j = (args.warmup + args.steps) % N_QUERIES
	sim = engine.DeviceSim.static(de, q_ids[j].numpy(), q_vecs[j].numpy())
	check_hits = dc.search(sim, ap, qp)[0]
	dc.set_shard(None, 0)                     # score_all and the local lists are per rank
	parity = sharded_parity(dc, sim, ap, qp, check_hits, rank, world, doc_lo)
	if world > 1:
		dc.set_shard(comm, doc_lo)
	sim.close()

	wall_s = t1 - t0
	times = torch.tensor([wall_s, dev_ms_total / 1e3, dp_ms_total / 1e3], dtype=torch.float64, device="cuda")
	counts = torch.tensor([float(dc.n_windows), float(dc.n_cells), float(launches_total)], dtype=torch.float64, device="cuda")
	if world > 1:
		dist.all_reduce(times, op=dist.ReduceOp.MAX)
		dist.all_reduce(counts, op=dist.ReduceOp.SUM)
	# per-rank means (diagnostic: which rank the others wait for in the collective)
	mine = torch.tensor([dp_ms_total / args.steps, gather_ms_total / args.steps, dev_ms_total / args.steps, (t1 - t0) * 1e3 / args.steps],
		dtype=torch.float64, device="cuda")
	per_rank = [mine.clone() for _ in range(world)]
	if world > 1:
		dist.all_gather(per_rank, mine)
	per_rank = [[round(float(x), 4) for x in r.cpu()] for r in per_rank]
	wall_s, dev_s, dp_s = [float(x) for x in times.cpu()]
	n_sent, n_cells, launches_all = [float(x) for x in counts.cpu()]

	if rank == 0:
		peaks = {}
		try:
			peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
		except Exception:
			pass
		hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
		peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
		sm_max = float(peaks.get("sm_max_mhz", 1965.0))
		sms = _cabi.lib().vx_device_sm_count()

		value = n_sent * args.steps / dev_s            # device-timed (CUDA events), inputs resident
		e2e = n_sent * args.steps / wall_s             # public API, host buffers, H2D/D2H inside
		# dominant kernel = the WSB DP: algorithmic bytes = 4 B per token id + 4 B per sentence
		alg_bytes = 4.0 * dc.n_cells + 4.0 * dc.n_windows
		dp_ms = dp_s * 1e3 / args.steps
		achieved = alg_bytes / (dp_ms * 1e-3) / 1e9
		lane_ops = float(ops_per_sentence(lens, T_QUERY).sum())
		issue_peak = sms * 128 * sm_max * 1e6
		issue_ach = lane_ops / (dp_ms * 1e-3)

		traffic, traffic_note = None, None
		try:
			tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
			ratio = (tr["dram_bytes_read"] + tr["dram_bytes_write"]) / tr["algorithmic_bytes"]
			traffic = ratio * alg_bytes
			traffic_note = "ncu dram bytes / algorithmic bytes = %.3f on the %d-sentence capture (%s), scaled to this launch" % (
				ratio, tr["workload_sentences"], tr["source"])
		except Exception:
			pass

		line = {
			"metric": "sentences/sec (WSB alignment, 12-token query, top-100)",
			"value": value, "unit": "sentences/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
			"ms_per_step": dev_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak" if args.weak else "strong",
			"vs_baseline": None, "dtype": "f32", "data": "synthetic",
			"config": {"workload": workload_name(args), "l2": "inputs larger than L2 (token tiles %.1f GB per GPU)" % (4.0 * dc.n_cells / 1e9),
				"queries": f"{N_QUERIES} distinct queries in rotation", "parallelism": f"sentence shards x{world}, one NCCL all-gather of the k result rows + device merge inside vx_search" if world > 1 else "1 GPU"},
			"sharded_parity": parity,
			"value_includes": "device time (CUDA events) of similarity build + DP + top-k + traceback" + (" + all-gather + merge" if world > 1 else ""),
			"gcups": n_cells * T_QUERY * args.steps / dev_s / 1e9,
			"e2e": {"value": e2e, "unit": "sentences/s",
				"h2d_bytes_per_step": int(T_QUERY * 4 + T_QUERY * DIM * 4 + (dc.max_len + 1 + T_QUERY + 1) * 4),
				"d2h_bytes_per_step": int(TOPK * (24 + 6 * T_QUERY) + 16), "ms_per_step": wall_s * 1e3 / args.steps},
			"gpu_launches": int(launches_all),
			# for the static-embedding configs the binding roofline of the DP kernel is FP32 issue
			# (achieved = algorithmic FP32 add / max operations per second, peak = SMs x 128 lanes x clock); the HBM view of the
			# same kernel (algorithmic bytes, ncu DRAM traffic) is reported beside it.
			"roofline": {"bound": "fp32-issue", "kernel": "wsb_score_kernel (vx_wsb.cu), all length-class launches of one query",
				"achieved": issue_ach / 1e12, "peak": issue_peak / 1e12, "unit": "TFLOP/s", "frac": issue_ach / issue_peak,
				"traffic": traffic, "traffic_note": traffic_note,
				"ops_per_launch": lane_ops, "ops_definition": "2 FP32 ops per DP candidate + 1 per Local cell (SURVEY.md 8d): 12600 per 25 x 12 sentence",
				"peak_source": "SMs (%d) x 128 FP32 lanes x %.0f MHz (MEASURED_PEAKS.json sm_max_mhz); the definition counts 2 lane-ops per candidate, the kernel spends one FADD2 + one FMNMX3 per TWO candidates (packed add.rn.f32x2), so the fraction can exceed 1: its ceiling is 2.0 (alu pipe: one FMNMX3 per scheduler every 2 cycles; measured 2.2-2.3 cycles, i.e. about 1.8)" % (sms, sm_max),
				"ms_per_launch": dp_ms,
				"note": "not HBM-bound and no tensor-core work: HBM view in hbm_roofline"},
			"hbm_roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
				"traffic": traffic, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src},
			"phases_ms": dict({k: v for k, v in dc.last_timing().items()}, gather_merge_ms_mean=gather_ms_total / args.steps),
			"per_rank_ms": {"columns": ["dp", "gather_merge (incl. waiting for the slowest rank)", "device", "wall"], "rows": per_rank},
			"setup_s": setup_s,
			"top_score": float(last_hits["score"][0]) if len(last_hits) else None,
		}
		if clocks:
			line["clocks"] = clocks
		if world == 1 and not args.no_cpu_baseline:
			threads = os.cpu_count() or 1
			r, dt, kind = cpu_baseline(args.cpu_sample, threads)
			line["cpu_baseline"] = {"value": r, "unit": "sentences/s", "cores": threads, "kind": kind,
				"sample": f"{args.cpu_sample} sentences of the same synthetic workload, {dt:.1f} s",
				"what": "pyalign solver.h (unmodified, xtensor stand-in) + numpy S + top-k" if kind == "reference" else "oracle port"}
			rp, dtp, _ = cpu_port_rate(args.cpu_sample, threads)
			line["cpu_port"] = {"value": rp, "unit": "sentences/s", "cores": threads,
				"what": "oracle/vx_oracle.c (optimised C restatement), same sample"}
		if world == 1 and not args.no_secondary:
			line["secondary"] = secondary_configs(emb)
		emit_line(json.dumps(line))

	if world > 1:
		dist.destroy_process_group()


def secondary_configs(emb50k):
	"""The other BASELINE.json configs on this GPU, a few steps each (scripts/bench_configs.py): so that the driver sees
	them too.  Times are medians of device phases (CUDA events); each entry names its own bound."""
	import importlib.util
	spec = importlib.util.spec_from_file_location("bench_configs", os.path.join(ROOT, "scripts", "bench_configs.py"))
	bc = importlib.util.module_from_spec(spec)
	spec.loader.exec_module(bc)
	from vectorian_b200 import _cabi, gaps
	bc.COLLECT = []
	out = {}
	try:
		bc.run_static("C2 SW Local linear(0.2) T=8 1M V=200k", _cabi.VX_LOCAL, gaps.LinearGapCost(0.2), 8, 1_000_000, 200_000, 300, steps=8, cand_per_cell=3.5)
		r = bc.COLLECT[-1]
		out["c2"] = {"workload": r["config"], "sentences_per_s_device": r["sentences_per_s_device"], "sentences_per_s_e2e": r["sentences_per_s_e2e"],
			"device_ms": r["device_ms"], "phases_ms": r["ms"], "roofline": {"bound": "fp32-issue (SURVEY 8d currency; the kernel itself sits at the L1TEX wavefront limit of the S gather, profiles/r1_linear)",
			"frac": r["dp_issue_frac"]}, "sim_hbm_frac": r["sim_hbm_frac"]}
		bc.run_contextual("C4 shard: NW Global linear(0.1) contextual d=128 T=8, 200k sentences (5 M tokens; 768->128 PCA taken as given)", _cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), 8, 200_000, 128, steps=8)
		r = bc.COLLECT[-1]
		out["c4_shard"] = {"workload": r["config"], "sentences_per_s_device": r["sentences_per_s_device"], "sentences_per_s_e2e": r["sentences_per_s_e2e"],
			"phases_ms": r["ms"], "roofline": {"bound": "hbm", "frac": r["gemm_dp_hbm_frac_lower_bound"],
			"note": "fused tcgen05 GEMM + DP kernel: algorithmic bytes tokens x (4 d + 4 TP) / kernel time vs MEASURED_PEAKS hbm_gbs"}}
		bc.run_sweep("C5 on one GPU: 64 queries x {linear 0,.1,.2,.5} x {Local,Global}, T=12, 10M sentences", 10_000_000, embedding=emb50k)
		r = bc.COLLECT[-1]
		out["c5_1gpu"] = {"workload": r["config"], "batch_wall_ms": r["batch_wall_ms"], "batch_device_ms": r["batch_device_ms"],
			"single_calls_wall_ms": r["single_calls_wall_ms"], "sentence_alignments_per_s_e2e": r["sentence_alignments_per_s_e2e"], "launches": r["launches"]}
	except Exception as e:   # the headline line must not depend on the secondary block
		out["error"] = repr(e)
	return out


def emit_line(text):
	"""The JSON line goes to the REAL stdout; everything else written to fd 1 while the bench runs (e.g. the
	"NCCL version ..." banner NCCL prints on stdout under NCCL_DEBUG=VERSION) has been redirected to stderr, so that
	stdout holds exactly one line."""
	os.write(_REAL_STDOUT, (text + "\n").encode())


_REAL_STDOUT = 1


if __name__ == "__main__":
	sys.stdout.flush()
	_REAL_STDOUT = os.dup(1)
	os.dup2(2, 1)
	main()
