#!/usr/bin/env python
"""Summarise one gpurun_out/<tag>/ directory (scripts/gpu_round.sh) into profiles/<name>/.

  python scripts/ncu_summary.py gpurun_out/<tag> profiles/<name>

Writes launches.csv (the ncu launch list, our kernels only), launches_summary.md (per-kernel
count / total / share of the step), kernel_full.csv (selected --set full metrics of the captured DP
kernel, per launch) and copies bench.json / bench_ref.json.
"""
import csv
import io
import os
import shutil
import subprocess
import sys
from collections import OrderedDict

WANT = [
	"gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
	"gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
	"sm__throughput.avg.pct_of_peak_sustained_elapsed",
	"smsp__issue_active.avg.pct_of_peak_sustained_active",
	"sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
	"sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
	"sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
	"sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
	"sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
	"sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
	"sm__warps_active.avg.pct_of_peak_sustained_active",
	"smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
	"launch__registers_per_thread", "launch__grid_size", "launch__block_size",
	"launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
	"launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
	"lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
	"l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
	"sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]

OURS = ("score_", "topk_", "trace_", "sim_", "normalize_", "pack_", "unpermute", "ctx_", "vx_", "batch_", "wsb_")


def main():
	src, dst = sys.argv[1], sys.argv[2]
	os.makedirs(dst, exist_ok=True)
	lp = os.path.join(src, "launches.csv")
	if os.path.exists(lp):
		with open(lp) as f:
			lines = [l for l in f if l.startswith('"')]
		rd = csv.DictReader(io.StringIO("".join(lines)))
		agg = OrderedDict()
		kept = []
		for r in rd:
			name = r["Kernel Name"]
			short = name.split("(")[0].replace("void ", "")
			if not any(o in short for o in OURS):
				continue
			ns = float(r["Metric Value"].replace(",", ""))
			kept.append((r["ID"], short, r["Grid Size"], r["Block Size"], ns))
			a = agg.setdefault(short, [0, 0.0])
			a[0] += 1
			a[1] += ns
		with open(os.path.join(dst, "launches.csv"), "w") as f:
			f.write("id,kernel,grid,block,gpu__time_duration_ns\n")
			for k in kept:
				f.write('%s,"%s","%s","%s",%.0f\n' % k)
		total = sum(a[1] for a in agg.values()) or 1.0
		with open(os.path.join(dst, "launches_summary.md"), "w") as f:
			f.write("ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)\n\n")
			f.write("| kernel | launches | total us | share |\n|---|---:|---:|---:|\n")
			for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
				f.write("| `%s` | %d | %.1f | %.1f%% |\n" % (k, a[0], a[1] / 1e3, 100 * a[1] / total))
	rep = os.path.join(src, "prof.ncu-rep")
	if os.path.exists(rep):
		out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
		rows = list(csv.reader(io.StringIO(out)))
		hdr, units, data = rows[0], rows[1], rows[2:]
		with open(os.path.join(dst, "kernel_full.csv"), "w") as f:
			f.write("metric,unit," + ",".join("launch%d" % i for i in range(len(data))) + "\n")
			ki = hdr.index("Kernel Name")
			f.write("Kernel Name,," + ",".join('"%s"' % d[ki].split("(")[0] for d in data) + "\n")
			for w in WANT:
				if w in hdr:
					i = hdr.index(w)
					f.write("%s,%s,%s\n" % (w, units[i], ",".join(d[i] for d in data)))
	for name in ("bench.json", "bench_ref.json", "smi.txt"):
		p = os.path.join(src, name)
		if os.path.exists(p):
			shutil.copy(p, os.path.join(dst, name))


if __name__ == "__main__":
	main()
