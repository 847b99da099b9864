#!/usr/bin/env python
"""Brief of an .ncu-rep: per captured launch the handful of metrics that decide which bound a DP kernel sits at, and its top
warp-stall reasons.  `python scripts/ncu_brief.py file.ncu-rep`."""
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
	'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
	'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
	'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
	'launch__registers_per_thread', 'launch__grid_size', 'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts.sum',
	'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum']


def main():
	out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
	rows = list(csv.reader(io.StringIO(out)))
	hdr, units = rows[0], rows[1]
	for r in rows[2:]:
		d = dict(zip(hdr, r))
		print(d['Kernel Name'])
		for w in WANT:
			if w in d:
				print('  %-70s %s %s' % (w, d[w], units[hdr.index(w)]))
		st = []
		for h in hdr:
			if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio') and d[h]:
				st.append((float(d[h].replace(',', '')), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
		st.sort(reverse=True)
		print('  stalls per issue: ' + ', '.join('%s %.2f' % (n, v) for v, n in st[:6]))


if __name__ == '__main__':
	main()
