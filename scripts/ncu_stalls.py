#!/usr/bin/env python
"""Per-instruction stall samples of one kernel instance in an `ncu --page source --csv --print-source sass` dump.
usage: ncu_stalls.py dump.csv <section index> [min share]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]); thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
sections = []; hdr = None; data = []
for r in rows:
	if r and r[0] == 'Kernel Name':
		if data: sections.append((hdr, data))
		data = []; hdr = None; continue
	if r and r[0] == 'Address': hdr = r; continue
	if hdr and len(r) == len(hdr): data.append(r)
sections.append((hdr, data))
hdr, data = sections[which]
ia = hdr.index('Source'); isamp = hdr.index('# Samples'); iex = hdr.index('Instructions Executed')
stall = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[isamp]) for r in data)
print('sections', len(sections), 'instrs', len(data), 'total samples', tot)
agg = {}
for r in data:
	for i, h in stall:
		agg[h] = agg.get(h, 0) + int(r[i] or 0)
print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
for n, r in enumerate(data):
	s = int(r[isamp])
	if s > tot * thr:
		top = sorted(((int(r[i] or 0), h) for i, h in stall), reverse=True)[:2]
		print(n, r[ia][:56].ljust(56), s, r[iex], top)
