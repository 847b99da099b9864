#!/usr/bin/env python
"""Opcode counts of the innermost hot loop of a kernel: `python scripts/sass_loop.py <substring of the mangled name> [...]`.

Reads `cuobjdump -sass` of the built library, takes for each matching kernel the backward branch with the largest
span (the row loop of the DP kernels) and prints how many instructions of each opcode it holds -- the quickest way to
see what a DP cell really costs in issue slots (register moves, address arithmetic, predicate set-up) before going to the GPU.
"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'vectorian_b200', 'libvectorian_b200.so')
LINE = re.compile(r'^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);')


def kernels(text):
	name, body = None, []
	for ln in text.splitlines():
		if 'Function :' in ln:
			if name:
				yield name, body
			name, body = ln.split('Function :')[1].strip(), []
			continue
		m = LINE.match(ln)
		if m and name:
			body.append((int(m.group(1), 16), m.group(2).strip()))
	if name:
		yield name, body


def opcode(ins):
	parts = ins.split()
	if parts[0].startswith('@'):
		parts = parts[1:]
	return parts[0].split('.')[0]


def main():
	text = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True).stdout
	for name, body in kernels(text):
		if not any(w in name for w in sys.argv[1:]):
			continue
		best = None
		for addr, ins in body:
			m = re.search(r'\bBRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)', ins)
			if m:
				dst = int(m.group(1), 16)
				if dst < addr and (best is None or addr - dst > best[1] - best[0]):
					best = (dst, addr)
		if not best:
			print(name, ': no loop')
			continue
		ops = collections.Counter(opcode(i) for a, i in body if best[0] <= a <= best[1])
		n = sum(ops.values())
		print('%s\n  loop %#x..%#x: %d instructions  %s' % (name, best[0], best[1], n,
			'  '.join('%s %d' % kv for kv in ops.most_common(14))))


if __name__ == '__main__':
	main()
