#!/usr/bin/env python
"""Scoreboards of a kernel's loads: for every LDG / LDS of the hot loop the write barrier it signals, and every instruction
that waits on a barrier -- shows how far ahead of its first use ptxas really issues a prefetch (it reorders loads freely and
shares the six barriers between them).  `python scripts/sass_sb.py <mangled-name substring>`."""
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'vectorian_b200', 'libvectorian_b200.so')


def main():
	text = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True).stdout
	name = None
	lines = text.splitlines()
	i = 0
	while i < len(lines):
		ln = lines[i]
		if 'Function :' in ln:
			name = ln.split('Function :')[1].strip()
			if sys.argv[1] in name:
				print(name)
		m = re.match(r'\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/', ln)
		if m and name and sys.argv[1] in name and i + 1 < len(lines):
			m2 = re.match(r'\s+/\* (0x[0-9a-f]+) \*/', lines[i + 1])
			if m2:
				ctrl = (int(m2.group(1), 16) >> 41) & 0x1fffff
				wbar, wait = (ctrl >> 5) & 7, (ctrl >> 11) & 0x3f
				t = m.group(2)
				if 'LDG' in t or 'LDS' in t or wait or 'BRA' in t:
					print('%s %-66s %s %s' % (m.group(1), t[:66], 'sets SB%d' % wbar if wbar != 7 else '        ',
						'waits ' + ','.join('SB%d' % b for b in range(6) if wait >> b & 1) if wait else ''))
				i += 2
				continue
		i += 1


if __name__ == '__main__':
	main()
