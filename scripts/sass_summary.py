#!/usr/bin/env python
"""Per-kernel counts of the SASS opcodes that show what the kernels are built from (tcgen05 MMA / TMA / tensor-memory
loads and stores / packed fp32 adds / 3-input max): `cuobjdump -sass` over libvectorian_b200.so -> profiles/sass_summary.md.

  python scripts/sass_summary.py [library] [output.md]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OPS = ["UTCHMMA", "UTMALDG", "UTCBAR", "LDTM", "STTM", "FADD2", "FMNMX3", "FADD", "FFMA", "FMNMX", "LDG", "LDS", "STS", "SHFL", "SYNCS", "ATOMG", "RED"]


def demangle(names):
	out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
	return dict(zip(names, out))


def main():
	lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vectorian_b200", "libvectorian_b200.so")
	dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "sass_summary.md")
	text = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
	counts = collections.OrderedDict()
	cur = None
	for line in text.split("\n"):
		m = re.search(r"Function : (\S+)", line)
		if m:
			cur = m.group(1)
			counts[cur] = collections.Counter()
			continue
		m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
		if m and cur:
			counts[cur][m.group(1)] += 1
			counts[cur]["_total"] += 1
	names = demangle(list(counts))
	rows = []
	for k, c in counts.items():
		name = names.get(k, k).replace("void ", "").replace("(anonymous namespace)::", "")
		name = re.sub(r"_GLOBAL__N__\w+::", "", name)
		name = re.sub(r"\(.*", "", name)
		rows.append((name, c))
	rows.sort(key=lambda r: r[0])
	with open(dst, "w") as f:
		f.write("# SASS opcode counts per kernel (sm_100a)\n\n")
		f.write("`python scripts/sass_summary.py` = `cuobjdump -sass vectorian_b200/libvectorian_b200.so`, static instruction counts.\n")
		f.write("UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load (cp.async.bulk.tensor), UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st\n")
		f.write("(tensor memory), FADD2 = add.rn.f32x2, FMNMX3 = 3-input max.f32, SYNCS = mbarrier.\n\n")
		f.write("| kernel | total | " + " | ".join(OPS) + " |\n|---|---:|" + "---:|" * len(OPS) + "\n")
		for name, c in rows:
			f.write("| `%s` | %d | %s |\n" % (name, c["_total"], " | ".join(str(c[o]) if c[o] else "" for o in OPS)))
	print("wrote", dst, len(rows), "kernels")


if __name__ == "__main__":
	main()
