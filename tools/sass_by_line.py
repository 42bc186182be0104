"""Developer tool: static SASS size of one kernel by source region (no GPU needed).
  cuobjdump -xelf all lib.so; nvdisasm -gi <cubin> > dis.txt; python tools/sass_by_line.py dis.txt <mangled kernel name> [depth]
Every instruction is attributed to the chain of `inlined at` frames nvdisasm prints in front of it; the table groups by the outermost
`depth` frames (default 2: the line of the kernel body and the line of the function inlined there)."""
import collections
import re
import sys

dis, kern = sys.argv[1], sys.argv[2]
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 2
inside = False
cur = None
chain, prev_annot = [], False
count = collections.Counter()
total = 0
ins = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(\S+)")
loc = re.compile(r'"([^"]+)", line (\d+)')
with open(dis) as f:
    for line in f:
        if line.startswith("\t.section") or line.startswith(".section"):
            inside = (".text." + kern) in line
            continue
        if not inside:
            continue
        if "//## File" in line:  # consecutive annotation lines are one chain, innermost frame first
            a, b = loc.findall(line)[0]
            chain = (chain if prev_annot else []) + [(a.rsplit("/", 1)[-1], int(b))]
            cur = tuple(reversed(chain))[:depth]  # outermost first
            prev_annot = True
            continue
        prev_annot = False
        m = ins.match(line)
        if m:
            count[cur] += 1
            total += 1
print("total SASS instructions", total, "= %.1f KB" % (total * 16 / 1024))
for k, v in sorted(count.items(), key=lambda kv: -kv[1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 60]:
    print("%5d  %5.1f%%  %s" % (v, 100.0 * v / total, " <- ".join("%s:%d" % fr for fr in (k or ()))))
