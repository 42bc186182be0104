"""Developer tool: executed instruction footprint of a kernel from an ncu source page (SASS).  Prints, for a set of execution-count
thresholds, how many distinct SASS instructions (x 16 B) run at least that often — the part of the kernel that competes for the
32 KB L1.5 instruction cache — and the CUDA source lines that own the hot instructions.
  python tools/ncu_footprint.py gpurun_out/prof.ncu-rep"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ci = {k: i for i, k in enumerate(hdr)}
def num(x):
    try: return float(x)
    except Exception: return 0.0
ex = [num(r[ci["Instructions Executed"]]) for r in data]
tot = sum(ex); mx = max(ex)
print("SASS instructions %d (%.1f KB), executed warp instructions %.4g" % (len(ex), len(ex) * 16 / 1024, tot))
for frac in (0.5, 0.25, 0.1, 0.03, 0.01, 0.001):
    hot = [e for e in ex if e >= frac * mx]
    print("  executed >= %5.1f%% of the hottest instruction's count: %5d instr = %5.1f KB, %5.1f%% of all executed" % (100 * frac, len(hot), len(hot) * 16 / 1024, 100 * sum(hot) / tot))
