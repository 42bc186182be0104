"""Developer tool: per-basic-block view of an ncu source page (SASS): consecutive instructions with the same executed
count are one block; prints blocks by share of executed warp instructions, with their stall samples and opcode mix.
  python tools/ncu_blocks.py gpurun_out/prof.ncu-rep [min_share_pct]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ci = {k: i for i, k in enumerate(hdr)}
def num(x):
    try: return float(x)
    except Exception: return 0.0
blocks = []; cur = None
for r in data:
    ex = num(r[ci["Instructions Executed"]]); th = num(r[ci["Thread Instructions Executed"]]); smp = num(r[ci["# Samples"]])
    op = r[ci["Source"]].split()[0] if r[ci["Source"]] else "?"
    if op.startswith("@"): op = r[ci["Source"]].split()[1]
    if cur is None or cur["ex"] != ex:
        cur = {"ex": ex, "n": 0, "th": 0.0, "smp": 0.0, "first": r[ci["Address"]], "ops": collections.Counter(), "firstsrc": r[ci["Source"]]}
        blocks.append(cur)
    cur["n"] += 1; cur["th"] += th; cur["smp"] += smp; cur["ops"][op.split(".")[0]] += 1
tot = sum(b["ex"] * b["n"] for b in blocks); tsmp = sum(b["smp"] for b in blocks)
print("total warp instructions %.4g, samples %d" % (tot, tsmp))
for b in blocks:
    share = 100 * b["ex"] * b["n"] / tot
    if share >= min_share:
        print("%s  n=%3d  exec=%.3g  share=%5.1f%%  samples=%5.1f%%  avg_threads=%4.1f  %s" % (b["first"][-5:], b["n"], b["ex"], share, 100 * b["smp"] / max(tsmp, 1),
              b["th"] / max(b["ex"] * b["n"], 1), " ".join("%s:%d" % kv for kv in b["ops"].most_common(8))))
