"""Developer tool: per CUDA source line view of an ncu capture (needs -lineinfo and --import-source on): executed warp
instructions, distinct SASS instructions (code footprint), stall samples and average active threads per line.
  python tools/ncu_lines.py gpurun_out/prof.ncu-rep [min_share_pct] [hot_exec_threshold]"""
import csv, subprocess, sys, collections, os
rep = sys.argv[1]; min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.7
HOT = float(sys.argv[3]) if len(sys.argv) > 3 else 1e6
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
cur_file = None; hdr = None; agg = collections.OrderedDict(); seen = set()
def num(x):
    try: return float(x)
    except Exception: return 0.0
key = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = os.path.basename(r[1]); continue
    if r[0] == "Line No": hdr = r; ci = {}; [ci.setdefault(k, i) for i, k in enumerate(hdr)]; addr_i = hdr.index("Address"); continue
    if hdr is None or len(r) != len(hdr): continue
    if r[0].isdigit():  # a source line: the SASS rows that follow belong to it
        key = (cur_file, int(r[0])); agg.setdefault(key, {"src": r[1].strip(), "n": 0, "hot": 0, "ex": 0.0, "th": 0.0, "smp": 0.0, "noinst": 0.0, "wait": 0.0}); continue
    if key is None or not r[addr_i].startswith("0x") or r[addr_i] in seen: continue
    seen.add(r[addr_i])
    a = agg[key]; e = num(r[ci["Instructions Executed"]])
    a["n"] += 1; a["ex"] += e; a["th"] += num(r[ci["Thread Instructions Executed"]]); a["smp"] += num(r[ci["# Samples"]])
    a["hot"] += 1 if e >= HOT else 0
    a["noinst"] += num(r[ci["stall_no_inst"]]); a["wait"] += num(r[ci["stall_wait"]])
tot = sum(a["ex"] for a in agg.values()); tsmp = sum(a["smp"] for a in agg.values())
print("total executed warp instructions %.4g, SASS instructions %d (hot, executed >= %g times: %d = %.1f KB), samples %d" % (tot, sum(a["n"] for a in agg.values()), HOT, sum(a["hot"] for a in agg.values()), sum(a["hot"] for a in agg.values()) * 16 / 1024, tsmp))
print("%-22s %5s %4s %6s %7s %7s %6s %6s %5s  %s" % ("file:line", "sass", "hot", "share", "samples", "no_inst", "wait", "exec/i", "thr", "source"))
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1]["ex"]):
    sh = 100 * a["ex"] / tot
    if sh < min_share: continue
    print("%-22s %5d %4d %5.1f%% %6.1f%% %6.1f%% %5.1f%% %6.2g %5.1f  %s" % ("%s:%d" % (f, ln), a["n"], a["hot"], sh, 100 * a["smp"] / max(tsmp, 1), 100 * a["noinst"] / max(tsmp, 1), 100 * a["wait"] / max(tsmp, 1),
                                                             a["ex"] / a["n"], a["th"] / max(a["ex"], 1), a["src"][:110]))
