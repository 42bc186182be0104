"""Developer tool: the large-sample query parity of tests/test_gpu_parity.py::test_march_kernels_vs_reference_8k_rays as a table of
RATES (run on the GPU box; copy the output to profiles/).  Checker = oracle/_ref (the compiled, unmodified reference).

  python tools/parity_rates.py [rays, default 8192] [medium names, default tests/support.py PARITY_MEDIA] > gpurun_out/parity_rates.txt
"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import support as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
print("# GPU (libscn_b200.so through the C ABI) vs %s on %d seeded rays per case (tests/support.py: camera_rays seed 23, scene seed 9)" % (
    "oracle/_ref = the compiled UNMODIFIED reference" if S.have_ref() else "oracle/scn_oracle.c (no _ref on this box)", n))
print("# decision = ok/exited differ or hit > 1 march step away; step_moved = same decision, |dt| > 1e-4 rel (floor 1e-3) but within a step (H-sign);")
print("# matched = the rest: relative errors of t (floor 1e-3), gradient (floor: rms), the 7 conditioning coefficients (floor: rms of each)")
print("%-36s %-9s %8s %5s %5s %9s %9s %9s %9s %9s %9s %4s" % ("medium", "wavefront", "hit_frac", "dec", "moved", "t_max", "t_p999", "grad_max", "grad_p999", "ctx_max", "ctx_p999", "T!="))
rows = []
for name in (sys.argv[2:] or S.PARITY_MEDIA):
    for secondary in (False, True):
        r = S.query_parity_case(name, secondary, n=n)
        rows.append(r)
        print("%-36s %-9s %8.3f %5d %5d %9.2e %9.2e %9.2e %9.2e %9.2e %9.2e %4d" % (
            r["medium"], r["wavefront"], r["hit_fraction"], r["decision"], r["step_moved"], r["t_rel_max"], r["t_rel_p999"], r["grad_rel_max"], r["grad_rel_p999"],
            max(r["ctx_rel_max"]), max(r["ctx_rel_p999"]), r["transmittance_mismatch"]), flush=True)
print("# json")
for r in rows:
    print(json.dumps(r))
