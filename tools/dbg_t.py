"""Developer tool: worst hit-distance disagreements between the GPU query path and the CPU oracle on a seeded wavefront."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import support as S
name = sys.argv[1]; sec = "secondary" in sys.argv[2:]; generic = "generic" in sys.argv[2:]
rays, st = S.camera_rays(8192, seed=23, secondary=sec)
gpu, orc = S.GpuImpl(S.MEDIA[name]), S.checker(S.MEDIA[name])
if generic:
    gpu.m.force_generic(True)
a, b = gpu.sample_distance(rays, st, 9), orc.sample_distance(rays, st, 9)
m = (a["ok"] == b["ok"]) & (a["exited"] == b["exited"]) & (a["ok"] != 0)
err = np.abs(a["t"].astype(np.float64) - b["t"]) / np.maximum(np.abs(b["t"]), 1e-3)
err[~m] = 0
idx = np.argsort(-err)[:6]
print(name, "secondary" if sec else "primary", "generic kernels" if generic else gpu.m.info()["query_kernel"])
for i in idx:
    print("  ray %5d  t_gpu %.9g  t_cpu %.9g  abs %.3g  rel %.3g  exited %d  tmin %.4g tmax %.4g" % (i, a["t"][i], b["t"][i], abs(float(a["t"][i]) - float(b["t"][i])), err[i], b["exited"][i], rays["tmin"][i], rays["tmax"][i]))
print("  hist of abs err among hits:", np.percentile(np.abs(a["t"].astype(np.float64) - b["t"])[m & (b["exited"] == 0)], [50, 90, 99, 99.9, 100]))
