"""Developer tool: the workload `compute-sanitizer` is run over (tools/gpu_job.sh sanitize).

One process, small wavefronts through the C ABI so that every kernel family runs under the tool:
  * k_march      one level     (C1 medium, primary + secondary + transmittance)
  * k_march      two levels    (C5 medium: multi-resolution, Renewal+ conditioning)
  * k_cond1d -> k_march1d -> k_finish1d with the stream-ordered scratch (C3 medium)
  * the generic kernels (a rotation-anisotropy medium), k_fast_noise3d, k_nee
  * one NEE render (C4 scene: raygen, march1d, nee_setup, 2 x transmittance, nee_gather, shade, clamp, accumulate)

  python tools/sanitize_workload.py [log2 rays, default 14]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import support as S
from sparse_conv_gpis_tungsten_b200 import render as R

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 14)
for name in ("c1_sphere_3d_none", "c5_multires_nonstat_renewalplus", "c3_1d_renewal", "aniso_rot_iso_const_none"):
    g = S.GpuImpl(S.MEDIA[name])
    for secondary in (False, True):
        rays, st = S.camera_rays(n if name != "aniso_rot_iso_const_none" else min(n, 2048), seed=11, secondary=secondary)
        q = g.sample_distance(rays, st, 1)
        T = g.transmittance(rays, st, 1)
        print("%s %s: hit fraction %.3f, T mean %.3f" % (name, "sec" if secondary else "pri", float(1.0 - np.mean(q["exited"])), float(np.mean(T))), flush=True)
g = S.GpuImpl(S.MEDIA["c2_micro_iso"])
p, d = S.eval_points(n)
print("noise eval:", float(np.abs(g.eval_noise(p, d, 1)).mean()), flush=True)
sc = S.appendix_a_scene(S.c4_scene_medium(), 32, 2, max_bounces=3, quad=True)
r = R.WavefrontRenderer(sc, device=0)
r.render(2)
torch.cuda.synchronize()
print("nee render mean radiance %.4f" % float(R.gather_image_rgb(r).mean()), flush=True)
r.close()
