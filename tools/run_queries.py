"""Developer tool: run the medium-query kernels on a seeded wavefront and print queries/s and noise evals/s
(CUDA events).  Used as the short command for `ncu` captures of the march kernels.

  python tools/run_queries.py [medium-name] [log2 rays] [repeats] [primary|secondary] [generic]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import support as S

name = sys.argv[1] if len(sys.argv) > 1 else "c1_sphere_3d_none"
n = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 18)
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
secondary = len(sys.argv) > 4 and sys.argv[4] == "secondary"

g = S.GpuImpl(S.MEDIA[name])
if "generic" in sys.argv[4:]:
    g.m.force_generic(True)  # the one-thread-per-ray kernels, for before / after rates
print(name, g.m.info()["query_kernel"], flush=True)
rays, st = S.camera_rays(n, seed=11, secondary=secondary)
stream = torch.cuda.current_stream()
out = None
for k in range(reps + 1):
    r, s = g._rays_state(rays, st)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    out = g.m.sample_distance(r, s, scene_seed=1, count_evals=True, out=out, stream=stream)
    b.record(stream)
    torch.cuda.synchronize()
    if k == 0:
        continue
    dt = a.elapsed_time(b) * 1e-3
    ev = int(out["n_evals"].view(torch.int32).to(torch.int64).sum().item())
    print("%s %s rays=%d  %.3f ms  %.3e queries/s  %.3e evals/s  %.1f evals/query  hit=%.3f" % (
        name, "sec" if secondary else "pri", n, dt * 1e3, n / dt, ev / dt, ev / n, float((out["exited"] == 0).float().mean())), flush=True)
r, s = g._rays_state(rays, st)
T = g.m.transmittance(r, s, scene_seed=1)
torch.cuda.synchronize()
print("transmittance mean %.4f" % float(T.mean()))
