import os, sys, torch
sys.path.insert(0, "tests")
import support as S
mode = sys.argv[1]
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if mode == "dist":
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dist.barrier()
g = S.GpuImpl(S.MEDIA["c1_sphere_3d_none"], device=local)
rays, st = S.camera_rays(1 << 20, seed=11)
stream = torch.cuda.current_stream()
out = None; ts = []
for k in range(6):
    r, s = g._rays_state(rays, st)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream); out = g.m.sample_distance(r, s, scene_seed=1, count_evals=True, out=out, stream=stream); b.record(stream); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
print(mode, "rank", local, "env OMP", os.environ.get("OMP_NUM_THREADS"), "ms", [round(t, 2) for t in ts[1:]], flush=True)
