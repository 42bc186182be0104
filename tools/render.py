"""Render a restricted GPIS scene with the wavefront stage, tiles sharded over the ranks of a
torchrun launch, one NCCL all_gather of the framebuffer slices at the end.  Prints one JSON line (rank 0).

  python tools/render.py --config c1 --res 256 --spp 16 [--out gpurun_out/c1.pfm]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/render.py --config c5 --res 2048 --spp 256
"""
import argparse, json, os, sys, time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support as S
from sparse_conv_gpis_tungsten_b200 import render as R

CONFIGS = {"c1": "c1_sphere_3d_none", "c5": "c5_multires_nonstat_renewalplus", "renewal": "iso_3d_renewal", "world": "world_3d_none"}


def scene(config, res, max_bounces):
    return {"media": [S.MEDIA[CONFIGS[config]]],
            "primitives": [{"type": "cube", "transform": {"scale": [3, 3, 3]}, "bsdf": {"type": "forward"}, "int_medium": "gpis"},
                           {"type": "infinite_sphere", "emission": 1.0, "sample": True}],
            "camera": {"type": "pinhole", "resolution": [res, res], "fov": 35, "transform": {"position": [0, 0, 5], "look_at": [0, 0, 0], "up": [0, 1, 0]}},
            "integrator": {"type": "path_tracer", "min_bounces": 0, "max_bounces": max_bounces},
            "renderer": {"spp": 1, "adaptive_sampling": False}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c1", choices=sorted(CONFIGS)); ap.add_argument("--res", type=int, default=256); ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--max-bounces", type=int, default=8); ap.add_argument("--out", default=None); ap.add_argument("--warmup-spp", type=int, default=1)
    ap.add_argument("--max-paths", type=int, default=1 << 22); ap.add_argument("--count-queries", action="store_true")
    a = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc = scene(a.config, a.res, a.max_bounces)
    r = R.WavefrontRenderer(sc, device=local, rank=rank, world=world, max_paths=a.max_paths)
    if a.warmup_spp:
        w = R.WavefrontRenderer(sc, device=local, rank=rank, world=world, max_paths=a.max_paths); w.render(a.warmup_spp); w.close()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r.render(a.spp, count_queries=a.count_queries); img = R.gather_image(r, world); e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64); q = torch.tensor([float(r.queries)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(q)
    if rank == 0:
        paths = a.res * a.res * a.spp
        line = {"workload": "wavefront render %s %dx%d @%dspp, max_bounces %d" % (a.config, a.res, a.res, a.spp, a.max_bounces), "n_gpus": world, "ms": float(ms.item()),
                "camera_paths_per_s": paths / (ms.item() * 1e-3), "mean_radiance": float(img.mean().item()), "tiles": r.n_tiles}
        if a.count_queries:
            line["medium_queries"] = q.item(); line["medium_queries_per_s"] = q.item() / (ms.item() * 1e-3)
        print(json.dumps(line), flush=True)
        if a.out:
            R.write_pfm(a.out, img.cpu().numpy())
    r.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
