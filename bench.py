#!/usr/bin/env python
"""bench.py — headline benchmark of the GPIS medium-query path (BASELINE.json metric, configs[1]).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU code on the host cores

Workload: 2^26 points uniform in [-8,8]^3 (seed 1234), unit directions (seed 1235); medium =
stationary squared-exponential sigma=1 lengthScale=1 aniso=[1,1,1] localScale=3, single_realization, seed 1,
impulse_density 3, isotropic-ray sampling; one step = value+gradient of evaluateNoise3D at every point.
`value` = noise evaluations / s with inputs resident in HBM (CUDA events on the launching stream);
`e2e` = the same through the C ABI's host-buffer entry point (pinned host memory, H2D + kernel + D2H timed);
`roofline` = SURVEY §8(d) algorithmic lane-ops per evaluation x evaluations / s against the FP32 issue peak at the
clock measured during the run (the path is issue-bound: 40 B and ~6e3 lane-ops per evaluation);
`cpu_baseline` = the compiled reference (oracle/_ref) or the oracle port on the host cores, bounded sample.
Each rank of an N-GPU run owns its own 2^26 points (weak scaling, no data-path collective).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_POINTS = 1 << 26
DENSITY = 3
ISO = True
METRIC = "noise_evals_per_sec"
UNIT = "evals/s"


def w_issue(n):  # SURVEY.md §8(d), frozen before the first measurement
    return 681 + 1803 * n


def w_flop(n):
    return 80 + 412 * n


def workload_name(density=DENSITY, iso=ISO):
    return "C2 noise-eval microbench: 2^26 points, value+gradient, 3D, impulse_density=%d, %s sampling" % (density, "isotropic-ray" if iso else "world-space")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.stop_flag = threading.Event()
        self.rows = []

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        # one nvidia-smi process in loop mode (a sample every 20 ms) instead of one process per sample: the timed region is short
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        for line in self.proc.stdout:
            if self.stop_flag.is_set():
                break
            line = line.strip()
            if line:
                self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        self.stop_flag.set()
        proc = getattr(self, "proc", None)
        if proc is not None:
            try:
                proc.terminate()  # the exact process this sampler started
                proc.wait(timeout=3)
            except Exception:
                pass
        self.join(timeout=3)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, nme in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_impl(medium_json, threads):
    import support as S
    kind = "ref" if S.have_ref() else "orc"
    return S.CpuImpl(kind, medium_json, threads=threads), ("reference" if kind == "ref" else "port")


def host_points(n, offset=0):
    """The first n points of the seeded stream, generated on the host (numpy mirrors of seeds 1234/1235)."""
    import support as S
    return S.eval_points(n, seed=1234 + 2 * offset)


def time_cpu(impl, p, d, repeats=1):
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        impl.eval_noise(p, d, 1)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best


def cpu_reference_render(scene, threads):
    """Times the unmodified reference renderer (oracle/_ref, ref_render) on `scene`; returns (camera paths / s, seconds)."""
    import ctypes as C
    import tempfile
    import support as S
    L = C.CDLL(S.REF_LIB)
    L.ref_last_error.restype = C.c_char_p
    res = scene["camera"]["resolution"]
    spp = scene["renderer"]["spp"]
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "scene.json")
        json.dump(scene, open(path, "w"))
        w, h, sec = C.c_int(), C.c_int(), C.c_double()
        rgb = np.zeros((res[1], res[0], 3), np.float32)
        rc = L.ref_render(path.encode(), threads, C.c_uint32(0xBA5EBA11), C.byref(w), C.byref(h), rgb.ctypes.data_as(C.c_void_p), C.c_size_t(rgb.size), C.byref(sec))
        if rc != 0:
            raise RuntimeError(L.ref_last_error().decode())
    return res[0] * res[1] * spp / sec.value, sec.value


def run_reference(args):
    """The reference arm: the reference's own CPU implementation of the path (oracle/_ref when it was compiled from
    /root/reference, else the oracle port), all host threads, bounded sample of the same workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import support as S
    threads = os.cpu_count() or 1
    mj = S.microbench_medium(DENSITY, ISO)
    impl, kind = cpu_impl(mj, threads)
    sample = 1 << 21
    p, d = host_points(sample)
    for _ in range(max(args.warmup, 1)):
        impl.eval_noise(p[: sample // 8], d[: sample // 8], 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        impl.eval_noise(p, d, 1)
    dt = time.perf_counter() - t0
    v = sample * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(), "points_per_step": sample, "note": "bounded sample of the 2^26-point workload per step"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": "2^21 of the 2^26 points per step, value+gradient"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import support as S
    from sparse_conv_gpis_tungsten_b200 import SparseConvolutionNoiseMedium, new_state

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this repo has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = args.points
    mj = S.microbench_medium(DENSITY, ISO)
    med = SparseConvolutionNoiseMedium.from_json(mj, device=local)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + 2 * rank)  # every rank owns its own 2^26 points
    P = [(torch.rand(n, generator=gen, device=dev) * 16 - 8) for _ in range(3)]
    gen.manual_seed(1235 + 2 * rank)
    Dm = torch.randn(3, n, generator=gen, device=dev)
    Dm = Dm / Dm.norm(dim=0, keepdim=True)
    D = [Dm[0].contiguous(), Dm[1].contiguous(), Dm[2].contiguous()]
    del Dm
    outs = [torch.empty(n, device=dev) for _ in range(4)]
    stream = torch.cuda.current_stream()

    def step():
        med.eval_noise_into(P, D, 1, outs, stream=stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.15)  # let nvidia-smi come up; its first samples then fall inside the timed region
    sampler.rows.clear()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    sampler.stop()
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * n * args.steps / (ms_max * 1e-3)

    # ---- end to end through the host-buffer entry point (rank-local; max over ranks)
    ne = min(n, args.e2e_points)
    hp = [x[:ne].cpu().pin_memory() for x in P]
    hd = [x[:ne].cpu().pin_memory() for x in D]
    ho = [torch.empty(ne, pin_memory=True) for _ in range(4)]
    med.eval_noise_into(hp, hd, 1, ho)  # grows the staging arena once
    med.eval_noise_into(hp, hd, 1, ho)
    if world > 1:
        dist.barrier()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        med.eval_noise_into(hp, hd, 1, ho)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * ne * e2e_steps / float(te.item())
    same = all(torch.equal(a, b[:ne].cpu()) for a, b in zip(ho, outs))

    # ---- end to end with PAGEABLE host buffers (what a caller gets who does not page-lock: every cudaMemcpyAsync is staged by the driver)
    e2e_pageable = None
    if not args.no_extras:
        npg = min(ne, 1 << 24)
        pp = [x[:npg].cpu() for x in P]; pd_ = [x[:npg].cpu() for x in D]; po = [torch.empty(npg) for _ in range(4)]
        med.eval_noise_into(pp, pd_, 1, po)
        t0 = time.perf_counter()
        med.eval_noise_into(pp, pd_, 1, po)
        torch.cuda.synchronize()
        e2e_pageable = {"value": npg / (time.perf_counter() - t0), "unit": UNIT, "points_per_step": npg, "note": "rank 0's figure; ordinary (pageable) numpy/torch host memory"}
        del pp, pd_, po

    # ---- medium queries / s (BASELINE metric, first half): 2^20 rays through the Appendix-A cube per BASELINE render config, primary
    #      (camera rays) and secondary (rays carrying Renewal state), median of 5 launches, device-resident; and the same C1 wavefront
    #      end to end through scn_sample_distance_batch_host from pinned host SoA buffers (max over ranks)
    queries, e2e_queries = None, None
    if not args.no_extras:
        queries = {}
        nq = 1 << 20
        qclock = ClockSampler(local); qclock.start()
        for key, name in (("C1", "c1_sphere_3d_none"), ("C3", "c3_1d_renewal"), ("C4", "c4_1d_renewalplus_xy_mis"), ("C5", "c5_multires_nonstat_renewalplus")):
            g = S.GpuImpl(S.MEDIA[name], device=local)
            row = {"medium": name, "kernel": g.m.info()["query_kernel"], "rays": nq}
            for secondary in (False, True):
                rr, ss = S.camera_rays(nq, seed=11, secondary=secondary)
                r, s_ = g._rays_state(rr, ss)
                out = g.m.sample_distance(r, s_, scene_seed=1, count_evals=True)
                torch.cuda.synchronize()
                times = []
                for _ in range(5):
                    r, s_ = g._rays_state(rr, ss)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(stream); out = g.m.sample_distance(r, s_, scene_seed=1, out=out, stream=stream); b.record(stream); torch.cuda.synchronize()
                    times.append(a.elapsed_time(b) * 1e-3)
                qs = float(np.median(times))
                tq = torch.tensor([qs], device=dev, dtype=torch.float64)
                if world > 1:
                    allq = [torch.zeros_like(tq) for _ in range(world)]
                    dist.all_gather(allq, tq)
                    row[("secondary" if secondary else "primary") + "_per_rank_ms"] = [round(float(t_.item()) * 1e3, 3) for t_ in allq]
                    dist.all_reduce(tq, op=dist.ReduceOp.MAX)
                qs = float(tq.item())
                ev = int(out["n_evals"].view(torch.int32).to(torch.int64).sum().item())
                tag = "secondary" if secondary else "primary"
                row[tag + "_queries_per_s"] = world * nq / qs
                row[tag + "_noise_evals_per_s"] = world * ev / qs
                row[tag + "_evals_per_query"] = ev / nq
                if not secondary:
                    out_primary = {"t": out["t"].clone()}
            queries[key] = row
            if key == "C1":  # end to end: pinned host SoA -> scn_sample_distance_batch_host -> pinned host SoA
                gh = S.GpuImpl(S.MEDIA[name], device=local, host=True)
                rr, ss = S.camera_rays(nq, seed=11)
                hr0, hs0 = gh._rays_state(rr, ss)  # numpy SoA
                pin = lambda a_: torch.from_numpy(a_.view(np.int32) if a_.dtype == np.uint32 else a_).pin_memory()
                hr = {k: pin(v) for k, v in hr0.items()}
                tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8}
                from sparse_conv_gpis_tungsten_b200.medium import SAMPLE_FIELDS
                ho_q = {k: torch.empty(nq, dtype=tdt[dt]).pin_memory() for k, dt in SAMPLE_FIELDS}
                tt = []
                for k in range(4):
                    hs = {kk: pin(v.copy()) for kk, v in hs0.items()}  # the state is updated in place: a fresh copy per step
                    if world > 1:
                        dist.barrier()
                    t0 = time.perf_counter()
                    gh.m.sample_distance(hr, hs, scene_seed=1, out=ho_q)
                    tt.append(time.perf_counter() - t0)
                te2 = torch.tensor([float(np.median(tt[1:]))], device=dev, dtype=torch.float64)
                if world > 1:
                    dist.all_reduce(te2, op=dist.ReduceOp.MAX)
                nbytes = lambda d_: int(sum(v.numel() * v.element_size() for v in d_.values()))
                e2e_queries = {"value": world * nq / float(te2.item()), "unit": "queries/s", "h2d_bytes_per_step": nbytes(hr) + nbytes({k: v for k, v in hs.items() if k != "scheme" and not k.startswith("ctx_")}),
                               "d2h_bytes_per_step": nbytes(ho_q) + nbytes(hs), "rays_per_step": nq, "matches_device": bool(torch.equal(ho_q["t"], out_primary["t"].cpu())),
                               "api": "scn_sample_distance_batch_host (C1 medium, 2^20 camera rays; pinned host SoA rays + state in, samples + state out; one stream per stage)"}
                gh.close()
            g.close()
        qclock.stop()
        qck = torch.tensor([qclock.summary()["sm_mhz"] or 0.0], device=dev, dtype=torch.float64)
        if world > 1:
            allc = [torch.zeros_like(qck) for _ in range(world)]
            dist.all_gather(allc, qck)
            queries["per_rank_sm_mhz"] = [float(t_.item()) for t_ in allc]
        else:
            queries["per_rank_sm_mhz"] = [float(qck.item())]
        queries["clock_reasons"] = qclock.summary()["reasons"]

    # ---- tile-sharded wavefront renders at the BASELINE.json sizes: every rank renders its tiles (round-robin ownership), one NCCL
    #      all_gather of the slices.  Strong scaling: the same image at every N.
    renders = None
    if not args.no_extras:
        from sparse_conv_gpis_tungsten_b200 import render as R

        def timed_render(sc, spp, label, sharding="interleaved"):
            wr = R.WavefrontRenderer(sc, device=local, rank=rank, world=world, sharding=sharding); wr.render(1); R.gather_image_rgb(wr, world); wr.close()  # warm-up (kernels, NCCL all_gather connections)
            wr = R.WavefrontRenderer(sc, device=local, rank=rank, world=world, sharding=sharding)
            wr.prepare(spp)  # wavefront buffers allocated outside the timed region
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            a, b, c_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            # (no nvidia-smi polling inside this region: a 20 ms poll loop contends with the render's ~500 launches per second for the driver
            #  and cost the C4 render 12 % when it was tried; the clocks of the run are sampled in the headline and the query legs)
            a.record(stream); wr.render(spp); c_.record(stream); img = R.gather_image_rgb(wr, world); b.record(stream); torch.cuda.synchronize()
            tr = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
            own = torch.tensor([a.elapsed_time(c_)], device=dev, dtype=torch.float64)  # this rank's tiles, before the gather
            per_rank = [float(own[0].item())]
            if world > 1:  # the render is as slow as its slowest rank: load balance and per-GPU speed are both in here
                allt = [torch.zeros_like(own) for _ in range(world)]
                dist.all_gather(allt, own)
                per_rank = [float(t_[0].item()) for t_ in allt]
                dist.all_reduce(tr, op=dist.ReduceOp.MAX)
            res = sc["camera"]["resolution"]
            out = {"workload": "%s %dx%d @%dspp, max_bounces %d, nee %s, 16x16 tiles over %d GPU(s) (%s), one all_gather" % (label, res[0], res[1], spp, sc["integrator"]["max_bounces"], "on" if wr.nee else "off", world, wr.sharding),
                   "ms": float(tr.item()), "camera_paths": res[0] * res[1] * spp, "camera_paths_per_s": res[0] * res[1] * spp / (float(tr.item()) * 1e-3), "mean_radiance": float(img.mean().item()),
                   "per_rank_render_ms": [round(v, 3) for v in per_rank]}
            wr.close()
            return out

        q = args.render_scale  # 1 = the BASELINE.json sizes; smaller values shrink the sample counts (developer runs), stated in the workload string
        spp_of = lambda full: max(1, int(round(full * q)))
        renders = {
            "C1": timed_render(S.appendix_a_scene(S.MEDIA["c1_sphere_3d_none"], 256, 16), 16, "C1 scene (Appendix A.1)"),
            "C3": timed_render(S.appendix_a_scene(S.MEDIA["c3_1d_renewal"], 512, spp_of(64)), spp_of(64), "C3 scene (Appendix A.2: 1D sampling, renewal, ensemble)"),
            "C4": timed_render(S.appendix_a_scene(S.c4_scene_medium(), 1024, spp_of(128), quad=True), spp_of(128), "C4 scene (Appendix A.3: 1D Renewal+ XY, copper conductor, quad light, mis)"),
            "C5": timed_render(S.appendix_a_scene(S.MEDIA["c5_multires_nonstat_renewalplus"], 2048, spp_of(256)), spp_of(256), "C5 scene (Appendix A.4: multi-resolution non-stationary, renewal+)"),
        }
        if world > 1:  # the contiguous ranges of SURVEY 8(e) next to the round-robin ownership used above, C5 scene at 1/16 of the samples
            sc5 = S.appendix_a_scene(S.MEDIA["c5_multires_nonstat_renewalplus"], 2048, 16)
            renders["C5_sharding"] = {"interleaved_ms": timed_render(sc5, 16, "C5", "interleaved")["ms"], "contiguous_ms": timed_render(sc5, 16, "C5", "contiguous")["ms"], "workload": "C5 scene 2048x2048 @16spp"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    clocks = sampler.summary()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    props = torch.cuda.get_device_properties(local)
    sms = props.multi_processor_count
    f_mhz = clocks["sm_mhz"] or peaks.get("sm_max_mhz") or 1965.0
    per_gpu = value / world
    issue_peak = sms * 4 * 32 * f_mhz * 1e6            # lane-ops / s at the clock seen during the run
    fp32_peak = sms * 128 * 2 * f_mhz * 1e6
    ach_issue = per_gpu * w_issue(DENSITY)
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed ncu --set full capture, scaled to this launch size
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["k_fast_noise3d<ISO=%d, NIMP=%d>" % (int(ISO), DENSITY)]
        traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * (n / tj["evals_per_launch"])
    except Exception:
        pass
    roofline = {"bound": "fp32-issue", "achieved": ach_issue / 1e12, "peak": issue_peak / 1e12, "unit": "Tlane-op/s", "frac": ach_issue / issue_peak,
                "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read+write; algorithmic = 40 B x evals_per_launch)", "kernel": "k_fast_noise3d<ISO=%d, NIMP=%d>" % (int(ISO), DENSITY), "avg_launch_ms": ms_max / args.steps,
                "algorithmic": {"lane_ops_per_eval": w_issue(DENSITY), "flops_per_eval": w_flop(DENSITY), "bytes_per_eval": 40, "evals_per_launch": n},
                "fp32_flops": {"achieved": per_gpu * w_flop(DENSITY) / 1e12, "peak": fp32_peak / 1e12, "unit": "TFLOP/s", "frac": per_gpu * w_flop(DENSITY) / fp32_peak},
                "hbm": {"achieved": per_gpu * 40 / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": per_gpu * 40 / 1e9 / hbm_peak,
                        "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback"},
                "peak_source": "SMs x 4 schedulers x 32 lanes x SM clock sampled during the timed region (SURVEY.md 8(d)); sm_count=%d, MHz=%.0f" % (sms, f_mhz)}

    # ---- density sweep of the microbench (BASELINE configs[1]: impulse_density 1-64, isotropic vs world space); one timed pass each
    sweep = None
    if not args.no_extras:
        sweep = {}
        for iso in (True, False):
            for dens in (1, 2, 4, 8, 16, 32, 64):
                m2 = SparseConvolutionNoiseMedium.from_json(S.microbench_medium(dens, iso), device=local)
                m2.eval_noise_into(P, D, 1, outs, stream=stream)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream); m2.eval_noise_into(P, D, 1, outs, stream=stream); b.record(stream); torch.cuda.synchronize()
                ev = n / (a.elapsed_time(b) * 1e-3)
                sweep["%s_N%d" % ("iso" if iso else "world", dens)] = {"evals_per_s": ev, "issue_frac": ev * w_issue(dens) / issue_peak}
                m2.close()

    # ---- what ncu counted for each kernel family (committed captures, profiles/ncu_issue_active.json): the model fraction above charges
    #      SURVEY's frozen 6090 lane-ops per evaluation, the kernel executes fewer; for the march kernels only the ncu figure applies
    ncu_counted = None
    try:
        ncu_counted = json.load(open(os.path.join(ROOT, "profiles", "ncu_issue_active.json")))
    except Exception:
        pass
    if ncu_counted:
        roofline["issue_active_ncu"] = {k: v.get("issue_active_pct") for k, v in ncu_counted.items() if isinstance(v, dict)}
        kf = ncu_counted.get(roofline["kernel"])
        if kf and kf.get("lane_instr_per_eval"):
            roofline["frac_ncu_counted"] = per_gpu * kf["lane_instr_per_eval"] / issue_peak
            roofline["note"] = "frac = SURVEY 8(d) model (6090 lane-ops/eval, frozen); frac_ncu_counted = the %d lane-instructions/eval ncu counted x this run's rate; issue_active_ncu = smsp__issue_active of the committed captures" % kf["lane_instr_per_eval"]

    # ---- CPU baseline on this box's host cores (bounded samples)
    threads = os.cpu_count() or 1
    if renders is not None and S.have_ref():
        # the reference's own multithreaded renderer (Scene::load -> PathTraceIntegrator through oracle/ref_harness.cpp: ref_render) on a
        # bounded sample of each scene, all host threads
        for key, sc, what in (("C1", S.appendix_a_scene(S.MEDIA["c1_sphere_3d_none"], 64, 4), "64x64 @4spp"),
                              ("C3", S.appendix_a_scene(S.MEDIA["c3_1d_renewal"], 64, 16), "64x64 @16spp"),
                              ("C4", S.appendix_a_scene(S.c4_scene_medium(), 64, 16, quad=True), "64x64 @16spp"),
                              ("C5", S.appendix_a_scene(S.MEDIA["c5_multires_nonstat_renewalplus"], 48, 4), "48x48 @4spp")):
            try:
                ref_paths_s, ref_sec = cpu_reference_render(sc, threads)
                renders[key]["cpu_reference"] = {"camera_paths_per_s": ref_paths_s, "cores": threads, "seconds": ref_sec, "kind": "reference",
                                                 "sample": "the same scene at %s through the unmodified reference renderer (scene_bvh off)" % what}
                renders[key]["speedup_vs_cpu_reference"] = renders[key]["camera_paths_per_s"] / ref_paths_s
            except Exception as ex:  # the checker must never take the bench line down
                renders[key]["cpu_reference"] = {"unavailable": str(ex)[:200]}
    impl, kind = cpu_impl(mj, threads)
    cs = 1 << 22
    cp, cd = host_points(cs)
    impl.eval_noise(cp[: cs // 16], cd[: cs // 16], 1)
    cdt = time_cpu(impl, cp, cd, repeats=1)
    cpu_v = cs / cdt
    # and the checker: the same sample through the GPU must agree with the CPU arm
    gp = [torch.from_numpy(np.ascontiguousarray(cp[:65536, j])).to(dev) for j in range(3)]
    gd = [torch.from_numpy(np.ascontiguousarray(cd[:65536, j])).to(dev) for j in range(3)]
    go = torch.stack(med.eval_noise(gp, gd, 1), 1).cpu().numpy()
    want = impl.eval_noise(np.ascontiguousarray(cp[:65536]), np.ascontiguousarray(cd[:65536]), 1)
    parity = float(S.rel_err(go, want).max())

    config = {"workload": workload_name(), "points_per_gpu_per_step": n, "impulse_density": DENSITY, "isotropic_3D_sampling": ISO,
              "l2": "inputs (%.1f GB) and outputs (%.1f GB) per step exceed the 126 MB L2" % (24 * n / 1e9, 16 * n / 1e9), "parallelism": "replicas x%d (weak)" % world,
              "parity_vs_cpu_max_rel_err": parity}
    if queries is not None:  # the other half of BASELINE's metric: medium queries / s per render config (whole job: all ranks' rays / max time)
        config["medium_queries"] = queries
    if renders is not None:  # strong scaling: the same image at every N
        config["renders"] = renders
    if sweep is not None:
        config["density_sweep"] = sweep
    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 24 * ne, "d2h_bytes_per_step": 16 * ne, "points_per_step": ne, "matches_device": bool(same),
           "api": "scn_eval_batch_host (pinned host buffers; H2D, kernel and D2H each on their own stream, chunks chained by events)"}
    if e2e_pageable is not None:
        e2e["pageable"] = e2e_pageable
    if e2e_queries is not None:
        e2e["queries"] = e2e_queries
    launches = args.steps
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": threads, "kind": kind, "sample": "first 2^22 of the 2^26 points, value+gradient, one pass"},
            "e2e": e2e, "gpu_launches": launches}
    if world > 1:
        dist.destroy_process_group()
    return line


class StdoutToStderr:
    """Anything libraries print on fd 1 during the run (NCCL's version banner, ...) goes to stderr: stdout carries ONE JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=N_POINTS)
    ap.add_argument("--e2e-points", type=int, default=N_POINTS)
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--render-scale", type=float, default=1.0, help="1 = BASELINE.json sample counts (C3 64, C4 128, C5 256 spp); smaller for developer runs")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    with StdoutToStderr():
        line = run_gpu(args)
    if isinstance(line, dict):
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
