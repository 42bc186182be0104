"""Developer tool: what the host link gives (H2D alone, D2H alone, both at once, pinned memory) next to the end-to-end
rate of scn_eval_batch_host for a sweep of pipeline-stage caps (SCN_HOST_CHUNK_LOG2) and two batch sizes.  `python tools/pcie_probe.py`."""
import os, subprocess, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def link():
    import torch
    n = 1 << 28
    h_in, h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True), torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d_in, d_out = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.zeros(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, down, reps=8):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
            if down:
                with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize(); return reps * n / (time.perf_counter() - t0) / 1e9
    run(True, True, 2)
    print("H2D alone %.1f GB/s, D2H alone %.1f GB/s, concurrent %.1f GB/s each way" % (run(True, False), run(False, True), run(True, True)))


def e2e():
    import torch, support as S
    from sparse_conv_gpis_tungsten_b200.medium import SparseConvolutionNoiseMedium as Medium
    med = Medium.from_json(S.MEDIA["c1_sphere_3d_none"])
    n = 1 << 26
    g = torch.Generator().manual_seed(1)
    hp = [((torch.rand(n, generator=g) - 0.5) * 40).pin_memory() for _ in range(3)]
    d = torch.randn(3, n, generator=g); d /= d.norm(dim=0)
    hd = [d[i].contiguous().pin_memory() for i in range(3)]
    ho = [torch.empty(n, pin_memory=True) for _ in range(4)]
    med.eval_noise_into(hp, hd, 1, ho); med.eval_noise_into(hp, hd, 1, ho)
    for m in (n, n >> 4):
        for l in (20, 21, 22, 23):
            os.environ["SCN_HOST_CHUNK_LOG2"] = str(l)
            a = [[x[:m] for x in v] for v in (hp, hd, ho)]
            ts = []
            for _ in range(6):
                t0 = time.perf_counter(); med.eval_noise_into(a[0], a[1], 1, a[2]); ts.append(time.perf_counter() - t0)
            print("2^%d points, stage cap 2^%d: best %.3f median %.3f worst %.3f G evals/s" % (m.bit_length() - 1, l, m / min(ts) / 1e9, m / sorted(ts)[3] / 1e9, m / max(ts) / 1e9), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "e2e": e2e()
    else:
        link()
        e2e()
