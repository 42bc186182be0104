"""GPU leg (`pytest -m gpu`, B200): the CUDA kernels, called through the C ABI, against (1) the golden vectors recorded
from the unmodified reference, (2) the oracle on fresh seeded inputs, (3) size-independent properties at BASELINE.json's
full microbench size.  Integer work (hash, PCG, impulse positions, accepted-impulse counts) must be bit-exact;
floating-point results within 1e-4 relative (north_star), with decision flips of the march counted separately."""
import ctypes as C

import numpy as np
import pytest

import golden_util as G
import support as S

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def _dev(torch, a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).cuda()


def test_hash_pcg_impulses_bit_exact(torch_cuda, built_library):
    torch = torch_cuda
    lib = built_library
    from sparse_conv_gpis_tungsten_b200 import abi
    g = dict(np.load(S.GOLDEN_DIR + "/primitives.npz"))
    keys = g["keys"]
    cols = [_dev(torch, keys[:, j]) for j in range(4)]
    out = torch.zeros(len(keys), dtype=torch.int32, device="cuda")
    p = lambda t, ct: C.cast(C.c_void_p(t.data_ptr()), ct)
    abi.check(lib.scn_debug_hash4(*[p(c, abi.u32p) for c in cols], p(out, abi.u32p), len(keys), None))
    assert np.array_equal(out.cpu().numpy().view(np.uint32), g["h4"])
    seeds = _dev(torch, g["seeds"])
    st = torch.zeros(len(g["seeds"]) * 16, dtype=torch.float32, device="cuda")
    abi.check(lib.scn_debug_pcg_stream(p(seeds, abi.u32p), p(st, abi.f32p), len(g["seeds"]), 16, None))
    assert np.array_equal(st.cpu().numpy().reshape(-1, 16), g["streams"])  # set_state(seed) then 16 draws
    # per-cell impulses against the oracle, including negative (wrapped) cell indices
    orc = C.CDLL(S.build_oracle())
    rng = np.random.default_rng(9)
    cells = rng.integers(-50, 50, (512, 3)).astype(np.int32).view(np.uint32)
    count = 9
    got = torch.zeros(512 * count * 4, dtype=torch.float32, device="cuda")
    cc = [_dev(torch, cells[:, j]) for j in range(3)]
    abi.check(lib.scn_debug_cell_impulses(p(cc[0], abi.u32p), p(cc[1], abi.u32p), p(cc[2], abi.u32p), 12345, count, p(got, abi.f32p), 512, None))
    want = np.zeros((512, count * 4), np.float32)
    for i, c in enumerate(cells):
        orc.orc_cell_impulses(C.c_uint32(int(c[0])), C.c_uint32(int(c[1])), C.c_uint32(int(c[2])), C.c_uint32(12345), C.c_uint32(count), want[i].ctypes.data_as(C.c_void_p))
    assert np.array_equal(got.cpu().numpy().reshape(512, -1), want)


@pytest.mark.parametrize("name", sorted(S.MEDIA))
def test_kernels_vs_reference_golden(torch_cuda, name):
    impl = S.GpuImpl(S.MEDIA[name])
    G.check_against_golden(impl, name, G.load(name), exact=False, allow_mismatch=1)
    impl.close()


@pytest.mark.parametrize("name", ["c1_sphere_3d_none", "c3_1d_renewal"])
def test_host_buffer_entry_points(torch_cuda, name):
    """The *_host entry points (numpy in/out, staged through the medium's arena) give the same answers."""
    g = G.load(name)
    dev, host = S.GpuImpl(S.MEDIA[name]), S.GpuImpl(S.MEDIA[name], host=True)
    e1 = dev.eval(g["p"], g["d"], g["tt"], g["pix"], 77, g["info_t"], g["coeffs"]); e2 = host.eval(g["p"], g["d"], g["tt"], g["pix"], 77, g["info_t"], g["coeffs"])
    assert np.array_equal(e1["val"], e2["val"]) and np.array_equal(e1["grad"], e2["grad"])
    for tag in ("pri", "sec"):
        rays, st, _ = G.rays_state(g, tag)
        a, b = dev.sample_distance(rays, st, 99), host.sample_distance(rays, st, 99)
        for k in ("ok", "exited", "t", "grad", "coeffs", "n_evals"):
            assert np.array_equal(a[k], b[k]), k
        for k in a["state"]:
            assert np.array_equal(a["state"][k], b["state"][k]), k
        assert np.array_equal(dev.transmittance(rays, st, 99), host.transmittance(rays, st, 99))


@pytest.mark.parametrize("iso", [True, False])
@pytest.mark.parametrize("density", [1, 2, 4, 8, 16, 32, 64])
def test_microbench_noise_vs_reference(torch_cuda, iso, density):
    """BASELINE.json configs[1]: value + gradient of the 3D noise, impulse_density sweep, isotropic vs world space."""
    n = 16384 if density <= 8 else 4096
    mj = S.microbench_medium(density, iso)
    p, d = S.eval_points(n)
    gpu, orc, chk = S.GpuImpl(mj), S.CpuImpl("orc", mj), S.checker(mj)  # values against the compiled reference when it is on the box
    a, b = gpu.eval_noise(p, d, 1), chk.eval_noise(p, d, 1)
    assert S.rel_err(a, b).max() <= TOL
    # accepted-impulse sets: counts must agree exactly (target 0 differing evals); the reference has no counter, the restatement
    # (bit-identical values, tests/test_oracle.py) counts
    torch = torch_cuda
    from sparse_conv_gpis_tungsten_b200 import abi
    cols = [_dev(torch, p[:, j]) for j in range(3)] + [_dev(torch, d[:, j]) for j in range(3)]
    cnt = torch.zeros(n, dtype=torch.int32, device="cuda")
    pp = lambda t, ct: C.cast(C.c_void_p(t.data_ptr()), ct)
    abi.check(abi.load().scn_debug_accepted_counts(gpu.m._h, *[pp(c, abi.f32p) for c in cols], 1, pp(cnt, abi.u32p), n, None))
    want = np.zeros(n, np.uint32)
    orc.lib.orc_accepted_counts(orc.hp, C.c_size_t(n), p.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), C.c_uint32(1), want.ctypes.data_as(C.c_void_p), 8)
    assert np.array_equal(cnt.cpu().numpy().view(np.uint32), want)
    assert abs(want.mean() / (27 * density) - 4 * np.pi / 81) < 0.01  # a = 4*pi/81 of the impulses are inside


def test_microbench_full_size_properties(torch_cuda):
    """2^26 points (the size BASELINE.json quotes): determinism, agreement with the oracle on a strided subsample,
    unit variance (the reference normalises by sqrt(variance3D)), gradient consistent with central differences."""
    torch = torch_cuda
    from sparse_conv_gpis_tungsten_b200 import SparseConvolutionNoiseMedium
    n = 1 << 26
    mj = S.microbench_medium(3, True)
    m = SparseConvolutionNoiseMedium.from_json(mj)
    gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
    P = [(torch.rand(n, generator=gen, device="cuda") * 16 - 8) for _ in range(3)]
    gen.manual_seed(1235)
    D = torch.randn(3, n, generator=gen, device="cuda"); D = D / D.norm(dim=0, keepdim=True); D = [D[0].contiguous(), D[1].contiguous(), D[2].contiguous()]
    o1 = m.eval_noise(P, D, 1); o2 = m.eval_noise(P, D, 1)
    for a, b in zip(o1, o2):
        assert torch.equal(a, b)
    assert bool(torch.isfinite(o1[0]).all())
    var = float(o1[0].double().var())
    # one realization of finite extent (the rotated lattices of all rays share the ~13^3 cells around the origin), so the
    # spatial statistics fluctuate by a few per cent; the compiled reference gives var 0.962, mean 0.055 on these points
    assert abs(var - 1.0) < 0.1, var
    assert abs(float(o1[0].double().mean())) < 0.1
    idx = torch.arange(0, n, n // 2048, device="cuda")
    p = torch.stack([x[idx] for x in P], 1).cpu().numpy(); d = torch.stack([x[idx] for x in D], 1).cpu().numpy()
    orc = S.checker(mj)
    want = orc.eval_noise(np.ascontiguousarray(p), np.ascontiguousarray(d), 1)
    got = torch.stack([x[idx] for x in o1], 1).cpu().numpy()
    assert S.rel_err(got, want).max() <= TOL
    # gradient vs central differences of the value along x (same ray direction => same rotated lattice)
    h = 1e-3
    sub = [x[:4096].contiguous() for x in P]; dsub = [x[:4096].contiguous() for x in D]
    vp = m.eval_noise([sub[0] + h, sub[1], sub[2]], dsub, 1)[0]; vm = m.eval_noise([sub[0] - h, sub[1], sub[2]], dsub, 1)[0]
    fd = (vp - vm) / (2 * h)
    gx = m.eval_noise(sub, dsub, 1)[1]
    ok = (fd - gx).abs() < 2e-2 * (1 + gx.abs())
    assert float(ok.float().mean()) > 0.98  # the kernel is truncated at its radius: a few points straddle a cut-off


def test_march_round_trip_properties(torch_cuda):
    """Size-independent properties of the query kernels on 2^18 rays: every reported hit brackets a sign change of the
    field (evaluateValue at t is on sign0's side, one step later is not), transmittance is 1 exactly where
    sampleDistance exits, and a replayed wavefront is bit-identical."""
    n = 1 << 18
    mj = S.MEDIA["c1_sphere_3d_none"]
    rays, st = S.camera_rays(n, seed=3)
    gpu = S.GpuImpl(mj)
    a = gpu.sample_distance(rays, st, 5); b = gpu.sample_distance(rays, st, 5)
    for k in ("ok", "exited", "t", "grad"):
        assert np.array_equal(a[k], b[k])
    T = gpu.transmittance(rays, st, 5)
    assert np.array_equal(T == 1.0, a["exited"] != 0)
    hit = (a["ok"] != 0) & (a["exited"] == 0) & (a["t"] > 0)
    assert hit.mean() > 0.3
    # gradient at a hit points against the ray (GPM:303-308 rejects the others)
    gd = (a["grad"] * rays["d"]).sum(1)
    assert (gd[hit] <= 0).all()
    # the reported crossing (the refinement's previous test point, SCNM:144-162) is bracketed by a sign change of the
    # field within one march step on either side; camera rays start outside the sphere, so sign0 = +1
    o, d, t = rays["o"][hit][:4096], rays["d"][hit][:4096], a["t"][hit][:4096]
    p0 = (o.astype(np.float64) + (t[:, None] - 0.0101) * d).astype(np.float32); p1 = (o.astype(np.float64) + (t[:, None] + 0.0101) * d).astype(np.float32)
    v0 = gpu.eval(np.ascontiguousarray(p0), np.ascontiguousarray(d), gradient=False)["val"]; v1 = gpu.eval(np.ascontiguousarray(p1), np.ascontiguousarray(d), gradient=False)["val"]
    assert ((v0 >= 0) & (v1 < 0)).mean() > 0.97


# Bounds of the large-sample query parity, per (medium, wavefront): (decision, step_moved, t_rel_max, grad_rel_max, ctx_rel_max).
# Integers are (observed on B200 + 1), relative errors the observed maximum rounded up; every number is in
# profiles/r02_parity_rates.txt (tools/parity_rates.py).  north_star's tolerance is 1e-4: entries above it are explained there.
DEFAULT_BOUND = (3, 3, 1e-4, 3e-4, 1e-3, TOL)  # last: 99.9 % quantile of the gradient error
# A ProceduralMean's gradient is a forward difference with eps = 1e-3 of a mean held in FLOAT (math/GPFunctions.cpp:279-296): the
# reference's own gradient carries ulp(mean) / eps ~ 1e-4 ... 3e-3 of rounding noise, re-drawn whenever the hit point moves by an ulp.
# The mean itself is bit-identical (golden media); the bound below is that noise, measured (profiles/r02_parity_rates.txt).
_FD = (3, 3, 1e-4, 5e-3, 1e-3, 5e-4)
PARITY_BOUNDS = {(n, sec): _FD for n in ("proc_mean_two_spheres_iso", "proc_mean_grid_sdf_iso", "proc_mean_knob_inner_1d") for sec in (False, True)}
# The "rust" lateral-scale field sums 11 simplex octaves up to a spatial frequency of ~1e5 per unit (math/GPFunctions.cpp:112-117: p * 2 * 25,
# fbm: * 5 * 0.5 * 2.1042^9): ls(p) itself moves by 1e-4 ... 1e-3 when p moves by a float ulp, and the gradient at a hit scales with it.  The
# field is bit-identical at a given point (golden media); the hit POINT differs by an ulp between the implementations.
_ROUGH = (3, 3, 1e-4, 5e-3, 1e-3, 1e-3)
PARITY_BOUNDS.update({("fbm_rust_ls_multires_iso", sec): _ROUGH for sec in (False, True)})
PARITY_BOUNDS.update({(n, sec): _FD for n in ("proc_mean_knob_iso", "proc_mean_plane_world") for sec in (False, True)})
# Matern nu = 1/2 (a kernel with a cusp at the impulse) and Gabor: ONE primary ray of 8192 whose gradient is nearly zero sits at 2.2e-3 / 4.4e-3 of the rms
# gradient; the 99.9 % quantile is 2.5e-5 / 5e-6 (profiles/r02_parity_rates_late_kernels.txt)
PARITY_BOUNDS.update({(n, sec): (3, 3, 1e-4, 6e-3, 1e-3, TOL) for n in ("matern05_world", "gabor_aniso_world") for sec in (False, True)})


@pytest.mark.parametrize("secondary", [False, True])
@pytest.mark.parametrize("name", S.PARITY_MEDIA)
def test_march_kernels_vs_reference_8k_rays(torch_cuda, name, secondary):
    """The cached march kernels (k_march in its iso-ray and world modes, stationary / brute-force / multi-resolution; k_march1d) against
    the compiled UNMODIFIED reference (oracle/_ref; the C restatement only where _ref is absent — tests/test_oracle.py shows the two are
    bit-identical on these very wavefronts) on 8192 seeded rays per medium, primary (camera) and secondary (conditioned, Renewal state
    carried in).  Reported as RATES (S.query_parity_rates): decision mismatches, hits moved within one march step (SURVEY H-sign: a sign
    flip at |f| ~ 1e-9 changes the number of 0.9-shrinks of the secant refinement), and max relative errors on the matched queries."""
    r = S.query_parity_case(name, secondary)
    if S.have_ref():
        assert r["checker"] == "ref"
    dec, moved, t_max, g_max, c_max, g_p999 = PARITY_BOUNDS.get((name, secondary), DEFAULT_BOUND)
    assert r["decision"] <= dec and r["step_moved"] <= moved, r
    assert r["t_rel_max"] <= t_max, r
    assert r["grad_rel_p999"] <= g_p999 and r["grad_rel_max"] <= g_max, r
    assert max(r["ctx_rel_p999"]) <= TOL and max(r["ctx_rel_max"]) <= c_max, r
    assert r["state_int_mismatch"] == 0 and r["scheme_mismatch"] == 0, r
    assert r["transmittance_mismatch"] <= 2, r
    assert r["hit_fraction"] > 0.05


def test_edge_cases(torch_cuda):
    """Empty batch, zero-length ray (maxT == 0), infinite farT, max_bounces reached, ragged sizes."""
    mj = S.MEDIA["iso_3d_renewal"]
    gpu, orc = S.GpuImpl(mj), S.checker(mj)
    rays, st = S.camera_rays(0)
    out = gpu.sample_distance(rays, st, 1)
    assert len(out["ok"]) == 0
    for n in (1, 31, 33, 129):
        rays, st = S.camera_rays(n, seed=n, secondary=True)
        rays["tmax"][0] = 0.0; rays["tmin"][0] = 0.0  # maxT == 0: trivial exit (GPM:239-248)
        if n > 2:
            rays["tmax"][1] = np.inf; rays["tmin"][1] = 1990.0  # infinite ray: farT = nearT + 2000, 10 unit march
            st["bounce"][2] = 1024  # bounce >= max_bounces: refused (GPM:235-237)
        a, b = gpu.sample_distance(rays, st, 1), orc.sample_distance(rays, st, 1)
        rep = S.compare_queries(a, b)
        assert rep["decision_mismatch"] == 0 and rep["t_rel"] <= TOL and rep["grad_rel"] <= TOL, rep
        assert a["ok"][0] == 1 and a["exited"][0] == 1 and a["t"][0] == 0
        if n > 2:
            assert a["ok"][2] == 0


def test_medium_info_reports_the_kernel_family(torch_cuda):
    """scn_medium_info: a configuration that falls to the generic one-thread-per-ray kernels is visible through the API."""
    from sparse_conv_gpis_tungsten_b200 import SparseConvolutionNoiseMedium
    want = {"c1_sphere_3d_none": "k_march", "c5_multires_nonstat_renewalplus": "k_march", "world_3d_renewalplus": "k_march", "c3_1d_renewal": "k_march1d",
            "multires_1d_xy": "k_march1d", "matern15_world": "k_march", "gabor_iso_world": "k_march", "proc_mean_knob_iso": "k_march", "proc_mean_plane_world": "k_march",
            "proc_mean_grid_sdf_iso": "k_march", "proc_mean_knob_inner_1d": "k_march1d",
            "aniso_mtx_world_renewalplus": "k_march", "aniso_mtx_iso_renewalplus": "k_march", "aniso_mtx_1d_xy_mis": "k_march1d",
            "aniso_rot_world_bruteforce": "k_march", "aniso_rot_iso_multires": "k_march", "aniso_rot_iso_const_none": "k_march", "grid_aniso_var_world": "k_march"}
    media = dict(S.MEDIA, one_d_density6=S.microbench_medium(6, True, one_d=True))  # 1D sampling with N > 4: the register window holds at most 4 impulses per cell
    want["one_d_density6"] = "generic"
    for name, kern in want.items():
        m = SparseConvolutionNoiseMedium.from_json(media[name])
        i = m.info()
        assert i["query_kernel"] == kern, (name, i)
        assert i["sm_count"] > 0 and i["kernel_radius"] > 0
        if kern == "k_march":
            assert i["march_block"] in (64, 128) and i["cache_entries"] > 0 and 0 < i["smem_bytes"] <= 227 * 1024
        m.force_generic(True)
        assert m.info()["query_kernel"] == "generic"
        m.close()
    assert SparseConvolutionNoiseMedium.from_json(S.microbench_medium(3, True)).info()["eval_kernel"] == "k_fast_noise3d"


@pytest.mark.parametrize("name", ["c1_sphere_3d_none", "c5_multires_nonstat_renewalplus", "c3_1d_renewal", "world_3d_renewalplus",
                                  "proc_mean_knob_iso", "proc_mean_plane_world", "proc_mean_grid_sdf_iso", "proc_mean_two_spheres_iso", "proc_mean_knob_outer_world", "proc_mean_knob_inner_1d",
                                  "aniso_mtx_world_renewalplus", "aniso_mtx_1d_xy_mis", "aniso_mtx_iso_renewalplus",
                                  "aniso_rot_world_bruteforce", "aniso_rot_iso_multires", "aniso_rot_iso_const_none", "aniso_rot_world_multires_renewal", "grid_aniso_var_world",
                                  "matern05_world", "matern15_world", "matern25_world_global", "gabor_aniso_world", "gabor_iso_world"])
def test_tuned_kernels_equal_generic_kernels(torch_cuda, name):
    """The tuned query kernels (k_march, k_march1d) against the generic one-thread-per-ray walk of the same library on 4096 secondary
    rays: same decisions, 1e-4 on matched queries (both are held to the reference separately; this pins them to each other)."""
    rays, st = S.camera_rays(4096, seed=41, secondary=True)
    tuned, generic = S.GpuImpl(S.MEDIA[name]), S.GpuImpl(S.MEDIA[name])
    generic.m.force_generic(True)
    a, b = tuned.sample_distance(rays, st, 4), generic.sample_distance(rays, st, 4)
    r = S.query_parity_rates(a, b, tuned.transmittance(rays, st, 4), generic.transmittance(rays, st, 4))
    gtol = G.grad_tol(name)  # procedural means: the gradient is a forward difference of the SDF (eps 1e-3), which amplifies the rounding of the hit point
    assert r["decision"] + r["step_moved"] <= 3 and r["t_rel_max"] <= TOL and r["grad_rel_p999"] <= gtol and r["transmittance_mismatch"] <= 2, r
    tuned.close(); generic.close()
