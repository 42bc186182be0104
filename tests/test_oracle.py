"""CPU leg: pins the oracle (oracle/scn_oracle.c) — against the golden vectors recorded from the unmodified reference
(always) and against the compiled reference itself when oracle/_ref is present (build container)."""
import ctypes as C
import os

import numpy as np
import pytest

import golden_util as G
import support as S


@pytest.fixture(scope="module")
def orc():
    return C.CDLL(S.build_oracle())


def test_survey_known_answers(orc):
    """values extracted from the reference headers with g++ 13."""
    orc.orc_xxhash32_1.restype = orc.orc_xxhash32_2.restype = orc.orc_xxhash32_4.restype = C.c_uint32
    assert orc.orc_xxhash32_1(C.c_uint32(0)) == 878055299
    assert orc.orc_xxhash32_2(C.c_uint32(5), C.c_uint32(9)) == 4223614050
    assert orc.orc_xxhash32_4(C.c_uint32(3), C.c_uint32(2), C.c_uint32(1), C.c_uint32(7)) == 2393193692
    out = np.zeros(4, np.float32)
    orc.orc_pcg_stream(C.c_uint64(2393193692 + 1), 4, out.ctypes.data_as(C.c_void_p))
    assert np.allclose(out, [0.996994734, 0.820410132, 0.886805892, 0.165692687], rtol=0, atol=1e-9)
    orc.orc_pcg_next3d(C.c_uint64(2393193692 + 1), out.ctypes.data_as(C.c_void_p))
    # next3D is filled z, y, x (g++ evaluates the constructor arguments right to left)
    assert np.allclose(out, [0.886805892, 0.820410132, 0.996994734, 0.165692687], rtol=0, atol=1e-9)
    imp = S.CpuImpl("orc", S.microbench_medium(3, True))
    p = np.array([[0.3, -0.7, 1.1], [-2.25, 0.5, 0.125], [1, 1, 1]], np.float32); d = np.tile(np.array([0.36, 0.48, 0.8], np.float32), (3, 1))
    want = np.array([[0.697328031, 0.99503243, 0.615222812, -2.10603499], [-0.273985505, -0.296739697, -0.257851452, 0.426979423],
                     [0.399594277, 2.31846929, -0.802524269, -2.88962364]], np.float32)
    assert np.array_equal(imp.eval_noise(p, d, 1), want)
    imp = S.CpuImpl("orc", S.microbench_medium(64, False))
    assert np.array_equal(imp.eval_noise(p[2:], d[2:], 1), np.array([[0.182936788, 0.732263088, 0.557286441, 0.17786321]], np.float32))


def test_primitives_golden(orc):
    g = dict(np.load(os.path.join(S.GOLDEN_DIR, "primitives.npz")))
    orc.orc_xxhash32_1.restype = orc.orc_xxhash32_2.restype = orc.orc_xxhash32_3.restype = orc.orc_xxhash32_4.restype = C.c_uint32
    u = lambda x: C.c_uint32(int(x))
    for k, h1, h2, h3, h4 in zip(g["keys"], g["h1"], g["h2"], g["h3"], g["h4"]):
        assert orc.orc_xxhash32_1(u(k[0])) == h1 and orc.orc_xxhash32_2(u(k[0]), u(k[1])) == h2
        assert orc.orc_xxhash32_3(u(k[0]), u(k[1]), u(k[2])) == h3 and orc.orc_xxhash32_4(u(k[0]), u(k[1]), u(k[2]), u(k[3])) == h4
    for s, st, n3, bm in zip(g["seeds"], g["streams"], g["next3d"], g["box_muller"]):
        a = np.zeros(16, np.float32); orc.orc_pcg_stream(C.c_uint64(int(s)), 16, a.ctypes.data_as(C.c_void_p)); assert np.array_equal(a, st)
        b = np.zeros(4, np.float32); orc.orc_pcg_next3d(C.c_uint64(int(s)), b.ctypes.data_as(C.c_void_p)); assert np.array_equal(b, n3)
        c = np.zeros(8, np.float64); orc.orc_box_muller(C.c_uint64(int(s)), 4, c.ctypes.data_as(C.c_void_p)); assert np.array_equal(c, bm)
    for i, k in enumerate(range(-20, 21)):
        sc = C.c_float(); ad = C.c_int(); orc.orc_multires_level(k, C.byref(sc), C.byref(ad))
        assert (sc.value, ad.value) == (np.float32(g["levels"][i, 0]), int(g["levels"][i, 1]))
        if -6 <= k <= 6:
            assert ad.value == k


@pytest.mark.parametrize("name", sorted(S.MEDIA))
def test_oracle_bit_exact_vs_golden(name):
    """The restatement reproduces the reference bit for bit on every entry point (same compiler family, no FMA)."""
    impl = S.CpuImpl("orc", S.MEDIA[name], threads=2)
    G.check_against_golden(impl, name, G.load(name), exact=True)
    impl.close()


@pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref is only built where /root/reference is mounted")
@pytest.mark.parametrize("name", ["c1_sphere_3d_none", "c4_1d_renewalplus_xy_mis", "c5_multires_nonstat_renewalplus", "nonstat_bruteforce_world_var"])
def test_oracle_vs_live_reference(name):
    """Fresh seeds (not the golden ones) against the reference compiled in this container."""
    mj = S.MEDIA[name]
    ref, orc_ = S.CpuImpl("ref", mj, threads=4), S.CpuImpl("orc", mj, threads=4)
    p, d = S.eval_points(300, seed=991, lo=-1.4, hi=1.4)
    pix = np.random.default_rng(17).integers(0, 999, (300, 4)).astype(np.uint32)
    a, b = orc_.eval(p, d, None, pix, 5), ref.eval(p, d, None, pix, 5)
    assert np.array_equal(a["val"], b["val"]) and np.array_equal(a["grad"], b["grad"])
    for sec in (False, True):
        rays, st = S.camera_rays(120, seed=23 + sec, secondary=sec)
        qa, qb = orc_.sample_distance(rays, st, 3), ref.sample_distance(rays, st, 3)
        for k in ("ok", "exited", "t", "coeffs"):
            assert np.array_equal(qa[k], qb[k]), k
        assert np.array_equal(orc_.transmittance(rays, st, 3), ref.transmittance(rays, st, 3))


@pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref is only built where /root/reference is mounted")
@pytest.mark.parametrize("name", S.PARITY_MEDIA)
def test_oracle_equals_reference_on_the_8k_parity_wavefronts(name):
    """The restatement equals the compiled reference BIT FOR BIT on the very 8192-ray wavefronts the GPU parity test uses
    (tests/test_gpu_parity.py::test_march_kernels_vs_reference_8k_rays): every decision, hit distance, gradient, the seven conditioning
    coefficients, the advanced state and the transmittance.  So a box without oracle/_ref that falls back to the restatement checks
    against the same numbers."""
    for secondary in (False, True):
        r = S.query_parity_case(name, secondary, impl=S.CpuImpl("orc", S.MEDIA[name]))
        assert r["checker"] == "ref"
        assert r["decision"] == 0 and r["step_moved"] == 0 and r["transmittance_mismatch"] == 0 and r["state_int_mismatch"] == 0 and r["scheme_mismatch"] == 0, r
        assert r["t_rel_max"] == 0 and r["grad_rel_max"] == 0 and max(r["ctx_rel_max"]) == 0, r
        assert r["matched"] > 0.2 * r["n"], r


def test_conditioning_residual_invariant():
    """The reference's own run-time self-check (SCN.cpp:469-474,511-515): after conditioning, the realization
    reproduces the target value (and gradient for Renewal+) at the ray origin to < 1e-2."""
    for name in ("iso_3d_renewalplus_aniso", "world_3d_renewalplus", "c5_multires_nonstat_renewalplus"):
        impl = S.CpuImpl("orc", S.MEDIA[name], threads=2)
        p, d = S.eval_points(64, seed=5, lo=-1.2, hi=1.2)
        rng = np.random.default_rng(1)
        pix = rng.integers(0, 99, (64, 4)).astype(np.uint32)
        tv = rng.normal(0, 0.02, 64).astype(np.float32); tg = rng.normal(0, 1, (64, 3)).astype(np.float32)
        c = impl.conditioning(p, d, tv, tg, pix, 9)
        e = impl.eval(p, d, None, pix, 9, None, c)
        assert np.abs(e["val"] - tv).max() < 1e-2
        assert np.linalg.norm(e["grad"] - tg, axis=1).max() < 1e-2
        impl.close()
