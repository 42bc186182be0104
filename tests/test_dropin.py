"""The drop-in compiled into the reference (oracle/gpu_medium_shim.cpp): the UNMODIFIED reference PathTracer
(integrators/path_tracer/PathTracer.cpp:68, integrators/TraceBase.cpp:118-122) renders the Appendix-A scenes twice — stock
"sparse_conv_noise" medium vs "sparse_conv_noise_gpu" (Medium::sampleDistance / transmittance served by libscn_b200.so through
scn_sample_distance_batch_host / scn_transmittance_batch_host) — with the SAME scene seed and the same per-tile samplers.  The two
images must agree per pixel up to float rounding except for a counted set of paths whose zero crossing sits on a sign
flip (SURVEY.md H-sign); no statistical bound is involved."""
import ctypes as C
import os

import numpy as np
import pytest

import support as S

needs_ref = pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref/libscn_ref.so not built (needs /root/reference at build time)")


def _has_gpu():
    import torch
    return torch.cuda.is_available()


@needs_ref
def test_reference_exceptions_cross_the_harness():
    """FAIL(...) inside the reference (media/GaussianProcessMedium.cpp:40) must come back as an error string, not a crash: the
    harness library shares the process's C++ runtime (oracle/Makefile STDCXX)."""
    L = C.CDLL(S.REF_LIB)
    L.ref_medium_create.restype = C.c_void_p
    L.ref_last_error.restype = C.c_char_p
    h = L.ref_medium_create(b'{"type":"sparse_conv_noise","correlation_context":"bogus"}')
    assert not h and b"Invalid correlation context" in L.ref_last_error()


@needs_ref
def test_gpu_medium_registers_and_fails_loudly_without_a_device():
    """"sparse_conv_noise_gpu" is a factory string of the reference once the shim is registered; without a CUDA device
    prepareForRender must fail with the library's message (no CPU fallback anywhere on this path)."""
    if _has_gpu():
        pytest.skip("a CUDA device is present; covered by the gpu tests")
    sc = S.appendix_a_scene(S.MEDIA["c1_sphere_3d_none"], 16, 1)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        S.ref_render(sc, gpu_medium=True)
    img, _ = S.ref_render(sc)  # the stock medium still renders in the same process
    assert np.isfinite(img).all() and img.mean() > 0


# name -> (medium, resolution, spp, quad light).  Sizes the reference finishes in seconds on the box's host cores.
DROPIN_SCENES = {
    "c1": ("c1_sphere_3d_none", 32, 8, False),
    "c3": ("c3_1d_renewal", 32, 8, False),
    "c4": ("c4", 32, 16, True),
    "c5": ("c5_multires_nonstat_renewalplus", 32, 4, False),
}
# per-pixel agreement: |gpu - cpu| <= PIX_TOL * max(|cpu|, PIX_FLOOR); pixels outside it hold a path that took the other side
# of a zero crossing (H-sign) — counted, bounded per scene by (observed on B200 + margin), profiles/r02_dropin_pathtracer_parity.txt
PIX_TOL, PIX_FLOOR = 1e-3, 0.05
MAX_FLIPPED_PIXELS = {"c1": 8, "c3": 8, "c4": 24, "c5": 12}


def dropin_pair(name, seed=0xBA5EBA11):
    med, res, spp, quad = DROPIN_SCENES[name]
    mj = S.c4_scene_medium() if med == "c4" else S.MEDIA[med]
    sc = S.appendix_a_scene(mj, res, spp, quad=quad)
    cpu, t_cpu = S.ref_render(sc, seed=seed)
    S.ref_gpu_medium_stats(reset=True)
    gpu, t_gpu = S.ref_render(sc, seed=seed, gpu_medium=True)
    calls = S.ref_gpu_medium_stats()
    return cpu, gpu, t_cpu, t_gpu, calls


def dropin_report(name, cpu, gpu):
    err = np.abs(gpu - cpu).max(axis=2)
    tol = PIX_TOL * np.maximum(np.abs(cpu).max(axis=2), PIX_FLOOR)
    bad = err > tol
    good = ~bad
    rel = err / np.maximum(np.abs(cpu).max(axis=2), PIX_FLOOR)
    return {"scene": name, "pixels": int(bad.size), "flipped_pixels": int(bad.sum()), "max_rel_matched": float(rel[good].max()) if good.any() else 0.0,
            "median_rel_matched": float(np.median(rel[good])) if good.any() else 0.0, "mean_cpu": float(cpu.mean()), "mean_gpu": float(gpu.mean()),
            "bit_identical_pixels": int((err == 0).sum())}


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("name", sorted(DROPIN_SCENES))
def test_reference_pathtracer_with_gpu_medium(name):
    cpu, gpu, t_cpu, t_gpu, calls = dropin_pair(name)
    rep = dropin_report(name, cpu, gpu)
    print(rep, "cpu %.1f s, gpu-medium %.1f s, calls %s" % (t_cpu, t_gpu, (calls,)))
    assert calls[0] > 0 and calls[2] >= 1, "the render did not go through libscn_b200.so"
    if name == "c4":
        assert calls[1] > 0, "no shadow ray reached scn_transmittance_batch_host"
    assert np.isfinite(gpu).all()
    assert rep["flipped_pixels"] <= MAX_FLIPPED_PIXELS[name], rep
    assert abs(rep["mean_gpu"] - rep["mean_cpu"]) <= 0.01 * rep["mean_cpu"], rep
