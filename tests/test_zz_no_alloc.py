"""scn_sample_distance_batch_static (include/scn.h; no counterpart in the reference): the 1D-sampling query over the library's file-scope
scratch array, for callers that may not allocate.  Called through the C ABI and compared with scn_sample_distance_batch on the same
wavefronts — several calls in a row, of different sizes, so that every call after the first runs over the previous call's leftovers.

Collected last on purpose.  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu

MEDIA_1D = ["c3_1d_renewal", "c4_1d_renewalplus_xy_mis"]
# (rays, secondary, seed): a large primary wavefront, then a smaller secondary one (conditioned: k_cond1d writes the coefficients), then a
# larger primary one again — the second and third run over scratch the earlier calls left behind
WAVEFRONTS = [(6000, False, 31), (2500, True, 32), (9000, False, 33), (2500, True, 32)]
TOL = 1e-4  # the parity tests' float tolerance (tests/test_gpu_parity.py); integers and flags must be identical


def _device_copy(torch, d):
    return {k: v.clone() for k, v in d.items()}


def _outputs(torch, M, n, dev):
    tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8}
    out = {k: torch.full((n,), 77, dtype=tdt[dt], device=dev) for k, dt in M.SAMPLE_FIELDS}
    out["n_evals"] = torch.zeros(n, dtype=torch.int32, device=dev)
    return out


def _compare(M, a, b, sa, sb, what):
    for k, dt in M.SAMPLE_FIELDS + (("n_evals", np.uint32),):
        x, y = a[k].cpu().numpy(), b[k].cpu().numpy()
        if dt == np.float32:
            ok = np.isfinite(x) == np.isfinite(y)
            assert ok.all(), "%s: %s finite mask differs on %d rays" % (what, k, int((~ok).sum()))
            f = np.isfinite(x)
            err = np.abs(x[f] - y[f]) / np.maximum(np.abs(y[f]), 1e-3)
            assert err.size == 0 or float(err.max()) <= TOL, "%s: %s rel err %g" % (what, k, float(err.max()))
        else:
            assert np.array_equal(x, y), "%s: %s differs on %d rays" % (what, k, int((x != y).sum()))
    for k in sa:  # the Renewal state and the conditioning coefficients, updated in place by both entry points
        x, y = sa[k].cpu().numpy(), sb[k].cpu().numpy()
        if x.dtype.kind == "f":
            f = np.isfinite(y)
            assert np.array_equal(np.isfinite(x), f), "%s: state %s finite mask" % (what, k)
            err = np.abs(x[f] - y[f]) / np.maximum(np.abs(y[f]), 1e-3)
            assert err.size == 0 or float(err.max()) <= TOL, "%s: state %s rel err %g" % (what, k, float(err.max()))
        else:
            assert np.array_equal(x, y), "%s: state %s differs" % (what, k)


@pytest.mark.parametrize("name", MEDIA_1D)
def test_static_scratch_matches_pool_entry_point(built_library, name):
    import torch
    from sparse_conv_gpis_tungsten_b200 import abi, medium as M
    gpu = S.GpuImpl(S.MEDIA[name])
    try:
        assert gpu.m.info()["query_kernel"] == "k_march1d"
        bitwise = 0
        for call, (n, secondary, seed) in enumerate(WAVEFRONTS):
            rays, st = S.camera_rays(n, seed=seed, secondary=secondary)
            r, s0 = gpu._rays_state(rays, st)
            s_pool, s_static = _device_copy(torch, s0), _device_copy(torch, s0)
            o_pool, o_static = _outputs(torch, M, n, gpu.dev), _outputs(torch, M, n, gpu.dev)
            gpu.m.sample_distance(r, s_pool, scene_seed=11, out=o_pool)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(gpu.dev)[0]
            gpu.m.sample_distance_static(r, s_static, o_static, scene_seed=11)
            torch.cuda.synchronize()
            free_after = torch.cuda.mem_get_info(gpu.dev)[0]
            if call > 0:  # the first call may load the kernels' module; later ones must leave the device's free memory alone
                assert free_after >= free_before, "call %d took %d bytes of device memory" % (call, free_before - free_after)
            _compare(M, o_static, o_pool, s_static, s_pool, "%s call %d (n=%d)" % (name, call, n))
            bitwise += all(torch.equal(o_static[k], o_pool[k]) for k in o_pool)
            assert int(o_pool["ok"].sum()) > 0  # a wavefront with crossings, not an all-miss one
        print("%s: %d of %d calls bit-identical to scn_sample_distance_batch" % (name, bitwise, len(WAVEFRONTS)))
    finally:
        gpu.close()


def test_static_scratch_limits(built_library):
    import torch
    from sparse_conv_gpis_tungsten_b200 import abi, medium as M
    n = abi.STATIC_SCRATCH_MAX_RAYS + 1
    gpu = S.GpuImpl(S.MEDIA["c3_1d_renewal"])
    try:
        rays, st = S.camera_rays(n, seed=5)
        r, s = gpu._rays_state(rays, st)
        with pytest.raises(abi.ScnError) as ei:
            gpu.m.sample_distance_static(r, s, _outputs(torch, M, n, gpu.dev))
        assert ei.value.code == abi.SCN_EINVAL and "SCN_STATIC_SCRATCH_MAX_RAYS" in str(ei.value)
        # exactly the maximum is served
        n = abi.STATIC_SCRATCH_MAX_RAYS
        rays, st = S.camera_rays(n, seed=6)
        r, s = gpu._rays_state(rays, st)
        s2 = _device_copy(torch, s)
        a, b = _outputs(torch, M, n, gpu.dev), _outputs(torch, M, n, gpu.dev)
        gpu.m.sample_distance_static(r, s, a)
        gpu.m.sample_distance(r, s2, out=b)
        torch.cuda.synchronize()
        _compare(M, a, b, s, s2, "max-size wavefront")
    finally:
        gpu.close()
    gpu3d = S.GpuImpl(S.MEDIA["c1_sphere_3d_none"])  # k_march: no scratch memory, nothing for this entry point to replace
    try:
        rays, st = S.camera_rays(64, seed=5)
        r, s = gpu3d._rays_state(rays, st)
        with pytest.raises(abi.ScnError) as ei:
            gpu3d.m.sample_distance_static(r, s, _outputs(torch, M, 64, gpu3d.dev))
        assert ei.value.code == abi.SCN_EUNSUPPORTED
    finally:
        gpu3d.close()
