"""CPU leg: the C-ABI library loads, exports every symbol include/scn.h declares, parses the reference's JSON like
the reference does, and refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes as C
import json
import os
import re

import pytest

import support as S
from sparse_conv_gpis_tungsten_b200 import abi
from sparse_conv_gpis_tungsten_b200.medium import MediumFactory, SparseConvolutionNoiseMedium as Medium


def test_header_symbols_exported(built_library):
    hdr = open(os.path.join(S.ROOT, "include", "scn.h")).read()
    declared = set(re.findall(r"\b(scn_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(abi.SYMBOLS), declared ^ set(abi.SYMBOLS)
    for name in declared:
        assert hasattr(built_library, name)
    assert built_library.scn_abi_version() == 5


def test_desc_layout_and_defaults(built_library):
    d = abi.MediumDesc()
    abi.check(built_library.scn_desc_default(C.byref(d)))
    # first and last members round-trip => ctypes layout == C layout (media/SparseConvolutionNoiseMedium.cpp:17-35 defaults)
    assert abs(d.step_size - 0.01) < 1e-9 and d.min_step == 8 and d.impulse_density == 3.0 and d.max_bounces == 1024
    assert d.kernel_scale == 3.0 and list(d.aniso) == [1.0, 1.0, 1.0] and d.sigma == 1.0
    assert d.var.end2 == 1.0 and d.ls.max == 500.0 and d.mean.radius == 1.0


def test_reference_json_accepted_unchanged():
    for name, mj in S.MEDIA.items():
        d = Medium.desc_from_json(json.dumps(mj))
        assert d.impulse_density == pytest.approx(mj["impulse_density"])
        assert bool(d.sampling_1d) == bool(mj.get("1D_sampling", False))
    d = Medium.desc_from_json(S.MEDIA["c5_multires_nonstat_renewalplus"])
    assert d.cov_type == 1 and d.multires == 1 and d.ls.type == 2 and d.ls.min == 0.5 and d.ls.max == 3.0
    assert d.correlation_context == abi.CTX_RENEWAL_PLUS
    d = Medium.desc_from_json(S.MEDIA["csg_two_spheres"])
    assert d.has_mean_additional == 1 and d.id_additional == 1 and d.mean_additional.center[0] == 0.5
    d = Medium.desc_from_json(S.MEDIA["linear_mean_world"])
    n = sum(x * x for x in d.mean.dir) ** 0.5
    assert abs(n - 1.0) < 1e-12  # direction normalised like LinearMean::fromJson


@pytest.mark.parametrize("patch,code,msg", [
    ({"correlation_context": "goldfish"}, abi.SCN_EJSON, "Invalid correlation context: 'goldfish'"),
    ({"1D_sampling_scheme": "foo"}, abi.SCN_EJSON, "Invalid sparse conv sampling scheme: 'foo'"),
    ({"intersect_method": "bar"}, abi.SCN_EJSON, "Invalid intersect method"),
    ({"type": "homogeneous"}, abi.SCN_EJSON, "not \"sparse_conv_noise\""),
])
def test_json_errors_match_reference_fail(patch, code, msg):
    mj = dict(S.MEDIA["c1_sphere_3d_none"]); mj.update(patch)
    with pytest.raises(abi.ScnError) as e:
        Medium.desc_from_json(mj)
    assert e.value.code == code and msg in str(e.value)


def test_missing_correlation_context_is_an_error():
    mj = dict(S.MEDIA["c1_sphere_3d_none"]); del mj["correlation_context"]
    with pytest.raises(abi.ScnError) as e:  # the reference's default string "goldfish" is invalid (GaussianProcessMedium.cpp:104-106)
        Medium.desc_from_json(mj)
    assert e.value.code == abi.SCN_EJSON


def test_unsupported_configurations_are_refused_not_approximated():
    mj = S.medium(covariance={"type": "rational_quadratic", "sigma": 1, "lengthScale": 1, "a": 1.5})
    with pytest.raises(abi.ScnError) as e:
        Medium.desc_from_json(mj)
    assert e.value.code == abi.SCN_EUNSUPPORTED
    # matern / gabor: the reference defines them for world-space 3D sampling without conditioning only (GPF.cpp:870-880, GPF.hpp:1336-1345)
    for patch in ({"isotropic_3D_sampling": True}, {"1D_sampling": True}, {"correlation_context": "renewal+", "single_realization": False}):
        mj = S.medium(isotropic_3D_sampling=False, covariance={"type": "matern", "sigma": 1, "lengthScale": 1, "v": 1.5}); mj.update(patch)
        with pytest.raises(abi.ScnError) as e:
            Medium.from_json(mj)
        assert e.value.code == abi.SCN_EUNSUPPORTED, patch
    with pytest.raises(abi.ScnError) as e:
        Medium.from_json(S.medium(isotropic_3D_sampling=False, covariance={"type": "matern", "v": 2.0}))
    assert e.value.code == abi.SCN_EINVAL
    # "useAnisoMtx" parses (nine numbers, row-major); a malformed matrix is a JSON error, the anisotropy field with 1D sampling is refused
    d = Medium.desc_from_json(S.medium(covariance={"type": "squared_exponential", "useAnisoMtx": True, "anisoMtx": [1, 0.5, 0, 0, 2, 0, 0, 0, 1]}))
    assert d.use_aniso_mtx == 1 and list(d.aniso_mtx)[:5] == [1.0, 0.5, 0.0, 0.0, 2.0]
    with pytest.raises(abi.ScnError) as e:
        Medium.desc_from_json(S.medium(covariance={"type": "squared_exponential", "useAnisoMtx": True, "anisoMtx": [1, 0, 0, 1]}))
    assert e.value.code == abi.SCN_EJSON
    d = Medium.desc_from_json(S.medium(covariance=S._nonstat(False, aniso={"type": "constant", "val": 0.3}), **{"1D_sampling": True}))
    assert d.aniso_field.type == 1
    with pytest.raises(abi.ScnError):
        Medium.desc_from_json("{ not json")
    with pytest.raises(abi.ScnError):
        MediumFactory.create({"type": "voxel"})


def test_regular_grid_fields_load_like_the_reference():
    """"regular_grid" fields (math/GPFunctions.hpp:346-401): "file" resolves against the base directory, "path" inside the grid JSON
    against that file's directory; values arrive as res^3 (x3 for the vector field) doubles owned by the descriptor until released."""
    import ctypes as C
    import os
    import numpy as np
    gdir = os.path.join(S.GOLDEN_DIR, "grids")
    mj = S.medium(covariance=S._nonstat(True, ls={"type": "regular_grid", "file": "ls6_quadratic.json"}, var={"type": "regular_grid", "file": "var8_linear.json"}))
    d = Medium.desc_from_json(mj, base_dir=gdir)
    try:
        assert d.ls.type == abi.FIELD_REGULAR_GRID and d.ls.grid.res == 6 and d.ls.grid.interp == abi.INTERP_QUADRATIC and d.ls.grid.owned == 1
        assert d.var.grid.res == 8 and d.var.grid.interp == abi.INTERP_LINEAR and list(d.var.grid.bounds_min) == [-1.6] * 3
        got = np.ctypeslib.as_array(d.ls.grid.values, shape=(6 * 6 * 6 * 3,))
        assert np.array_equal(got, np.fromfile(os.path.join(gdir, "ls6_quadratic-values.bin"), "<f8"))
    finally:
        abi.load().scn_desc_release(C.byref(d))
    assert d.ls.grid.owned == 0 and not d.ls.grid.values
    with pytest.raises(abi.ScnError) as e:  # relative path without its directory
        Medium.desc_from_json(mj, base_dir="/nonexistent")
    assert e.value.code == abi.SCN_EJSON and "cannot read" in str(e.value)
    # a grid "ls" in the brute-force form: the reference's maximum lateral scale of a grid field is 0 (GPFunctions.hpp:474-477) -> refused
    bad = S.medium(covariance=S._nonstat(False, ls={"type": "regular_grid", "file": "ls6_quadratic.json"}))
    db = Medium.desc_from_json(bad, base_dir=gdir)
    try:
        with pytest.raises(abi.ScnError) as e:
            Medium(db)
        assert e.value.code == abi.SCN_EUNSUPPORTED and "multiResolutionGrid" in str(e.value)
    finally:
        abi.load().scn_desc_release(C.byref(db))


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to create a medium (SCN_ECUDA); nothing routes to the oracle."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present: the refusal path cannot be observed")
    with pytest.raises(abi.ScnError) as e:
        Medium.from_json(S.MEDIA["c1_sphere_3d_none"])
    assert e.value.code == abi.SCN_ECUDA and "no CPU path" in str(e.value)
    src = open(os.path.join(S.ROOT, "sparse_conv_gpis_tungsten_b200", "medium.py")).read() + open(os.path.join(S.ROOT, "sparse_conv_gpis_tungsten_b200", "abi.py")).read()
    assert "oracle" not in src.replace("no fallback", "")


def test_null_and_bad_arguments(built_library):
    assert built_library.scn_desc_from_json(None, None) == abi.SCN_EINVAL
    d = abi.MediumDesc(); built_library.scn_desc_default(C.byref(d))
    h = abi.MediumP()
    assert built_library.scn_medium_create(C.byref(d), 0, C.byref(h)) == abi.SCN_EINVAL  # correlation_context unset
    assert b"mandatory" in built_library.scn_last_error()
