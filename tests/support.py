"""Shared test infrastructure: media configurations, seeded inputs, and one driver class over the three CPU-side
implementations that share the array-of-structs harness signatures:

  ref  oracle/_ref/libscn_ref.so        the UNMODIFIED reference compiled from /root/reference (container only)
  orc  oracle/_build/libscn_oracle.so   the plain-C restatement (the oracle; always buildable)
  sim  tests/hostsim/_build/...         the product's device header compiled for the host (logic check, CPU CI leg)

Only tests/, bench.py's cpu_baseline/reference legs and __graft_entry__.smoke() import this module.
"""
import copy
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "libscn_ref.so")
ORC_LIB = os.path.join(ORACLE_DIR, "_build", "libscn_oracle.so")
SIM_DIR = os.path.join(ROOT, "tests", "hostsim")
SIM_LIB = os.path.join(SIM_DIR, "_build", "libscn_hostsim.so")
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

import sys
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# ------------------------------------------------------------------------------------------------ media
_BASE = {
    "name": "gpis", "type": "sparse_conv_noise", "sigma_a": 0.0, "sigma_s": 1.0, "density": 1.0,
    "correlation_context": "none", "single_realization": True, "step_size": 0.01, "impulse_density": 3,
    "isotropic_3D_sampling": True, "1D_sampling": False, "seed": 1,
    "phase_function": {"type": "brdf", "bsdf": {"type": "lambert", "albedo": 0.8}},
    "gaussian_process": {"type": "standard", "mean": {"type": "spherical", "center": [0, 0, 0], "radius": 1.0},
                         "covariance": {"type": "squared_exponential", "sigma": 0.05, "lengthScale": 0.1, "aniso": [1, 1, 1]}},
}
_NONSTAT = {"type": "proc_nonstationary", "multiResolutionGrid": True,
            "cov": {"type": "squared_exponential", "sigma": 0.05, "lengthScale": 0.1, "aniso": [1, 1, 1]},
            "ls": {"type": "noise", "noise": "bottom_top", "min": 0.5, "max": 3.0, "start": -1.0, "end": 1.0}}


def medium(**kw):
    m = copy.deepcopy(_BASE)
    cov = kw.pop("covariance", None)
    mean = kw.pop("mean", None)
    m.update(kw)
    if cov is not None:
        m["gaussian_process"]["covariance"] = copy.deepcopy(cov)
    if mean is not None:
        m["gaussian_process"]["mean"] = copy.deepcopy(mean)
    return m


def microbench_medium(density=3, iso=True, one_d=False):
    """BASELINE.json configs[1] / SURVEY §8(d) C2: stationary SE sigma=1 l=1 aniso=1 localScale=3, single realization, seed 1."""
    return medium(impulse_density=density, isotropic_3D_sampling=iso, **{"1D_sampling": one_d},
                  covariance={"type": "squared_exponential", "sigma": 1, "lengthScale": 1, "aniso": [1, 1, 1], "localScale": 3})


def _nonstat(multires=True, **kw):
    c = copy.deepcopy(_NONSTAT)
    c["multiResolutionGrid"] = multires
    c.update(kw)
    return c


def _grid(name):
    """A regular_grid field over tests/golden/grids/<name>.json (absolute path: the reference resolves "file" against the scene's directory,
    the harness has none)."""
    return {"type": "regular_grid", "file": os.path.join(GOLDEN_DIR, "grids", name + ".json")}


# a rotated, sheared anisotropy matrix (row-major), well conditioned
_MTX_COV = {"type": "squared_exponential", "sigma": 0.05, "lengthScale": 0.1, "aniso": [1, 1, 1], "useAnisoMtx": True,
            "anisoMtx": [1.2, 0.3, -0.1, -0.2, 0.8, 0.25, 0.15, -0.1, 1.1]}

MEDIA = {
    # BASELINE.json configs
    "c1_sphere_3d_none": medium(),
    "c2_micro_iso": microbench_medium(3, True),
    "c2_micro_world": microbench_medium(3, False),
    "c3_1d_renewal": medium(correlation_context="renewal", single_realization=False, **{"1D_sampling": True, "1D_sampling_scheme": "uni"}),
    "c4_1d_renewalplus_xy_mis": medium(correlation_context="renewal+", single_realization=False,
                                        **{"1D_sampling": True, "1D_sampling_scheme": "mis", "1D_gradient_correlationXY": True}),
    "c5_multires_nonstat_renewalplus": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True)),
    # variants covering the remaining branches
    "world_3d_none": medium(isotropic_3D_sampling=False),
    "world_3d_renewalplus": medium(isotropic_3D_sampling=False, correlation_context="renewal+", single_realization=False),
    "iso_3d_renewal": medium(correlation_context="renewal", single_realization=False),
    "iso_3d_renewalplus_aniso": medium(correlation_context="renewal+", single_realization=False,
                                       covariance={"type": "squared_exponential", "sigma": 0.07, "lengthScale": 0.12, "aniso": [1, 0.5, 2], "localScale": 2.5}),
    "iso_3d_global": medium(correlation_context="global", single_realization=False),
    "1d_renewalplus_noxy": medium(correlation_context="renewal+", single_realization=False, **{"1D_sampling": True, "1D_sampling_scheme": "nee"}),
    "1d_none_single": medium(**{"1D_sampling": True}),
    "linear_mean_world": medium(isotropic_3D_sampling=False, impulse_density=5,
                                mean={"type": "linear", "reference_point": [0, -0.2, 0], "direction": [0.2, 1, 0.1], "scale": 0.8, "min": -0.5}),
    "homogeneous_mean_density8": medium(impulse_density=8.7, mean={"type": "homogeneous", "offset": 0.01}),
    "nonstat_bruteforce_iso": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(False)),
    "nonstat_bruteforce_world_var": medium(isotropic_3D_sampling=False, correlation_context="renewal", single_realization=False,
                                           covariance=_nonstat(False, var={"type": "noise", "noise": "left_right", "min": 0.5, "max": 2.0, "start": -1.5, "end": 1.5})),
    "multires_world": medium(isotropic_3D_sampling=False, correlation_context="renewal+", single_realization=False, covariance=_nonstat(True)),
    "multires_1d_xy": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True),
                             **{"1D_sampling": True, "1D_sampling_scheme": "mis", "1D_gradient_correlationXY": True}),
    "multires_1d_renewal": medium(correlation_context="renewal", single_realization=False, covariance=_nonstat(True, ls={"type": "noise", "noise": "bottom_top_left_right", "min": 0.4, "max": 1.5, "start": -1, "end": 1, "min2": 0.8, "max2": 1.6, "start2": -1.2, "end2": 1.1}),
                                  **{"1D_sampling": True}),
    "csg_two_spheres": medium(**{"gaussian_process": {"type": "csg",
                                 "left": {"type": "standard", "id": 0, "mean": {"type": "spherical", "center": [-0.4, 0, 0], "radius": 0.7},
                                          "covariance": {"type": "squared_exponential", "sigma": 0.05, "lengthScale": 0.1, "aniso": [1, 1, 1]}},
                                 "right": {"type": "standard", "id": 1, "mean": {"type": "spherical", "center": [0.5, 0.1, 0], "radius": 0.6},
                                           "covariance": {"type": "squared_exponential", "sigma": 0.05, "lengthScale": 0.1, "aniso": [1, 1, 1]}}}}),
    "absorption_only": medium(sigma_a=1.0, sigma_s=0.0),
    # alternative splat kernels: world-space 3D sampling, no conditioning
    "matern05_world": medium(isotropic_3D_sampling=False, covariance={"type": "matern", "sigma": 0.05, "v": 0.5, "lengthScale": 0.1, "aniso": [1, 1, 1]}),
    "matern15_world": medium(isotropic_3D_sampling=False, impulse_density=4, covariance={"type": "matern", "sigma": 0.05, "v": 1.5, "lengthScale": 0.12, "aniso": [1, 0.6, 1.5], "localScale": 2.5}),
    "matern25_world_global": medium(isotropic_3D_sampling=False, correlation_context="global", single_realization=False,
                                    covariance={"type": "matern", "sigma": 0.05, "v": 2.5, "lengthScale": 0.1, "aniso": [1, 1, 1]}),
    "gabor_aniso_world": medium(isotropic_3D_sampling=False, covariance={"type": "gabor_aniso", "sigma": 0.05, "a_inv": 0.15, "f_inv": 0.2, "omega": [1, 2, 0.5]}),
    "gabor_iso_world": medium(isotropic_3D_sampling=False, impulse_density=5, covariance={"type": "gabor_iso", "sigma": 0.05, "a_inv": 0.15, "f_inv": 0.25, "localScale": 3.5}),
    # rotation-anisotropic non-stationary kernel (math/GPFunctions.cpp:1678-1690): full-matrix splat kernel, 3D sampling
    "aniso_rot_world_bruteforce": medium(isotropic_3D_sampling=False, correlation_context="renewal+", single_realization=False,
                                         covariance=_nonstat(False, aniso={"type": "noise", "noise": "left_right", "min": 0.1, "max": 0.9, "start": -1.5, "end": 1.5})),
    "aniso_rot_iso_multires": medium(correlation_context="renewal+", single_realization=False,
                                     covariance=_nonstat(True, aniso={"type": "noise", "noise": "bottom_top", "min": 0.2, "max": 1.2, "start": -1.0, "end": 1.0})),
    "aniso_rot_iso_const_none": medium(covariance=_nonstat(False, aniso={"type": "constant", "val": 0.35})),
    "aniso_rot_world_multires_renewal": medium(isotropic_3D_sampling=False, correlation_context="renewal", single_realization=False, impulse_density=4,
                                               covariance=_nonstat(True, aniso={"type": "noise", "noise": "front_back", "min": 0.3, "max": 0.8, "start": -1.0, "end": 1.5})),
    "nonstat_bruteforce_1d_renewalplus": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(False),
                                                **{"1D_sampling": True, "1D_sampling_scheme": "mis"}),  # (correlationXY is not defined for brute force: SCN warns)
    # "useAnisoMtx": full anisotropy matrix of the stationary kernel (math/GPFunctions.cpp:668-675)
    "aniso_mtx_world_renewalplus": medium(isotropic_3D_sampling=False, correlation_context="renewal+", single_realization=False,
                                          covariance=dict(_MTX_COV)),
    "aniso_mtx_iso_renewalplus": medium(correlation_context="renewal+", single_realization=False, covariance=dict(_MTX_COV, localScale=2.5)),
    "aniso_mtx_1d_xy_mis": medium(correlation_context="renewal+", single_realization=False, covariance=dict(_MTX_COV),
                                  **{"1D_sampling": True, "1D_sampling_scheme": "mis", "1D_gradient_correlationXY": True}),
    "aniso_mtx_multires_iso": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True, cov=dict(_MTX_COV))),
    # "regular_grid" procedural fields (math/GPFunctions.hpp:241-330, 429-578): dense res^3 arrays read from disk, point / trilinear /
    # triquadratic lookup in double.  A grid "ls" exists in the multi-resolution form only (the reference's maxVal() of a grid field is 0).
    "grid_var_bruteforce_iso": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(False, var=_grid("var8_linear"))),
    "grid_ls_multires_iso": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True, ls=_grid("ls6_quadratic"))),
    "grid_ls_var_multires_1d": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True, ls=_grid("ls7_linear"), var=_grid("var5_quadratic")),
                                      **{"1D_sampling": True, "1D_sampling_scheme": "mis", "1D_gradient_correlationXY": True}),
    # "procedural" means (math/GPFunctions.hpp:1049-1087, GPFunctions.cpp:250-296): SDF functions of math/SdfFunctions.cpp and scalar fields
    # (a regular_grid "f" is a tabulated signed distance), config transform, finite-difference gradient
    "proc_mean_plane_world": medium(isotropic_3D_sampling=False, mean={"type": "procedural", "func": "plane", "scale": 0.9, "offset": 0.05,
                                                                       "transform": {"position": [0.1, 0.2, -0.1], "rotation": [20, -35, 30], "scale": [1.2, 0.8, 1.0]}}),
    "proc_mean_two_spheres_iso": medium(correlation_context="renewal+", single_realization=False,
                                        mean={"type": "procedural", "func": "two_spheres", "scale": 0.1, "transform": {"scale": [0.1, 0.1, 0.1], "position": [0, 0.05, 0]}}),
    "proc_mean_knob_inner_1d": medium(correlation_context="renewal", single_realization=False, mean={"type": "procedural", "f": {"type": "sdf", "function": "knob_inner"}, "min": -0.3},
                                      **{"1D_sampling": True}),
    "proc_mean_knob_iso": medium(correlation_context="renewal+", single_realization=False, mean={"type": "procedural", "func": "knob", "transform": {"position": [0, 0.1, 0]}}),
    "proc_mean_knob_outer_world": medium(isotropic_3D_sampling=False, mean={"type": "procedural", "func": "knob_outer", "scale": 1.1, "transform": {"rotation": [10, 40, 0]}}),
    "proc_mean_grid_sdf_iso": medium(correlation_context="renewal+", single_realization=False,
                                     mean={"type": "procedural", "f": _grid("torus12_sdf_linear"), "transform": {"rotation": [0, 0, 25]}}),
    # fbm fields "sandstone" / "rust" (math/GPFunctions.cpp:70-83, 104-117 over math/SdfFunctions.cpp:200-296: simplex noise on a sin() hash)
    "fbm_sandstone_var_bruteforce_iso": medium(correlation_context="renewal+", single_realization=False,
                                               covariance=_nonstat(False, var={"type": "noise", "noise": "sandstone", "min": 0.6, "max": 1.8})),
    "fbm_rust_var_world": medium(isotropic_3D_sampling=False, correlation_context="global", single_realization=False,
                                 covariance=_nonstat(False, var={"type": "noise", "noise": "rust", "min": 0.5, "max": 2.0})),
    "fbm_rust_ls_multires_iso": medium(correlation_context="renewal+", single_realization=False, covariance=_nonstat(True, ls={"type": "noise", "noise": "rust"})),
    "fbm_rust_ls_bruteforce_1d": medium(correlation_context="renewal", single_realization=False, covariance=_nonstat(False, ls={"type": "noise", "noise": "rust"}),
                                        **{"1D_sampling": True}),
    "grid_aniso_var_world": medium(isotropic_3D_sampling=False, correlation_context="renewal", single_realization=False,
                                   covariance=_nonstat(False, aniso=_grid("aniso4_point"), var=_grid("var5_quadratic"))),
}
IS_1D = {k: bool(v.get("1D_sampling")) for k, v in MEDIA.items()}


QUAD_LIGHT = {"type": "quad", "transform": {"position": [0, 4, 0], "scale": [2, 2, 2], "rotation": [180, 0, 0]}, "emission": 20.0, "bsdf": {"type": "null"}}


def c4_scene_medium():
    """the C4 medium (1D, Renewal+, correlationXY, mis) with a copper conductor as the phase bsdf."""
    m = copy.deepcopy(MEDIA["c4_1d_renewalplus_xy_mis"])
    m["phase_function"] = {"type": "brdf", "bsdf": {"type": "conductor", "material": "Cu"}}
    return m


def appendix_a_scene(medium_json, res, spp, max_bounces=8, quad=False, env=1.0):
    """SURVEY.md Appendix A.1 around `medium_json`: cube [-1.5,1.5]^3 holding the medium, constant environment, pinhole camera
    (A.3 with quad=True: plus the quad area light).  Box reconstruction filter (the wavefront stage's), no embree
    ("scene_bvh": false), no output files."""
    return {"media": [medium_json], "bsdfs": [],
            "primitives": [{"type": "cube", "transform": {"scale": [3, 3, 3]}, "bsdf": {"type": "forward"}, "int_medium": medium_json["name"]},
                           {"type": "infinite_sphere", "emission": env, "sample": True}] + ([copy.deepcopy(QUAD_LIGHT)] if quad else []),
            "camera": {"type": "pinhole", "resolution": [res, res], "fov": 35, "tonemap": "gamma", "reconstruction_filter": "box",
                       "transform": {"position": [0, 0, 5], "look_at": [0, 0, 0], "up": [0, 1, 0]}},
            "integrator": {"type": "path_tracer", "min_bounces": 0, "max_bounces": max_bounces},
            "renderer": {"spp": spp, "spp_step": spp, "adaptive_sampling": False, "stratified_sampler": False, "scene_bvh": False,
                         "output_file": "", "hdr_output_file": "", "overwrite_output_files": True}}


# ------------------------------------------------------------------------------------------------ seeded inputs
def eval_points(n, seed=1234, lo=-8.0, hi=8.0):
    """points uniform in [-8,8]^3 (seed 1234) and unit directions (seed 1235)."""
    p = np.random.default_rng(seed).uniform(lo, hi, (n, 3)).astype(np.float32)
    d = np.random.default_rng(seed + 1).normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return np.ascontiguousarray(p), np.ascontiguousarray(d)


def camera_rays(n, seed=7, half=1.5, secondary=False):
    """Rays through the 3-unit cube of Appendix A.1: pinhole at (0,0,5) looking at the origin (primary), or rays
    started inside the cube in random directions with a previous value/gradient to condition on (secondary)."""
    rng = np.random.default_rng(seed)
    if not secondary:
        o = np.tile(np.array([0, 0, 5], np.float32), (n, 1))
        tgt = rng.uniform(-1.3, 1.3, (n, 3)); tgt[:, 2] = 0
        d = tgt - o
    else:
        o = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
        d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    # slab intersection with [-half, half]^3
    with np.errstate(divide="ignore", invalid="ignore"):
        t0 = (-half - o) / d; t1 = (half - o) / d
    tn = np.minimum(t0, t1).max(axis=1); tf = np.maximum(t0, t1).min(axis=1)
    tn = np.maximum(tn, 0).astype(np.float32); tf = tf.astype(np.float32)
    miss = tf <= tn
    tf[miss] = tn[miss] + 0.5
    if secondary:
        tn[:] = 0.0
    rays = {"o": np.ascontiguousarray(o.astype(np.float32)), "d": np.ascontiguousarray(d), "tmin": tn, "tmax": tf,
            "u": rng.uniform(0, 1, n).astype(np.float32),
            "pix": np.ascontiguousarray(np.stack([rng.integers(0, 512, n), rng.integers(0, 512, n), rng.integers(0, 64, n),
                                                   rng.integers(0, 6, n) if secondary else np.zeros(n, np.int64)], 1).astype(np.uint32))}
    st = {"first_scatter": np.ones(n, np.uint8), "last_val": np.zeros(n, np.float32), "last_grad": np.zeros((n, 3), np.float64),
          "bounce": np.zeros(n, np.int32), "last_gp_id": np.zeros(n, np.int32), "info_t": np.zeros(n, np.float32)}
    if secondary:
        st["first_scatter"][:] = 0
        st["last_val"] = rng.normal(0, 0.02, n).astype(np.float32)
        g = rng.normal(size=(n, 3)); g /= np.linalg.norm(g, axis=1, keepdims=True)
        st["last_grad"] = np.ascontiguousarray((g * rng.uniform(0.5, 1.5, (n, 1))).astype(np.float32).astype(np.float64))
        st["bounce"] = rng.integers(1, 5, n).astype(np.int32)
        st["info_t"] = rng.uniform(0, 3, n).astype(np.float32)
    return rays, st


# ------------------------------------------------------------------------------------------------ library drivers
def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)
    return ORC_LIB


def build_hostsim():
    src = os.path.join(SIM_DIR, "hostsim.cpp")
    csrc = os.path.join(ROOT, "sparse_conv_gpis_tungsten_b200", "csrc")
    deps = [src, os.path.join(csrc, "scn_device.cuh"), os.path.join(csrc, "scn_params.cpp"), os.path.join(csrc, "scn_internal.h")]
    if os.path.exists(SIM_LIB) and all(os.path.getmtime(SIM_LIB) >= os.path.getmtime(d) for d in deps):
        return SIM_LIB
    os.makedirs(os.path.dirname(SIM_LIB), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O2", "-msse4.2", "-ffp-contract=off", "-fPIC", "-shared", "-o", SIM_LIB, src,
                    os.path.join(csrc, "scn_params.cpp")], check=True)
    return SIM_LIB


def desc_from_json(mj):
    from sparse_conv_gpis_tungsten_b200.medium import SparseConvolutionNoiseMedium
    return SparseConvolutionNoiseMedium.desc_from_json(mj)


class CpuImpl:
    """ref / orc / sim behind one interface (array-of-structs, numpy in, numpy out)."""

    def __init__(self, kind, medium_json, threads=None):
        self.kind = kind
        self.threads = threads or min(8, os.cpu_count() or 1)
        self.mj = medium_json
        js = json.dumps(medium_json).encode()
        if kind == "ref":
            if not os.path.exists(REF_LIB):
                raise FileNotFoundError(REF_LIB)
            self.lib = C.CDLL(REF_LIB)
            self.pre = "ref_"
            self.lib.ref_medium_create.restype = C.c_void_p
            self.lib.ref_last_error.restype = C.c_char_p
            self.h = self.lib.ref_medium_create(js)
            if not self.h:
                raise RuntimeError("reference refused the medium: %s" % self.lib.ref_last_error().decode())
        else:
            self.lib = C.CDLL(build_oracle() if kind == "orc" else build_hostsim())
            self.pre = "orc_" if kind == "orc" else "sim_"
            create = getattr(self.lib, self.pre + "medium_create")
            create.restype = C.c_void_p
            self.desc = desc_from_json(js)
            self.h = create(C.byref(self.desc))
            if not self.h:
                raise RuntimeError("%s refused the medium" % kind)
        self.hp = C.c_void_p(self.h)

    def close(self):
        if self.h:
            getattr(self.lib, self.pre + "medium_destroy")(self.hp)
            self.h = None

    def fn(self, name):
        return getattr(self.lib, self.pre + name)

    def eval_noise(self, p, d, seed):
        n = len(p)
        out = np.zeros((n, 4), np.float32)
        one_d = bool(self.mj.get("1D_sampling"))
        if one_d and self.kind == "ref":
            raise NotImplementedError("the reference harness has no 1D noise-level entry; use eval()")
        name = "eval_noise1d" if (one_d and self.kind == "orc") else "eval_noise3d"
        self.fn(name)(self.hp, C.c_size_t(n), _vp(p), _vp(d), C.c_uint32(seed), _vp(out), self.threads)
        return out

    def eval(self, p, d, t=None, pix=None, scene_seed=0, info_t=None, coeffs=None, value=True, gradient=True):
        n = len(p)
        if pix is None:
            pix = np.zeros((n, 4), np.uint32)
        val = np.zeros(n, np.float32) if value else None
        grad = np.zeros((n, 3), np.float32) if gradient else None
        gpid = np.zeros(n, np.int32) if value else None
        self.fn("eval")(self.hp, C.c_size_t(n), _vp(p), _vp(d), _vp(t), _vp(pix), C.c_uint32(scene_seed), _vp(info_t), _vp(coeffs),
                        _vp(val), _vp(grad), _vp(gpid), self.threads)
        return {"val": val, "grad": grad, "gp_id": gpid}

    def conditioning(self, p, d, target_val, target_grad, pix, scene_seed=0, info_t=None):
        n = len(p)
        out = np.zeros((n, 7), np.float32)
        self.fn("conditioning")(self.hp, C.c_size_t(n), _vp(p), _vp(d), _vp(target_val), _vp(target_grad), _vp(pix), C.c_uint32(scene_seed),
                                _vp(info_t), _vp(out), self.threads)
        return out

    def sample_distance(self, rays, state, scene_seed=0):
        n = len(rays["u"])
        st = {k: np.array(v, copy=True) for k, v in state.items()}
        out = {"ok": np.zeros(n, np.uint8), "exited": np.zeros(n, np.uint8), "t": np.zeros(n, np.float32), "cont_t": np.zeros(n, np.float32),
               "grad": np.zeros((n, 3), np.float64), "gp_id": np.zeros(n, np.int32), "scheme": np.zeros(n, np.uint8),
               "coeffs": np.zeros((n, 7), np.float32), "weight": np.zeros((n, 3), np.float32)}
        self.fn("sample_distance")(self.hp, C.c_size_t(n), _vp(rays["o"]), _vp(rays["d"]), _vp(rays["tmin"]), _vp(rays["tmax"]), _vp(rays["u"]),
                                   _vp(rays["pix"]), C.c_uint32(scene_seed),
                                   _vp(st["first_scatter"]), _vp(st["last_val"]), _vp(st["last_grad"]), _vp(st["bounce"]), _vp(st["last_gp_id"]), _vp(st["info_t"]),
                                   _vp(out["ok"]), _vp(out["exited"]), _vp(out["t"]), _vp(out["cont_t"]), _vp(out["grad"]), _vp(out["gp_id"]), _vp(out["scheme"]),
                                   _vp(out["coeffs"]), _vp(out["weight"]), self.threads)
        out["state"] = st
        return out

    def transmittance(self, rays, state, scene_seed=0):
        n = len(rays["u"])
        T = np.zeros(n, np.float32)
        self.fn("transmittance")(self.hp, C.c_size_t(n), _vp(rays["o"]), _vp(rays["d"]), _vp(rays["tmin"]), _vp(rays["tmax"]), _vp(rays["u"]),
                                 _vp(rays["pix"]), C.c_uint32(scene_seed),
                                 _vp(state["first_scatter"]), _vp(state["last_val"]), _vp(state["last_grad"]), _vp(state["bounce"]), _vp(state["last_gp_id"]),
                                 _vp(state["info_t"]), _vp(T), self.threads)
        return T

    def nee(self, d, normal, p, t_segment, pix, scene_seed=0, info_t=None, coeffs=None):
        n = len(p)
        pdf = np.zeros(n, np.float32); grad = np.zeros((n, 3), np.float32)
        self.fn("nee")(self.hp, C.c_size_t(n), _vp(d), _vp(normal), _vp(p), _vp(t_segment), _vp(pix), C.c_uint32(scene_seed), _vp(info_t), _vp(coeffs),
                       _vp(pdf), _vp(grad), self.threads)
        return pdf, grad

    def eval_count(self):
        if self.kind != "orc":
            return None
        f = self.lib.orc_last_eval_count
        f.restype = C.c_uint64
        return int(f())


def have_ref():
    return os.path.exists(REF_LIB)


# ------------------------------------------------------------------------------------------------ comparisons
def rel_err(a, b, floor=None):
    """|a-b| / max(|b|, floor); floor defaults to the rms of b (a value near a zero of the field has no relative scale)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    if floor is None:
        floor = float(np.sqrt(np.mean(b * b))) if b.size else 1.0
        floor = floor if floor > 0 else 1.0
    return np.abs(a - b) / np.maximum(np.abs(b), floor)


def compare_queries(a, b, tol=1e-4):
    """Parity report for two sampleDistance result sets (SURVEY H-sign): decision mismatches are counted separately
    from the 1e-4 tolerance on matched queries."""
    same = (a["ok"] == b["ok"]) & (a["exited"] == b["exited"])
    same &= np.abs(a["t"].astype(np.float64) - b["t"]) <= 1e-3 * np.maximum(1.0, np.abs(b["t"]))  # same march step
    rep = {"n": len(same), "decision_mismatch": int((~same).sum())}
    m = same & (a["ok"] != 0)
    rep["t_rel"] = float(rel_err(a["t"][m], b["t"][m], 1e-3).max()) if m.any() else 0.0
    g_floor = float(np.sqrt(np.mean(np.asarray(b["grad"][m], np.float64) ** 2))) if m.any() else 1.0
    rep["grad_rel"] = float(rel_err(a["grad"][m], b["grad"][m], max(g_floor, 1e-12)).max()) if m.any() else 0.0
    rep["ok_fraction"] = float((b["ok"] != 0).mean())
    rep["hit_fraction"] = float(((b["ok"] != 0) & (b["exited"] == 0)).mean())
    return rep


# ------------------------------------------------------------------------------------------------ the product on the GPU
class GpuImpl:
    """The CUDA library behind the same interface as CpuImpl (array-of-structs numpy in/out), through the package's
    host mirror and the C ABI.  `host=True` routes through the *_host entry points (numpy buffers)."""

    def __init__(self, medium_json, device=0, host=False):
        import torch
        from sparse_conv_gpis_tungsten_b200 import SparseConvolutionNoiseMedium
        self.torch = torch
        self.m = SparseConvolutionNoiseMedium.from_json(medium_json, device)
        self.dev = torch.device("cuda", device)
        self.host = host
        self.mj = medium_json

    def close(self):
        self.m.close()

    def _f(self, a):  # (n,k) numpy -> list of k contiguous 1-D buffers
        a = np.asarray(a)
        cols = [np.ascontiguousarray(a[:, j]) for j in range(a.shape[1])] if a.ndim == 2 else [np.ascontiguousarray(a)]
        if self.host:
            return cols
        out = []
        for c in cols:
            if c.dtype == np.uint32:
                c = c.view(np.int32)
            if c.dtype == np.float64:
                c = c.astype(np.float32)
            out.append(self.torch.from_numpy(c).to(self.dev))
        return out

    def _np(self, t):
        if isinstance(t, np.ndarray):
            return t
        return t.detach().cpu().numpy()

    def eval_noise(self, p, d, seed):
        outs = self.m.eval_noise(self._f(p), self._f(d), seed)
        return np.stack([self._np(o) for o in outs], 1)

    def eval(self, p, d, t=None, pix=None, scene_seed=0, info_t=None, coeffs=None, value=True, gradient=True):
        r = self.m.evaluate(self._f(p), self._f(d), t=None if t is None else self._f(t)[0], info_t=None if info_t is None else self._f(info_t)[0],
                            pix=None if pix is None else self._f(pix), scene_seed=scene_seed, ctx=None if coeffs is None else self._f(coeffs),
                            value=value, gradient=gradient)
        out = {"val": None, "grad": None, "gp_id": None}
        if value:
            out["val"] = self._np(r["val"]); out["gp_id"] = self._np(r["gp_id"])
        if gradient:
            out["grad"] = np.stack([self._np(r[k]) for k in ("gx", "gy", "gz")], 1)
        return out

    def conditioning(self, p, d, target_val, target_grad, pix, scene_seed=0, info_t=None):
        assert not self.host
        outs = self.m.conditioning(self._f(p), self._f(d), self._f(target_val)[0], self._f(target_grad), info_t=None if info_t is None else self._f(info_t)[0],
                                   pix=self._f(pix), scene_seed=scene_seed)
        return np.stack([self._np(o) for o in outs], 1)

    def _rays_state(self, rays, state):
        from sparse_conv_gpis_tungsten_b200 import medium as M
        r = dict(zip(("ox", "oy", "oz"), self._f(rays["o"]))); r.update(zip(("dx", "dy", "dz"), self._f(rays["d"])))
        for k in ("tmin", "tmax", "u"):
            r[k] = self._f(rays[k])[0]
        r.update(zip(("px", "py", "spp", "segment"), self._f(rays["pix"])))
        s = {"first_scatter": self._f(state["first_scatter"])[0], "last_val": self._f(state["last_val"])[0], "bounce": self._f(state["bounce"])[0],
             "last_gp_id": self._f(state["last_gp_id"])[0], "info_t": self._f(state["info_t"])[0]}
        lg = np.asarray(state["last_grad"], np.float32)
        s.update(zip(("last_gx", "last_gy", "last_gz"), self._f(lg)))
        n = len(rays["u"])
        s["scheme"] = self._f(np.zeros(n, np.uint8))[0]
        for k in M.CTX_NAMES:
            s[k] = self._f(np.zeros(n, np.float32))[0]
        return r, s

    def sample_distance(self, rays, state, scene_seed=0):
        from sparse_conv_gpis_tungsten_b200 import medium as M
        r, s = self._rays_state(rays, state)
        o = self.m.sample_distance(r, s, scene_seed=scene_seed, count_evals=True)
        out = {"ok": self._np(o["ok"]), "exited": self._np(o["exited"]), "t": self._np(o["t"]), "cont_t": self._np(o["continued_t"]),
               "grad": np.stack([self._np(o[k]) for k in ("gx", "gy", "gz")], 1).astype(np.float64), "gp_id": self._np(o["gp_id"]),
               "scheme": self._np(o["scheme"]), "coeffs": np.stack([self._np(s[k]) for k in M.CTX_NAMES], 1),
               "weight": np.repeat(self._np(o["weight"])[:, None], 3, 1), "n_evals": self._np(o["n_evals"]).view(np.uint32)}
        out["state"] = {"first_scatter": self._np(s["first_scatter"]), "last_val": self._np(s["last_val"]),
                        "last_grad": np.stack([self._np(s[k]) for k in ("last_gx", "last_gy", "last_gz")], 1).astype(np.float64),
                        "bounce": self._np(s["bounce"]), "last_gp_id": self._np(s["last_gp_id"]), "info_t": self._np(s["info_t"])}
        return out

    def transmittance(self, rays, state, scene_seed=0):
        r, s = self._rays_state(rays, state)
        return self._np(self.m.transmittance(r, s, scene_seed=scene_seed))

    def nee(self, d, normal, p, t_segment, pix, scene_seed=0, info_t=None, coeffs=None):
        assert not self.host
        kw = dict(info_t=None if info_t is None else self._f(info_t)[0], pix=self._f(pix), scene_seed=scene_seed, ctx=None if coeffs is None else self._f(coeffs))
        pdf = self.m.nee_pdf(self._f(d), self._f(normal), self._f(p), self._f(t_segment)[0], **kw)
        g = self.m.nee_grad(self._f(d), self._f(normal), self._f(p), **kw)
        return self._np(pdf), np.stack([self._np(x) for x in g], 1)


# ------------------------------------------------------------------------------------------------ the reference's own renderer
def ref_render(scene_json, seed=0xBA5EBA11, threads=None, gpu_medium=False, device=0):
    """Render `scene_json` with the UNMODIFIED reference integrator (oracle/ref_harness.cpp: ref_render) -> ([h, w, 3] float32, seconds).
    gpu_medium=True registers GpuSparseConvolutionNoiseMedium (oracle/gpu_medium_shim.cpp) and switches every
    "sparse_conv_noise" medium of the scene to "sparse_conv_noise_gpu": the same integrator, sampler and seed, with
    Medium::sampleDistance / transmittance served by libscn_b200.so."""
    import json
    import tempfile
    from sparse_conv_gpis_tungsten_b200 import abi
    L = C.CDLL(REF_LIB)
    L.ref_last_error.restype = C.c_char_p
    sc = copy.deepcopy(scene_json)
    if gpu_medium:
        L.ref_gpu_medium_last_error.restype = C.c_char_p
        if L.ref_gpu_medium_register(abi.LIB_PATH.encode(), C.c_int(device)) != 0:
            raise RuntimeError(L.ref_gpu_medium_last_error().decode())
        for m in sc["media"]:
            if m.get("type") == "sparse_conv_noise":
                m["type"] = "sparse_conv_noise_gpu"
    res = sc["camera"]["resolution"]
    w0, h0 = (res, res) if np.isscalar(res) else res
    rgb = np.zeros((h0, w0, 3), np.float32)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "scene.json")
        with open(path, "w") as f:
            json.dump(sc, f)
        w, h, sec = C.c_int(), C.c_int(), C.c_double()
        rc = L.ref_render(path.encode(), int(threads or os.cpu_count() or 1), C.c_uint32(seed), C.byref(w), C.byref(h), rgb.ctypes.data_as(C.c_void_p),
                          C.c_size_t(rgb.size), C.byref(sec))
    if rc != 0:
        raise RuntimeError(L.ref_last_error().decode())
    return rgb, sec.value


def ref_gpu_medium_stats(reset=False):
    """(sampleDistance calls, transmittance calls, media prepared) served by libscn_b200.so through the shim."""
    L = C.CDLL(REF_LIB)
    out = (C.c_ulonglong * 3)()
    L.ref_gpu_medium_stats(out, C.c_int(int(reset)))
    return tuple(int(v) for v in out)


# ------------------------------------------------------------------------------------------------ large-sample query parity
PARITY_MEDIA = ["c1_sphere_3d_none", "c3_1d_renewal", "c4_1d_renewalplus_xy_mis", "c5_multires_nonstat_renewalplus", "iso_3d_renewalplus_aniso",
                "world_3d_renewalplus", "multires_world", "nonstat_bruteforce_iso", "nonstat_bruteforce_world_var", "nonstat_bruteforce_1d_renewalplus",
                "multires_1d_xy", "multires_1d_renewal", "grid_var_bruteforce_iso", "grid_ls_multires_iso", "grid_ls_var_multires_1d",
                "proc_mean_two_spheres_iso", "proc_mean_grid_sdf_iso", "proc_mean_knob_inner_1d", "aniso_mtx_iso_renewalplus", "aniso_mtx_multires_iso",
                "fbm_sandstone_var_bruteforce_iso", "fbm_rust_ls_multires_iso",
                # the kernel variants added late in round 2 (full-matrix kernels, Matern / Gabor, procedural means on the cached marches): profiles/r02_parity_rates_late_kernels.txt
                "aniso_mtx_world_renewalplus", "aniso_mtx_1d_xy_mis", "aniso_rot_world_bruteforce", "aniso_rot_iso_multires", "aniso_rot_iso_const_none",
                "aniso_rot_world_multires_renewal", "matern05_world", "matern15_world", "gabor_aniso_world", "proc_mean_knob_iso", "proc_mean_plane_world"]


def checker(medium_json):
    """The CPU side of a parity test: the compiled UNMODIFIED reference (oracle/_ref) whenever it is present — it travels to the
    GPU box — and the C restatement (pinned bit for bit against it, tests/test_oracle.py) only where it is not."""
    return CpuImpl("ref" if have_ref() else "orc", medium_json)


def _q(x, q):
    return float(np.quantile(x, q)) if len(x) else 0.0


def query_parity_rates(a, b, Ta=None, Tb=None, step=0.01):
    """Mismatch RATES of two sampleDistance result sets over the same rays (a = under test, b = checker), SURVEY H-sign taxonomy:
      decision     ok / exited differ, or the hit lies more than one march step away
      step_moved   same decision, hit distance off by more than 1e-4 relative (floor 1e-3) but within one march step: a sign that
                   flipped at one test point of the secant refinement (|f| ~ 1e-9) changes the number of 0.9-shrinks
      matched      everything else: max / 99.9 % relative error of t, of the gradient (floor: rms of the checker's gradients) and of
                   each of the 7 conditioning coefficients (floor: rms of that coefficient)"""
    n = len(a["ok"])
    dt = np.abs(a["t"].astype(np.float64) - b["t"].astype(np.float64))
    same = (a["ok"] == b["ok"]) & (a["exited"] == b["exited"])
    far = dt > np.maximum(step * 1.01, 1e-3 * np.maximum(1.0, np.abs(b["t"])))
    decision = ~same | (same & (a["ok"] != 0) & far)
    live = same & ~far & (a["ok"] != 0)
    t_rel = dt / np.maximum(np.abs(b["t"]), 1e-3)
    moved = live & (t_rel > 1e-4)
    m = live & ~moved
    rep = {"n": n, "decision": int(decision.sum()), "step_moved": int(moved.sum()), "matched": int(m.sum()),
           "ok_fraction": float((b["ok"] != 0).mean()), "hit_fraction": float(((b["ok"] != 0) & (b["exited"] == 0)).mean()),
           "t_rel_max": float(t_rel[m].max()) if m.any() else 0.0, "t_rel_p999": _q(t_rel[m], 0.999),
           "t_abs_max_moved": float(dt[moved].max()) if moved.any() else 0.0}
    g_floor = max(float(np.sqrt(np.mean(np.asarray(b["grad"][m], np.float64) ** 2))) if m.any() else 1.0, 1e-12)
    ge = rel_err(a["grad"][m], b["grad"][m], g_floor).max(axis=1) if m.any() else np.zeros(0)
    rep["grad_rel_max"] = float(ge.max()) if len(ge) else 0.0
    rep["grad_rel_p999"] = _q(ge, 0.999)
    cmax, cp = [], []
    for k in range(7):
        ca, cb = a["coeffs"][m, k].astype(np.float64), b["coeffs"][m, k].astype(np.float64)
        fl = max(float(np.sqrt(np.mean(cb * cb))) if len(cb) else 0.0, 1e-6)
        e = np.abs(ca - cb) / np.maximum(np.abs(cb), fl)
        cmax.append(float(e.max()) if len(e) else 0.0); cp.append(_q(e, 0.999))
    rep["ctx_rel_max"] = cmax
    rep["ctx_rel_p999"] = cp
    st_a, st_b = a["state"], b["state"]
    rep["state_int_mismatch"] = int(((st_a["bounce"] != st_b["bounce"]) | (st_a["last_gp_id"] != st_b["last_gp_id"]) |
                                     (st_a["first_scatter"] != st_b["first_scatter"]))[~decision].sum())
    rep["scheme_mismatch"] = int((a["scheme"] != b["scheme"])[m].sum())
    if Ta is not None:
        rep["transmittance_mismatch"] = int((Ta != Tb).sum())
    return rep


def query_parity_case(name, secondary, n=8192, seed=23, scene_seed=9, impl=None):
    """One (medium, wavefront) case of the large-sample parity: the library (or `impl`) against the checker."""
    mj = MEDIA[name]
    rays, st = camera_rays(n, seed=seed, secondary=secondary)
    a_impl, b_impl = impl or GpuImpl(mj), checker(mj)
    a, b = a_impl.sample_distance(rays, st, scene_seed), b_impl.sample_distance(rays, st, scene_seed)
    rep = query_parity_rates(a, b, a_impl.transmittance(rays, st, scene_seed), b_impl.transmittance(rays, st, scene_seed),
                             step=float(mj.get("step_size", 0.01)))
    rep.update(medium=name, wavefront="secondary" if secondary else "primary", checker=b_impl.kind)
    if impl is None:
        a_impl.close()
    b_impl.close()
    return rep
