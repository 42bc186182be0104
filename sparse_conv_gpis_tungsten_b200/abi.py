"""ctypes binding of include/scn.h (the C ABI of libscn_b200.so).

This is the Python twin of the cgo/FFI stub a Tungsten maintainer would write (INTEGRATION.md shows the C++ one).
The library is loaded from the package directory; a missing or unloadable library is an ImportError-level failure
— there is no Python or CPU fallback behind these calls.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SCN_LIB_PATH") or os.path.join(HERE, "libscn_b200.so")  # the override is a developer switch for A/B builds

SCN_OK, SCN_EINVAL, SCN_EJSON, SCN_EUNSUPPORTED, SCN_ECUDA, SCN_ENOMEM = 0, -1, -2, -3, -4, -5
CTX_GLOBAL, CTX_RENEWAL_PLUS, CTX_RENEWAL, CTX_NONE = 0, 1, 2, 3
SCHEME_UNI, SCHEME_NEE, SCHEME_MIS = 0, 1, 2
EVAL_NOISE, EVAL_VALUE, EVAL_GRADIENT = 1, 2, 4

f32p, u32p, i32p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)


FIELD_REGULAR_GRID = 6
INTERP_POINT, INTERP_LINEAR, INTERP_QUADRATIC = 0, 1, 2


class GridDesc(C.Structure):
    _fields_ = [("values", C.POINTER(C.c_double)), ("bounds_min", C.c_double * 3), ("bounds_max", C.c_double * 3), ("res", C.c_uint32),
                ("interp", C.c_int32), ("owned", C.c_int32), ("_pad", C.c_int32)]


class FieldDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("_pad", C.c_int32), ("val", C.c_double * 3), ("min", C.c_double), ("max", C.c_double),
                ("start", C.c_double), ("end", C.c_double), ("min2", C.c_double), ("max2", C.c_double), ("start2", C.c_double), ("end2", C.c_double),
                ("grid", GridDesc)]


class MeanDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("offset", C.c_float), ("center", C.c_double * 3), ("radius", C.c_float), ("scale", C.c_float),
                ("ref", C.c_double * 3), ("dir", C.c_double * 3), ("min", C.c_float), ("sdf_func", C.c_int32), ("inv_transform", C.c_float * 12),
                ("f", FieldDesc)]


MEAN_PROCEDURAL = 3
SDF_FIELD, SDF_KNOB_INNER, SDF_TWO_SPHERES, SDF_PLANE, SDF_KNOB, SDF_KNOB_OUTER = 0, 1, 2, 3, 4, 5


class MediumDesc(C.Structure):
    _fields_ = [("step_size", C.c_float), ("min_step", C.c_uint32), ("seed", C.c_uint32), ("impulse_density", C.c_float),
                ("single_realization", C.c_int32), ("isotropic_3d_sampling", C.c_int32), ("sampling_1d", C.c_int32),
                ("sampling_scheme_1d", C.c_int32), ("correlation_xy", C.c_int32), ("surf_vol_phase_separate", C.c_int32),
                ("surf_vol_phase_amp_thresh", C.c_float),
                ("sigma_a", C.c_float), ("sigma_s", C.c_float), ("density", C.c_float), ("correlation_context", C.c_int32), ("max_bounces", C.c_int32),
                ("gp_id", C.c_int32), ("has_mean_additional", C.c_int32), ("id_additional", C.c_int32), ("_pad0", C.c_int32),
                ("mean", MeanDesc), ("mean_additional", MeanDesc),
                ("cov_type", C.c_int32), ("use_aniso_mtx", C.c_int32), ("sigma", C.c_float), ("length_scale", C.c_float), ("aniso", C.c_float * 3),
                ("kernel_scale", C.c_float), ("aniso_mtx", C.c_float * 9),
                ("multires", C.c_int32), ("ls", FieldDesc), ("var", FieldDesc),
                ("matern_v", C.c_float), ("gabor_a_inv", C.c_float), ("gabor_f_inv", C.c_float), ("gabor_omega", C.c_float * 3),
                ("_pad1", C.c_int32), ("aniso_field", FieldDesc)]


class Rays(C.Structure):
    _fields_ = [(k, f32p) for k in ("ox", "oy", "oz", "dx", "dy", "dz", "tmin", "tmax", "u")] + \
               [(k, u32p) for k in ("px", "py", "spp", "segment")] + [("scene_seed", C.c_uint32)]


class State(C.Structure):
    _fields_ = [("first_scatter", u8p), ("last_val", f32p), ("last_gx", f32p), ("last_gy", f32p), ("last_gz", f32p), ("bounce", i32p),
                ("last_gp_id", i32p), ("info_t", f32p), ("scheme", u8p), ("ctx", f32p * 7)]


class Samples(C.Structure):
    _fields_ = [("ok", u8p), ("exited", u8p), ("t", f32p), ("continued_t", f32p), ("gx", f32p), ("gy", f32p), ("gz", f32p),
                ("gp_id", i32p), ("scheme", u8p), ("weight", f32p), ("n_evals", u32p)]


class EvalIn(C.Structure):
    _fields_ = [(k, f32p) for k in ("px", "py", "pz", "dx", "dy", "dz", "t", "info_t")] + \
               [(k, u32p) for k in ("pix_x", "pix_y", "pix_spp", "pix_seg")] + [("scene_seed", C.c_uint32), ("noise_seed", C.c_uint32), ("ctx", f32p * 7)]


class EvalOut(C.Structure):
    _fields_ = [("val", f32p), ("gx", f32p), ("gy", f32p), ("gz", f32p), ("gp_id", i32p)]


class NeeIn(C.Structure):
    _fields_ = [(k, f32p) for k in ("dx", "dy", "dz", "nx", "ny", "nz", "px", "py", "pz", "t_segment", "info_t")] + \
               [(k, u32p) for k in ("pix_x", "pix_y", "pix_spp", "pix_seg")] + [("scene_seed", C.c_uint32), ("ctx", f32p * 7)]


class MediumInfo(C.Structure):
    _fields_ = [("query_kernel", C.c_int32), ("eval_kernel", C.c_int32), ("march_block", C.c_int32), ("cache_entries", C.c_int32), ("smem_bytes", C.c_uint32),
                ("device", C.c_int32), ("sm_count", C.c_int32), ("kernel_radius", C.c_float)]


STATIC_SCRATCH_MAX_RAYS = 65536  # SCN_STATIC_SCRATCH_MAX_RAYS (scn_sample_distance_batch_static)
KERNEL_NAMES = {0: "generic", 1: "k_march", 2: "k_march1d", 3: "k_fast_noise3d"}

# every symbol include/scn.h declares: name -> (restype, argtypes)
MediumP = C.c_void_p
SYMBOLS = {
    "scn_last_error": (C.c_char_p, []),
    "scn_abi_version": (C.c_int, []),
    "scn_desc_default": (C.c_int, [C.POINTER(MediumDesc)]),
    "scn_desc_from_json": (C.c_int, [C.c_char_p, C.POINTER(MediumDesc)]),
    "scn_desc_from_json_at": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(MediumDesc)]),
    "scn_desc_release": (None, [C.POINTER(MediumDesc)]),
    "scn_medium_create": (C.c_int, [C.POINTER(MediumDesc), C.c_int, C.POINTER(MediumP)]),
    "scn_medium_destroy": (None, [MediumP]),
    "scn_eval_batch": (C.c_int, [MediumP, C.POINTER(EvalIn), C.POINTER(EvalOut), C.c_size_t, C.c_int, C.c_void_p]),
    "scn_eval_batch_host": (C.c_int, [MediumP, C.POINTER(EvalIn), C.POINTER(EvalOut), C.c_size_t, C.c_int]),
    "scn_conditioning_batch": (C.c_int, [MediumP, C.POINTER(EvalIn), f32p, f32p, f32p, f32p, C.POINTER(f32p * 7), C.c_size_t, C.c_void_p]),
    "scn_sample_distance_batch": (C.c_int, [MediumP, C.POINTER(Rays), C.POINTER(State), C.POINTER(Samples), C.c_size_t, C.c_void_p]),
    "scn_sample_distance_batch_host": (C.c_int, [MediumP, C.POINTER(Rays), C.POINTER(State), C.POINTER(Samples), C.c_size_t]),
    "scn_sample_distance_batch_static": (C.c_int, [MediumP, C.POINTER(Rays), C.POINTER(State), C.POINTER(Samples), C.c_size_t, C.c_void_p]),
    "scn_transmittance_batch": (C.c_int, [MediumP, C.POINTER(Rays), C.POINTER(State), f32p, C.c_size_t, C.c_void_p]),
    "scn_transmittance_batch_host": (C.c_int, [MediumP, C.POINTER(Rays), C.POINTER(State), f32p, C.c_size_t]),
    "scn_nee_pdf_batch": (C.c_int, [MediumP, C.POINTER(NeeIn), f32p, C.c_size_t, C.c_void_p]),
    "scn_nee_grad_batch": (C.c_int, [MediumP, C.POINTER(NeeIn), f32p, f32p, f32p, C.c_size_t, C.c_void_p]),
    "scn_medium_info": (C.c_int, [MediumP, C.POINTER(MediumInfo)]),
    "scn_debug_force_generic": (C.c_int, [MediumP, C.c_int]),
    "scn_debug_hash4": (C.c_int, [u32p, u32p, u32p, u32p, u32p, C.c_size_t, C.c_void_p]),
    "scn_debug_pcg_stream": (C.c_int, [u32p, f32p, C.c_size_t, C.c_int, C.c_void_p]),
    "scn_debug_cell_impulses": (C.c_int, [u32p, u32p, u32p, C.c_uint32, C.c_uint32, f32p, C.c_size_t, C.c_void_p]),
    "scn_debug_accepted_counts": (C.c_int, [MediumP, f32p, f32p, f32p, f32p, f32p, f32p, C.c_uint32, u32p, C.c_size_t, C.c_void_p]),
}

_lib = None


class ScnError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("scn error %d: %s" % (code, message))
        self.code = code


def load():
    """dlopen libscn_b200.so and bind every declared symbol.  Raises if the library is absent (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU/Python fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.scn_abi_version() != 5:
        raise ImportError("libscn_b200.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc):
    if rc != SCN_OK:
        raise ScnError(rc, load().scn_last_error().decode("utf-8", "replace"))
