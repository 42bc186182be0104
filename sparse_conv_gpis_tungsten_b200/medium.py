"""Host-side mirror of the reference's Medium plugin interface for the sparse-convolution GPIS medium.

Reference (paths relative to its src/core/): media/Medium.hpp:104-111 (sampleDistance / transmittance / pdf),
media/SparseConvolutionNoiseMedium.cpp:57-73 (fromJson), math/SparseConvolutionNoise.hpp:23-89 (realization API),
MediumFactory string "sparse_conv_noise" (media/MediumFactory.cpp:13-22).

The reference answers one query per virtual call; this mirror answers one *wavefront* per call (SoA batches) and
forwards straight to the C ABI of libscn_b200.so (include/scn.h).  Arguments may be CUDA torch tensors (device entry
points, no copies) or numpy arrays / CPU torch tensors (the *_host entry points, which stage through the library's
arena).  Nothing here computes: there is no fallback if the CUDA library is missing.
"""
import ctypes as C
import json
import os

import numpy as np

from . import abi

try:  # torch is plumbing only (device memory, streams)
    import torch
except Exception:  # pragma: no cover
    torch = None

_CT = {np.dtype(np.float32): abi.f32p, np.dtype(np.uint32): abi.u32p, np.dtype(np.int32): abi.i32p, np.dtype(np.uint8): abi.u8p}
_TORCH_TO_NP = {}
if torch is not None:
    _TORCH_TO_NP = {torch.float32: np.dtype(np.float32), torch.int32: np.dtype(np.int32), torch.uint8: np.dtype(np.uint8), torch.bool: np.dtype(np.uint8)}
    if hasattr(torch, "uint32"):
        _TORCH_TO_NP[torch.uint32] = np.dtype(np.uint32)

MEDIUM_TYPE = "sparse_conv_noise"


def _is_torch(x):
    return torch is not None and isinstance(x, torch.Tensor)


class _Args:
    """Resolves arrays to typed pointers and checks that a call does not mix host and device memory."""

    def __init__(self, n=None):
        self.n = n
        self.on_device = None
        self.keep = []

    def ptr(self, x, dtype, name, optional=False):
        ct = _CT[np.dtype(dtype)]
        if x is None:
            if optional:
                return ct()
            raise ValueError("%s is required" % name)
        if _is_torch(x):
            want = np.dtype(dtype)
            have = _TORCH_TO_NP.get(x.dtype)
            if want == np.dtype(np.uint32) and x.dtype == torch.int32:
                have = want  # int32 storage reinterpreted as uint32
            if have != want:
                raise TypeError("%s: expected %s, got %s" % (name, want, x.dtype))
            if not x.is_contiguous() or x.dim() != 1:
                raise ValueError("%s must be a contiguous 1-D tensor" % name)
            dev, count, addr = x.is_cuda, x.numel(), x.data_ptr()
        else:
            a = x
            if not isinstance(a, np.ndarray) or a.dtype != np.dtype(dtype) or a.ndim != 1 or not a.flags["C_CONTIGUOUS"]:
                raise TypeError("%s must be a contiguous 1-D numpy array of %s" % (name, np.dtype(dtype)))
            dev, count, addr = False, a.size, a.ctypes.data
        if self.n is None:
            self.n = count
        elif count != self.n:
            raise ValueError("%s has %d elements, expected %d" % (name, count, self.n))
        if self.on_device is None:
            self.on_device = dev
        elif self.on_device != dev:
            raise ValueError("%s: host and device buffers mixed in one call" % name)
        self.keep.append(x)
        return C.cast(C.c_void_p(addr), ct)


def _stream_ptr(stream):
    if stream is None:
        if torch is not None and torch.cuda.is_available():
            return C.c_void_p(torch.cuda.current_stream().cuda_stream)
        return C.c_void_p(0)
    return C.c_void_p(stream.cuda_stream if hasattr(stream, "cuda_stream") else int(stream))


RAY_F = ("ox", "oy", "oz", "dx", "dy", "dz", "tmin", "tmax", "u")
RAY_U = ("px", "py", "spp", "segment")
STATE_FIELDS = (("first_scatter", np.uint8), ("last_val", np.float32), ("last_gx", np.float32), ("last_gy", np.float32), ("last_gz", np.float32),
                ("bounce", np.int32), ("last_gp_id", np.int32), ("info_t", np.float32))
CTX_NAMES = ("ctx_value_scale", "ctx_gs_x", "ctx_gs_y", "ctx_gs_z", "ctx_org_x", "ctx_org_y", "ctx_org_z")
SAMPLE_FIELDS = (("ok", np.uint8), ("exited", np.uint8), ("t", np.float32), ("continued_t", np.float32), ("gx", np.float32), ("gy", np.float32),
                 ("gz", np.float32), ("gp_id", np.int32), ("scheme", np.uint8), ("weight", np.float32))


def _alloc_like(ref, n, dtype):
    if _is_torch(ref):
        tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8, np.uint32: torch.int32}[dtype]
        return torch.empty(n, dtype=tdt, device=ref.device)
    return np.empty(n, dtype=dtype)


def new_state(n, like):
    """MediumState::reset() for n paths (media/Medium.hpp:71-80) plus zeroed conditioning coefficients."""
    st = {}
    for name, dt in STATE_FIELDS:
        a = _alloc_like(like, n, dt)
        if _is_torch(a):
            a.zero_()
        else:
            a[...] = 0
        st[name] = a
    if _is_torch(st["first_scatter"]):
        st["first_scatter"].fill_(1)
    else:
        st["first_scatter"][...] = 1
    st["scheme"] = _alloc_like(like, n, np.uint8)
    for name in CTX_NAMES:
        a = _alloc_like(like, n, np.float32)
        if _is_torch(a):
            a.zero_()
        else:
            a[...] = 0
        st[name] = a
    if _is_torch(st["scheme"]):
        st["scheme"].zero_()
    else:
        st["scheme"][...] = 0
    return st


class SparseConvolutionNoiseMedium:
    """Drop-in for the reference's SparseConvolutionNoiseMedium on the batched query path."""

    def __init__(self, desc, device=0):
        self._lib = abi.load()
        self.desc = desc
        self.device = device
        h = abi.MediumP()
        abi.check(self._lib.scn_medium_create(C.byref(desc), device, C.byref(h)))
        self._h = h

    # ---- construction (fromJson) -------------------------------------------------------------------------------
    @staticmethod
    def desc_from_json(medium_json, base_dir=None):
        """Parse the reference's medium JSON (a str or a dict) into a flat descriptor; raises ScnError like FAIL().  `base_dir` is the
        directory relative resource paths (regular_grid "file") resolve against — the scene file's directory in Tungsten.  A descriptor
        with regular_grid fields owns their host arrays: release it with scn_desc_release (from_json does)."""
        if not isinstance(medium_json, (str, bytes)):
            medium_json = json.dumps(medium_json)
        if isinstance(medium_json, str):
            medium_json = medium_json.encode()
        lib = abi.load()
        d = abi.MediumDesc()
        abi.check(lib.scn_desc_from_json_at(medium_json, None if base_dir is None else os.fsencode(base_dir), C.byref(d)))
        return d

    @classmethod
    def from_json(cls, medium_json, device=0, base_dir=None):
        d = cls.desc_from_json(medium_json, base_dir)
        try:
            return cls(d, device)
        finally:
            abi.load().scn_desc_release(C.byref(d))  # the medium holds its own device copy of every grid

    def info(self):
        """scn_medium_info: which kernel family serves this medium ("k_march", "k_march1d", "generic"; "k_fast_noise3d" for the noise
        evaluation), block size, cache capacity, shared memory per block, lattice cell size."""
        i = abi.MediumInfo()
        abi.check(self._lib.scn_medium_info(self._h, C.byref(i)))
        return {"query_kernel": abi.KERNEL_NAMES[i.query_kernel], "eval_kernel": abi.KERNEL_NAMES[i.eval_kernel], "march_block": i.march_block,
                "cache_entries": i.cache_entries, "smem_bytes": i.smem_bytes, "device": i.device, "sm_count": i.sm_count, "kernel_radius": i.kernel_radius}

    def force_generic(self, on=True):
        """Test hook (scn_debug_force_generic): serve the queries of this medium with the generic one-thread-per-ray kernels."""
        abi.check(self._lib.scn_debug_force_generic(self._h, int(bool(on))))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.scn_medium_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- evaluateNoise3D / evaluateValue / evaluateGradient --------------------------------------------------------
    def _eval_in(self, a, p, d, t=None, info_t=None, pix=None, scene_seed=0, noise_seed=0, ctx=None):
        ein = abi.EvalIn()
        for k, arr in zip(("px", "py", "pz"), p):
            setattr(ein, k, a.ptr(arr, np.float32, k))
        for k, arr in zip(("dx", "dy", "dz"), d):
            setattr(ein, k, a.ptr(arr, np.float32, k))
        ein.t = a.ptr(t, np.float32, "t", optional=True)
        ein.info_t = a.ptr(info_t, np.float32, "info_t", optional=True)
        if pix is not None:
            for k, arr in zip(("pix_x", "pix_y", "pix_spp", "pix_seg"), pix):
                setattr(ein, k, a.ptr(arr, np.uint32, k))
        ein.scene_seed = int(scene_seed) & 0xFFFFFFFF
        ein.noise_seed = int(noise_seed) & 0xFFFFFFFF
        if ctx is not None:
            for i, arr in enumerate(ctx):
                ein.ctx[i] = a.ptr(arr, np.float32, "ctx[%d]" % i)
        return ein

    def eval_noise(self, p, d, seed, stream=None):
        """evaluateNoise3D(p, dir, seed, ., conditioning=false) -> (val, gx, gy, gz) (SCN.cpp:101-124); 1D media: (val, 0, 0, gz)."""
        a = _Args()
        ein = self._eval_in(a, p, d, noise_seed=seed)
        outs = [_alloc_like(p[0], a.n, np.float32) for _ in range(4)]
        eo = abi.EvalOut()
        for k, arr in zip(("val", "gx", "gy", "gz"), outs):
            setattr(eo, k, a.ptr(arr, np.float32, k))
        if a.on_device:
            abi.check(self._lib.scn_eval_batch(self._h, C.byref(ein), C.byref(eo), a.n, abi.EVAL_NOISE, _stream_ptr(stream)))
        else:
            abi.check(self._lib.scn_eval_batch_host(self._h, C.byref(ein), C.byref(eo), a.n, abi.EVAL_NOISE))
        return tuple(outs)

    def eval_noise_into(self, p, d, seed, outs, stream=None):
        """Same as eval_noise with caller-owned outputs (no allocation: the bench's timed region uses this)."""
        a = _Args()
        ein = self._eval_in(a, p, d, noise_seed=seed)
        eo = abi.EvalOut()
        for k, arr in zip(("val", "gx", "gy", "gz"), outs):
            setattr(eo, k, a.ptr(arr, np.float32, k))
        if a.on_device:
            abi.check(self._lib.scn_eval_batch(self._h, C.byref(ein), C.byref(eo), a.n, abi.EVAL_NOISE, _stream_ptr(stream)))
        else:
            abi.check(self._lib.scn_eval_batch_host(self._h, C.byref(ein), C.byref(eo), a.n, abi.EVAL_NOISE))
        return outs

    def evaluate(self, p, d, t=None, info_t=None, pix=None, scene_seed=0, ctx=None, value=True, gradient=True, stream=None):
        """evaluateValue / evaluateGradient (SCN.cpp:73-99) -> dict(val, gp_id, gx, gy, gz)."""
        a = _Args()
        ein = self._eval_in(a, p, d, t=t, info_t=info_t, pix=pix, scene_seed=scene_seed, ctx=ctx)
        eo = abi.EvalOut()
        res = {}
        what = 0
        if value:
            what |= abi.EVAL_VALUE
            res["val"] = _alloc_like(p[0], a.n, np.float32)
            res["gp_id"] = _alloc_like(p[0], a.n, np.int32)
            eo.val = a.ptr(res["val"], np.float32, "val")
            eo.gp_id = a.ptr(res["gp_id"], np.int32, "gp_id")
        if gradient:
            what |= abi.EVAL_GRADIENT
            for k in ("gx", "gy", "gz"):
                res[k] = _alloc_like(p[0], a.n, np.float32)
                setattr(eo, k, a.ptr(res[k], np.float32, k))
        if a.on_device:
            abi.check(self._lib.scn_eval_batch(self._h, C.byref(ein), C.byref(eo), a.n, what, _stream_ptr(stream)))
        else:
            abi.check(self._lib.scn_eval_batch_host(self._h, C.byref(ein), C.byref(eo), a.n, what))
        return res

    def conditioning(self, p, d, target_val, target_grad, info_t=None, pix=None, scene_seed=0, stream=None):
        """SparseConvolutionNoiseRealization::conditioning (SCN.cpp:431-595) -> 7 coefficient arrays (device buffers only)."""
        a = _Args()
        ein = self._eval_in(a, p, d, info_t=info_t, pix=pix, scene_seed=scene_seed)
        tv = a.ptr(target_val, np.float32, "target_val")
        tg = [a.ptr(x, np.float32, "target_grad") for x in target_grad]
        if not a.on_device:
            raise ValueError("conditioning takes device buffers")
        outs = [_alloc_like(p[0], a.n, np.float32) for _ in range(7)]
        arr = (abi.f32p * 7)(*[a.ptr(o, np.float32, "ctx_out") for o in outs])
        abi.check(self._lib.scn_conditioning_batch(self._h, C.byref(ein), tv, tg[0], tg[1], tg[2], C.byref(arr), a.n, _stream_ptr(stream)))
        return outs

    # ---- Medium::sampleDistance / transmittance ------------------------------------------------------------------
    def _rays(self, a, rays, scene_seed):
        r = abi.Rays()
        for k in RAY_F:
            setattr(r, k, a.ptr(rays[k], np.float32, k))
        if rays.get("px") is not None:
            for k in RAY_U:
                setattr(r, k, a.ptr(rays[k], np.uint32, k))
        r.scene_seed = int(scene_seed) & 0xFFFFFFFF
        return r

    def _state(self, a, state):
        s = abi.State()
        for k, dt in STATE_FIELDS:
            setattr(s, k, a.ptr(state[k], dt, k))
        s.scheme = a.ptr(state.get("scheme"), np.uint8, "scheme", optional=True)
        if state.get(CTX_NAMES[0]) is not None:
            for i, k in enumerate(CTX_NAMES):
                s.ctx[i] = a.ptr(state[k], np.float32, k)
        return s

    def sample_distance(self, rays, state, scene_seed=0, out=None, count_evals=False, stream=None):
        """Medium::sampleDistance for a wavefront (GaussianProcessMedium.cpp:221-341).  `state` is updated in place.
        Returns dict(ok, exited, t, continued_t, gx, gy, gz, gp_id, scheme, weight[, n_evals])."""
        a = _Args()
        r = self._rays(a, rays, scene_seed)
        s = self._state(a, state)
        if out is None:
            out = {k: _alloc_like(rays["ox"], a.n, dt) for k, dt in SAMPLE_FIELDS}
            if count_evals:
                out["n_evals"] = _alloc_like(rays["ox"], a.n, np.uint32)
        o = abi.Samples()
        for k, dt in SAMPLE_FIELDS:
            setattr(o, k, a.ptr(out[k], dt, k))
        if out.get("n_evals") is not None:
            o.n_evals = a.ptr(out["n_evals"], np.uint32, "n_evals")
        if a.on_device:
            abi.check(self._lib.scn_sample_distance_batch(self._h, C.byref(r), C.byref(s), C.byref(o), a.n, _stream_ptr(stream)))
        else:
            abi.check(self._lib.scn_sample_distance_batch_host(self._h, C.byref(r), C.byref(s), C.byref(o), a.n))
        return out

    def sample_distance_static(self, rays, state, out, scene_seed=0, stream=None):
        """scn_sample_distance_batch_static (no counterpart in the reference): sample_distance for 1D-sampling media over the library's
        file-scope scratch array — device buffers only, every output buffer supplied by the caller (nothing is allocated here or in the
        library), at most abi.STATIC_SCRATCH_MAX_RAYS rays, and calls that overlap on the GPU race."""
        a = _Args()
        r = self._rays(a, rays, scene_seed)
        s = self._state(a, state)
        o = abi.Samples()
        for k, dt in SAMPLE_FIELDS:
            setattr(o, k, a.ptr(out[k], dt, k))
        if out.get("n_evals") is not None:
            o.n_evals = a.ptr(out["n_evals"], np.uint32, "n_evals")
        if not a.on_device:
            raise ValueError("sample_distance_static takes device buffers only")
        abi.check(self._lib.scn_sample_distance_batch_static(self._h, C.byref(r), C.byref(s), C.byref(o), a.n, _stream_ptr(stream)))
        return out

    def transmittance(self, rays, state, scene_seed=0, out=None, stream=None):
        """Medium::transmittance for a wavefront (GaussianProcessMedium.cpp:343-393): 1 where the ray exits uncrossed, else 0."""
        a = _Args()
        r = self._rays(a, rays, scene_seed)
        s = self._state(a, state)
        if out is None:
            out = _alloc_like(rays["ox"], a.n, np.float32)
        T = a.ptr(out, np.float32, "T")
        if a.on_device:
            abi.check(self._lib.scn_transmittance_batch(self._h, C.byref(r), C.byref(s), T, a.n, _stream_ptr(stream)))
        else:
            abi.check(self._lib.scn_transmittance_batch_host(self._h, C.byref(r), C.byref(s), T, a.n))
        return out

    @staticmethod
    def pdf(n):
        """Medium::pdf is the constant 1 for GPIS media (GaussianProcessMedium.cpp:395-398)."""
        return np.ones(n, dtype=np.float32)

    # ---- 1D next-event estimation -----------------------------------------------------------------------------------
    def _nee_in(self, a, d, normal, p, t_segment, info_t, pix, scene_seed, ctx):
        n = abi.NeeIn()
        for k, arr in zip(("dx", "dy", "dz"), d):
            setattr(n, k, a.ptr(arr, np.float32, k))
        for k, arr in zip(("nx", "ny", "nz"), normal):
            setattr(n, k, a.ptr(arr, np.float32, k))
        for k, arr in zip(("px", "py", "pz"), p):
            setattr(n, k, a.ptr(arr, np.float32, k))
        n.t_segment = a.ptr(t_segment, np.float32, "t_segment", optional=True)
        n.info_t = a.ptr(info_t, np.float32, "info_t", optional=True)
        if pix is not None:
            for k, arr in zip(("pix_x", "pix_y", "pix_spp", "pix_seg"), pix):
                setattr(n, k, a.ptr(arr, np.uint32, k))
        n.scene_seed = int(scene_seed) & 0xFFFFFFFF
        if ctx is not None:
            for i, arr in enumerate(ctx):
                n.ctx[i] = a.ptr(arr, np.float32, "ctx[%d]" % i)
        return n

    def nee_pdf(self, d, normal, p, t_segment, info_t=None, pix=None, scene_seed=0, ctx=None, stream=None):
        """noise.neePDF (SCN.cpp:661-743); device buffers."""
        a = _Args()
        n = self._nee_in(a, d, normal, p, t_segment, info_t, pix, scene_seed, ctx)
        out = _alloc_like(p[0], a.n, np.float32)
        abi.check(self._lib.scn_nee_pdf_batch(self._h, C.byref(n), a.ptr(out, np.float32, "pdf"), a.n, _stream_ptr(stream)))
        return out

    def nee_grad(self, d, normal, p, info_t=None, pix=None, scene_seed=0, ctx=None, stream=None):
        """noise.neeGrad (SCN.cpp:652-658); device buffers."""
        a = _Args()
        n = self._nee_in(a, d, normal, p, None, info_t, pix, scene_seed, ctx)
        outs = [_alloc_like(p[0], a.n, np.float32) for _ in range(3)]
        ptrs = [a.ptr(o, np.float32, "g") for o in outs]
        abi.check(self._lib.scn_nee_grad_batch(self._h, C.byref(n), ptrs[0], ptrs[1], ptrs[2], a.n, _stream_ptr(stream)))
        return outs


class MediumFactory:
    """String registry like the reference's MediumFactory (media/MediumFactory.cpp:13-22), for the one type built here."""

    @staticmethod
    def create(medium_json, device=0):
        j = json.loads(medium_json) if isinstance(medium_json, (str, bytes)) else medium_json
        if j.get("type") != MEDIUM_TYPE:
            raise abi.ScnError(abi.SCN_EJSON, "Invalid medium: '%s'" % j.get("type"))
        return SparseConvolutionNoiseMedium.from_json(j, device)
