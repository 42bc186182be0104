"""Host side of the wavefront batching stage (include/scn_render.h): the loop PathTracer::traceSample runs per path
(integrators/path_tracer/PathTracer.cpp:62-169) run per WAVEFRONT, and the tile scheduling of PathTraceIntegrator
(integrators/path_tracer/PathTraceIntegrator.cpp:26-41, 135-163, 227-246) with GPUs in place of pool threads: every rank owns a
contiguous range of the 16 x 16 tiles and the framebuffer slice of those tiles; the only collective is one all_gather of the
slices at the end of the render (`gather_image`).  torch is plumbing here: device buffers, streams, torch.distributed.

Nothing in this module computes radiance: rays, shading and accumulation are the CUDA kernels of csrc/scn_render.cu, the medium
query is scn_sample_distance_batch.  There is no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import abi
from .medium import CTX_NAMES, RAY_F, RAY_U, SAMPLE_FIELDS, STATE_FIELDS, SparseConvolutionNoiseMedium

TILE = 16


class WfScene(C.Structure):
    _fields_ = [("cam_pos", C.c_float * 3), ("cam_look_at", C.c_float * 3), ("cam_up", C.c_float * 3), ("fov_deg", C.c_float),
                ("res_x", C.c_uint32), ("res_y", C.c_uint32), ("cube_center", C.c_float * 3), ("cube_half", C.c_float * 3),
                ("emission", C.c_float), ("albedo", C.c_float), ("max_bounces", C.c_int32), ("scene_seed", C.c_uint32),
                ("nee", C.c_int32), ("phase_type", C.c_int32), ("albedo_rgb", C.c_float * 3), ("eta", C.c_float * 3), ("k", C.c_float * 3),
                ("env_rgb", C.c_float * 3), ("env_sample", C.c_int32), ("has_quad", C.c_int32), ("quad_first", C.c_int32),
                ("quad_base", C.c_float * 3), ("quad_edge0", C.c_float * 3), ("quad_edge1", C.c_float * 3), ("quad_rgb", C.c_float * 3)]


class WfPaths(C.Structure):
    _fields_ = [("rays", abi.Rays), ("state", abi.State), ("samples", abi.Samples), ("throughput", abi.f32p), ("radiance", abi.f32p), ("alive", abi.u8p),
                ("throughput_gb", abi.f32p * 2), ("radiance_gb", abi.f32p * 2), ("scattered", abi.u8p)]


class WfShadow(C.Structure):
    _fields_ = [("rays", abi.Rays), ("state", abi.State), ("T", abi.f32p), ("contrib", abi.f32p * 3)]


PHASE_LAMBERT, PHASE_MIRROR, PHASE_CONDUCTOR = 0, 1, 2
# bsdfs/ComplexIorData.hpp (the materials the BASELINE configs name)
COMPLEX_IOR = {"Cu": ((0.2004376970, 0.9240334304, 1.1022119527), (3.9129485033, 2.4528477015, 2.1421879552))}


RENDER_SYMBOLS = {
    "scn_wf_tile_count": (C.c_uint32, [C.c_uint32, C.c_uint32]),
    "scn_wf_raygen": (C.c_int, [C.POINTER(WfScene), abi.MediumP, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(WfPaths), C.c_size_t, C.c_void_p]),
    "scn_wf_raygen_strided": (C.c_int, [C.POINTER(WfScene), abi.MediumP, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(WfPaths), C.c_size_t,
                                        C.c_void_p]),
    "scn_wf_shade": (C.c_int, [C.POINTER(WfScene), abi.MediumP, C.POINTER(WfPaths), C.c_size_t, C.c_int32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
    "scn_wf_accumulate": (C.c_int, [abi.f32p, abi.f32p, C.c_size_t, C.c_uint32, C.c_void_p]),
    "scn_wf_nee_setup": (C.c_int, [C.POINTER(WfScene), abi.MediumP, C.POINTER(WfPaths), C.POINTER(WfShadow), C.POINTER(WfShadow), C.c_size_t, C.c_int32, C.c_void_p]),
    "scn_wf_nee_gather": (C.c_int, [C.POINTER(WfPaths), C.POINTER(WfShadow), C.POINTER(WfShadow), C.c_size_t, C.c_void_p]),
    "scn_wf_clamp_samples": (C.c_int, [C.POINTER(WfPaths), C.c_size_t, C.c_void_p]),
    "scn_wf_render": (C.c_int, [C.POINTER(WfScene), abi.MediumP, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(WfPaths),
                                C.POINTER(WfShadow), C.POINTER(WfShadow), C.POINTER(abi.f32p * 3), C.c_void_p]),
}


def load():
    lib = abi.load()
    for name, (res, args) in RENDER_SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def scene_from_json(scene):
    """The restricted scene out of a Tungsten scene dict: pinhole camera, one cube holding the medium,
    constant infinite_sphere, lambert albedo of the medium's brdf phase function, integrator max_bounces."""
    s = WfScene()
    cam = scene["camera"]
    tr = cam.get("transform", {})
    s.cam_pos[:] = tr.get("position", [0, 0, 5]); s.cam_look_at[:] = tr.get("look_at", [0, 0, 0]); s.cam_up[:] = tr.get("up", [0, 1, 0])
    s.fov_deg = float(cam.get("fov", 60.0))
    res = cam.get("resolution", [64, 64])
    s.res_x, s.res_y = (int(res), int(res)) if np.isscalar(res) else (int(res[0]), int(res[1]))
    cube = next(p for p in scene["primitives"] if p["type"] == "cube")
    ctr = cube.get("transform", {})
    scale = ctr.get("scale", [1, 1, 1])
    scale = [scale] * 3 if np.isscalar(scale) else scale
    s.cube_center[:] = ctr.get("position", [0, 0, 0]); s.cube_half[:] = [0.5 * float(v) for v in scale]
    env = next((p for p in scene["primitives"] if p["type"] == "infinite_sphere"), None)
    em = env.get("emission", 1.0) if env else 0.0
    s.emission = float(em if np.isscalar(em) else np.mean(em))
    med = scene["media"][0]
    alb = med.get("phase_function", {}).get("bsdf", {}).get("albedo", 1.0)
    s.albedo = float(alb if np.isscalar(alb) else np.mean(alb))
    s.max_bounces = int(scene.get("integrator", {}).get("max_bounces", 64))
    s.scene_seed = int(scene.get("renderer", {}).get("seed", 0xBA5EBA11)) & 0xFFFFFFFF
    # ---- nee stage: phase bsdf, RGB emitters, one quad light (include/scn_render.h)
    rgb = lambda v, d=1.0: [float(v)] * 3 if np.isscalar(v) else [float(x) for x in v]
    bsdf = med.get("phase_function", {}).get("bsdf", {})
    kind = bsdf.get("type", "lambert")
    s.albedo_rgb[:] = rgb(bsdf.get("albedo", 1.0 if kind != "lambert" else 0.5))
    if kind == "lambert":
        s.phase_type = PHASE_LAMBERT
        s.albedo_rgb[:] = rgb(alb)
    elif kind == "mirror":
        s.phase_type = PHASE_MIRROR
    elif kind == "conductor":
        s.phase_type = PHASE_CONDUCTOR
        if "eta" in bsdf and "k" in bsdf:
            s.eta[:] = rgb(bsdf["eta"]); s.k[:] = rgb(bsdf["k"])
        else:
            name = bsdf.get("material", "Cu")
            if name not in COMPLEX_IOR:
                raise ValueError("conductor material %r is not in this stage's table" % name)
            s.eta[:], s.k[:] = COMPLEX_IOR[name]
    else:
        raise ValueError("phase bsdf %r is not covered by the wavefront stage (lambert, mirror, conductor)" % kind)
    s.env_rgb[:] = rgb(em) if env else [0.0] * 3
    s.env_sample = int(bool(env.get("sample", True))) if env else 0
    prims = scene["primitives"]
    quad = next((p for p in prims if p["type"] == "quad" and "emission" in p), None)
    s.has_quad = int(quad is not None)
    if quad is not None:
        base, e0, e1 = quad_geometry(quad.get("transform", {}))
        s.quad_base[:], s.quad_edge0[:], s.quad_edge1[:] = base, e0, e1
        s.quad_rgb[:] = rgb(quad["emission"])
        s.quad_first = int(env is None or prims.index(quad) < prims.index(env))
    specular = s.phase_type != PHASE_LAMBERT
    s.nee = int(bool(scene.get("integrator", {}).get("scn_nee", specular or quad is not None)))
    return s


def quad_geometry(tr):
    """Quad::prepareForRender (primitives/Quad.cpp:296-304) over the JSON transform (io/JsonUtils.cpp: scale, then rotation
    "rotation" in degrees through Mat4f::rotYXZ, then position; io/JsonPtr.cpp:108-183): base corner and the two edges."""
    pos = np.array(tr.get("position", [0, 0, 0]), np.float64)
    sc = tr.get("scale", [1, 1, 1]); sc = np.array([sc] * 3 if np.isscalar(sc) else sc, np.float64)
    r = np.radians(np.array(tr.get("rotation", [0, 0, 0]), np.float64))
    c, sn = np.cos(r), np.sin(r)
    R = np.array([[c[1] * c[2] - sn[1] * sn[0] * sn[2], -c[1] * sn[2] - sn[1] * sn[0] * c[2], -sn[1] * c[0]],
                  [c[0] * sn[2], c[0] * c[2], -sn[0]],
                  [sn[1] * c[2] + c[1] * sn[0] * sn[2], -sn[1] * sn[2] + c[1] * sn[0] * c[2], c[1] * c[0]]])  # Mat4f::rotYXZ (math/Mat4f.cpp:118-130)
    M = R @ np.diag(sc)  # io/JsonPtr.cpp:161-174: axes scaled, then rotated
    e0, e1 = M @ np.array([1.0, 0, 0]), M @ np.array([0, 0, 1.0])
    base = pos - 0.5 * e0 - 0.5 * e1
    return [float(v) for v in base], [float(v) for v in e0], [float(v) for v in e1]


def tile_range(n_tiles, rank, world):
    """Contiguous tile range of a rank: tiles in diceTiles order split into `world` near-equal ranges."""
    base, rem = divmod(n_tiles, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def tile_set(n_tiles, rank, world, sharding="contiguous"):
    """(tile_first, tile_stride, tile_count) of a rank.  "contiguous" is `tile_range`; "interleaved" deals the tiles round-robin
    (rank, rank + world, ...), which SURVEY.md §8(e) asks for when contiguous ranges leave some GPUs with background-only tiles."""
    if sharding == "contiguous" or world == 1:
        first, count = tile_range(n_tiles, rank, world)
        return first, 1, count
    if sharding != "interleaved":
        raise ValueError("sharding must be 'contiguous' or 'interleaved'")
    return rank, world, (n_tiles - rank + world - 1) // world if rank < n_tiles else 0


class WavefrontRenderer:
    """One GPU's share of a render: tiles tile_first + k * tile_stride, k < tile_count."""

    def __init__(self, scene_json, device=0, rank=0, world=1, max_paths=1 << 22, nee=None, sharding="contiguous"):
        self.lib = load()
        self.scene = scene_from_json(scene_json)
        if nee is not None:
            self.scene.nee = int(bool(nee))
        self.nee = bool(self.scene.nee)
        self.dev = torch.device("cuda", device)
        self.medium = SparseConvolutionNoiseMedium.from_json(scene_json["media"][0], device)
        self.n_tiles = int(self.lib.scn_wf_tile_count(self.scene.res_x, self.scene.res_y))
        self.sharding = sharding if world > 1 else "contiguous"
        self.tile_first, self.tile_stride, self.tile_count = tile_set(self.n_tiles, rank, world, self.sharding)
        self.n_pixels = self.tile_count * TILE * TILE
        self.max_paths = max_paths
        self.framebuffer = torch.zeros(max(self.n_pixels, 1), dtype=torch.float32, device=self.dev)  # red (or grey) plane
        self.framebuffer_gb = [torch.zeros(max(self.n_pixels, 1), dtype=torch.float32, device=self.dev) for _ in range(2)] if self.nee else None
        self.spp_done = 0
        self.queries = 0
        self._buf_n = 0

    def _alloc(self, n):
        if n == self._buf_n:
            return
        tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8, np.uint32: torch.int32}
        mk = lambda dt: torch.empty(n, dtype=tdt[dt], device=self.dev)
        self.rays = {k: mk(np.float32) for k in RAY_F}
        self.rays.update({k: mk(np.uint32) for k in RAY_U})
        self.state = {k: mk(dt) for k, dt in STATE_FIELDS}
        self.state["scheme"] = mk(np.uint8)
        self.state.update({k: mk(np.float32) for k in CTX_NAMES})
        self.samples = {k: mk(dt) for k, dt in SAMPLE_FIELDS}
        self.throughput, self.radiance, self.alive = mk(np.float32), mk(np.float32), mk(np.uint8)
        p = WfPaths()
        ptr = lambda t, ct: C.cast(C.c_void_p(t.data_ptr()), ct)
        for k in RAY_F:
            setattr(p.rays, k, ptr(self.rays[k], abi.f32p))
        for k in RAY_U:
            setattr(p.rays, k, ptr(self.rays[k], abi.u32p))
        p.rays.scene_seed = self.scene.scene_seed
        ct = {np.float32: abi.f32p, np.int32: abi.i32p, np.uint8: abi.u8p}
        for k, dt in STATE_FIELDS:
            setattr(p.state, k, ptr(self.state[k], ct[dt]))
        p.state.scheme = ptr(self.state["scheme"], abi.u8p)
        for i, k in enumerate(CTX_NAMES):
            p.state.ctx[i] = ptr(self.state[k], abi.f32p)
        for k, dt in SAMPLE_FIELDS:
            setattr(p.samples, k, ptr(self.samples[k], ct[dt]))
        p.throughput, p.radiance, p.alive = ptr(self.throughput, abi.f32p), ptr(self.radiance, abi.f32p), ptr(self.alive, abi.u8p)
        if self.nee:
            self.thr_gb, self.rad_gb, self.scattered = [mk(np.float32), mk(np.float32)], [mk(np.float32), mk(np.float32)], mk(np.uint8)
            for j in range(2):
                p.throughput_gb[j] = ptr(self.thr_gb[j], abi.f32p); p.radiance_gb[j] = ptr(self.rad_gb[j], abi.f32p)
            p.scattered = ptr(self.scattered, abi.u8p)
            self.shadow = []
            for _ in range(2):  # light-sampled and phase-sampled shadow rays (volumeLightSample / volumePhaseSample)
                h = WfShadow()
                rays = {k: mk(np.float32) for k in RAY_F}; rays.update({k: mk(np.uint32) for k in RAY_U})
                state = {k: mk(dt) for k, dt in STATE_FIELDS}
                for k in RAY_F:
                    setattr(h.rays, k, ptr(rays[k], abi.f32p))
                for k in RAY_U:
                    setattr(h.rays, k, ptr(rays[k], abi.u32p))
                h.rays.scene_seed = self.scene.scene_seed
                for k, dt in STATE_FIELDS:
                    setattr(h.state, k, ptr(state[k], ct[dt]))
                T, contrib = mk(np.float32), [mk(np.float32) for _ in range(3)]
                h.T = ptr(T, abi.f32p)
                for j in range(3):
                    h.contrib[j] = ptr(contrib[j], abi.f32p)
                self.shadow.append({"c": h, "rays": rays, "state": state, "T": T, "contrib": contrib})
        self.paths = p
        self._buf_n = n

    def prepare(self, spp):
        """Allocate the wavefront buffers a render(spp) call will use (so that a timed render does not include cudaMalloc)."""
        if self.n_pixels:
            self._alloc(self.n_pixels * max(1, min(spp, self.max_paths // self.n_pixels)))

    def render(self, spp, count_queries=False, stepwise=False):
        """Add `spp` samples per pixel to this rank's framebuffer slice.  Samples are processed in chunks so that a wavefront
        holds at most max_paths paths.  The loop over chunks and bounces runs inside the library (scn_wf_render: one call, no host
        synchronisation); `stepwise` (and `count_queries`, which needs a host read per bounce) drives the same kernels launch by
        launch through the per-stage entry points instead — the tests compare the two."""
        if self.n_pixels == 0:
            return
        stream = torch.cuda.current_stream(self.dev)
        sp = C.c_void_p(stream.cuda_stream)
        chunk = max(1, min(spp, self.max_paths // self.n_pixels))
        if not (count_queries or stepwise):
            self._alloc(self.n_pixels * chunk)
            fbs = (abi.f32p * 3)()
            planes = [self.framebuffer] + (self.framebuffer_gb if self.nee else [])
            for j, fb in enumerate(planes):
                fbs[j] = C.cast(C.c_void_p(fb.data_ptr()), abi.f32p)
            L, Q = (self.shadow if self.nee else (None, None))
            abi.check(self.lib.scn_wf_render(C.byref(self.scene), self.medium._h, self.tile_first, self.tile_stride, self.tile_count, self.spp_done, spp, chunk,
                                             C.byref(self.paths), C.byref(L["c"]) if L else None, C.byref(Q["c"]) if Q else None, C.byref(fbs), sp))
            self.spp_done += spp
            return
        done = 0
        while done < spp:
            c = min(chunk, spp - done)
            n = self.n_pixels * c
            self._alloc(n)
            abi.check(self.lib.scn_wf_raygen_strided(C.byref(self.scene), self.medium._h, self.tile_first, self.tile_stride, self.tile_count, self.spp_done + done, c,
                                                     C.byref(self.paths), n, sp))
            for bounce in range(self.scene.max_bounces):
                if count_queries:
                    self.queries += int(self.alive.sum().item())
                self.medium.sample_distance(self.rays, self.state, scene_seed=self.scene.scene_seed, out=self.samples, stream=stream)
                if self.nee and bounce < self.scene.max_bounces - 1:  # volumeEstimateDirect at every scatter of this bounce
                    L, Q = self.shadow
                    abi.check(self.lib.scn_wf_nee_setup(C.byref(self.scene), self.medium._h, C.byref(self.paths), C.byref(L["c"]), C.byref(Q["c"]), n, bounce, sp))
                    for h in (L, Q):
                        self.medium.transmittance(h["rays"], h["state"], scene_seed=self.scene.scene_seed, out=h["T"], stream=stream)
                    abi.check(self.lib.scn_wf_nee_gather(C.byref(self.paths), C.byref(L["c"]), C.byref(Q["c"]), n, sp))
                abi.check(self.lib.scn_wf_shade(C.byref(self.scene), self.medium._h, C.byref(self.paths), n, bounce, self.spp_done + done, c, self.tile_first, sp))
            if self.nee:
                abi.check(self.lib.scn_wf_clamp_samples(C.byref(self.paths), n, sp))
            abi.check(self.lib.scn_wf_accumulate(C.cast(C.c_void_p(self.radiance.data_ptr()), abi.f32p), C.cast(C.c_void_p(self.framebuffer.data_ptr()), abi.f32p),
                                                 self.n_pixels, c, sp))
            if self.nee:
                for j in range(2):
                    abi.check(self.lib.scn_wf_accumulate(C.cast(C.c_void_p(self.rad_gb[j].data_ptr()), abi.f32p), C.cast(C.c_void_p(self.framebuffer_gb[j].data_ptr()), abi.f32p),
                                                         self.n_pixels, c, sp))
            done += c
        self.spp_done += spp

    def local_image(self, channel=0):
        """This rank's tiles as [tile_count, 16, 16] mean radiance (channel 0 = red or grey; 1, 2 = green, blue of an nee render)."""
        fb = self.framebuffer if channel == 0 or not self.nee else self.framebuffer_gb[channel - 1]
        return (fb[: self.n_pixels] / max(self.spp_done, 1)).view(self.tile_count, TILE, TILE)

    def close(self):
        self.medium.close()


def assemble_image(tiles, res_x, res_y):
    """[n_tiles, 16, 16] in diceTiles order -> [res_y, res_x] (edge tiles cropped)."""
    tx, ty = (res_x + TILE - 1) // TILE, (res_y + TILE - 1) // TILE
    img = tiles.view(ty, tx, TILE, TILE).permute(0, 2, 1, 3).reshape(ty * TILE, tx * TILE)
    return img[:res_y, :res_x].contiguous()


def gather_image_rgb(renderer, world=1):
    """[res_y, res_x, 3] on every rank with ONE all_gather: the three planes of an nee (RGB) render travel stacked; a grey render
    gathers its one plane and repeats it."""
    if not getattr(renderer, "nee", False):
        g = gather_image(renderer, world)
        return torch.stack([g, g, g], -1)
    local = torch.stack([renderer.local_image(ch) for ch in range(3)], 1)  # [tile_count, 3, 16, 16]
    tiles = _gather_tiles(renderer, local, world)
    return torch.stack([assemble_image(tiles[:, ch].contiguous(), renderer.scene.res_x, renderer.scene.res_y) for ch in range(3)], -1)


def _gather_tiles(renderer, local, world):
    """The one collective of a render: all ranks' framebuffer slices ([tile_count, ...] per rank) -> every tile in diceTiles order
    on every rank (NCCL all_gather over NVLink).  Slices are padded to the largest tile count so that the gather is a single call."""
    if world <= 1:
        return local
    import torch.distributed as dist
    cap = (renderer.n_tiles + world - 1) // world
    pad = torch.zeros((cap,) + tuple(local.shape[1:]), dtype=torch.float32, device=local.device)
    pad[: local.shape[0]] = local
    if dist.get_backend() == "nccl":
        flat = torch.empty((world * cap,) + tuple(local.shape[1:]), dtype=torch.float32, device=local.device)
        dist.all_gather_into_tensor(flat, pad)  # one NCCL all_gather straight into the assembled buffer
        out = list(flat.view((world, cap) + tuple(local.shape[1:])).unbind(0))
    else:
        out = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(out, pad)
    if getattr(renderer, "sharding", "contiguous") == "interleaved":  # tile t lives at out[t % world][t // world]
        return torch.stack(out, 1).reshape((world * cap,) + tuple(local.shape[1:]))[: renderer.n_tiles]
    return torch.cat([out[r][: tile_range(renderer.n_tiles, r, world)[1]] for r in range(world)], 0)


def gather_image(renderer, world=1, channel=0):
    """One plane of the image on every rank ([res_y, res_x]; 4 B per pixel through the all_gather)."""
    local = renderer.local_image(channel) if channel else renderer.local_image()
    return assemble_image(_gather_tiles(renderer, local, world), renderer.scene.res_x, renderer.scene.res_y)


def _as_rgb(img):
    a = np.asarray(img.detach().cpu() if hasattr(img, "detach") else img, dtype=np.float32)
    if a.ndim == 2:  # a grey render: the reference's framebuffer is RGB, the three channels are equal
        a = np.repeat(a[:, :, None], 3, axis=2)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("image must be [h, w] or [h, w, 3]")
    return a


def write_pfm(path, img):
    """The reference's HDR output (io/ImageIO.cpp:531-547 savePfm): "PF", "w h", "-1" (the stream prints -1.0 as -1), then the rows
    bottom-up as little-endian float32 RGB.  `img` is [h, w, 3] (or [h, w] grey), top row first."""
    a = _as_rgb(img)
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1\n" % (a.shape[1], a.shape[0]))
        f.write(np.ascontiguousarray(a[::-1]).astype("<f4").tobytes())


TONEMAPS = ("linear", "gamma", "reinhard", "filmic", "pbrt")


def tonemap(img, op="gamma"):
    """Camera::tonemap (cameras/Camera.hpp:159-162: negative values clamped to 0) over Tonemap::tonemap (cameras/Tonemap.hpp:25-50), float32
    like the reference (powf: evaluated in double and rounded, which is what glibc's powf does)."""
    with np.errstate(invalid="ignore", over="ignore", divide="ignore"):  # inf / NaN radiance propagates like in the reference
        return _tonemap(np.maximum(_as_rgb(img), np.float32(0)), op)


def _tonemap(c, op):
    f32 = np.float32
    pw = lambda x, e: np.power(x.astype(np.float64), float(f32(e))).astype(np.float32)
    if op == "linear":
        return c
    if op == "gamma":
        return pw(c, f32(1.0) / f32(2.2))
    if op == "reinhard":
        return pw(c / (c + f32(1.0)), f32(1.0) / f32(2.2))
    if op == "filmic":
        x = np.maximum(f32(0), c - f32(0.004))
        return (x * (f32(6.2) * x + f32(0.5))) / (x * (f32(6.2) * x + f32(1.7)) + f32(0.06))
    if op == "pbrt":
        return np.where(c < f32(0.0031308), f32(12.92) * c, f32(1.055) * pw(c, f32(1.0) / f32(2.4)) - f32(0.055)).astype(np.float32)
    raise ValueError("Invalid tonemap operator: '%s'" % op)


def to_ldr(img, op="gamma"):
    """Integrator::writeBuffers (integrators/Integrator.cpp:50-51): Vec3c(clamp(Vec3i(tonemap(c) * 255.0f), 0, 255)) — truncation, not rounding."""
    t = tonemap(img, op) * np.float32(255.0)
    t = np.where(np.isfinite(t), t, np.where(t > 0, np.float32(2e9), np.float32(-2e9)))  # int conversion of +-inf / NaN is undefined in the reference; clamp like 0 / 255
    return np.clip(t.astype(np.int64), 0, 255).astype(np.uint8)


def write_png(path, img, op="gamma"):
    """The reference's LDR output (io/ImageIO.cpp:590-611 savePng over Integrator.cpp:50-56): 8-bit RGB of the tonemapped image.  The
    encoder differs from lodepng (zlib stream, filter choice), the decoded pixels do not."""
    import struct
    import zlib
    ldr = to_ldr(img, op)
    h, w = ldr.shape[:2]
    raw = b"".join(b"\x00" + ldr[y].tobytes() for y in range(h))  # filter type 0 on every scanline
    chunk = lambda tag, data: struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw, 9)) + chunk(b"IEND", b""))


def write_outputs(scene_json, img, directory="."):
    """Integrator::saveOutputs for a finished render: renderer.output_file (PNG, the camera's "tonemap" operator, default "gamma":
    cameras/Camera.cpp:22) and renderer.hdr_output_file (PFM) relative to `directory`, like the reference writes them next to the
    scene file.  Returns the paths written."""
    import os
    r = scene_json.get("renderer", {})
    out = []
    if r.get("output_file"):
        if not r["output_file"].lower().endswith(".png"):
            raise ValueError("output_file: only PNG is covered (io/ImageIO.cpp:625-631)")
        write_png(os.path.join(directory, r["output_file"]), img, scene_json.get("camera", {}).get("tonemap", "gamma")); out.append(os.path.join(directory, r["output_file"]))
    if r.get("hdr_output_file"):
        if not r["hdr_output_file"].lower().endswith(".pfm"):
            raise ValueError("hdr_output_file: only PFM is covered (OpenEXR is not built: io/ImageIO.cpp:613-623)")
        write_pfm(os.path.join(directory, r["hdr_output_file"]), img); out.append(os.path.join(directory, r["hdr_output_file"]))
    return out
