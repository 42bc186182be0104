// scn_render.cu — wavefront batching stage around the medium query (include/scn_render.h): camera rays, analytic cube,
// Lambert "brdf" phase function about the sampled GP gradient, constant environment, Russian roulette.  One thread per path;
// every kernel reads and writes the SoA buffers scn_sample_distance_batch works on, so a bounce is
//   k_wf_shade (previous bounce) -> k_march -> k_wf_shade ...   with no host round trip and no compaction: dead paths are refused
// by the medium query at once (bounce >= max_bounces, GPM:235-237) and the persistent warps of k_march skip over them.
#include <cuda_runtime.h>

#include <math.h>

#include "../../include/scn_render.h"
#include "scn_device.cuh"

namespace scn {

constexpr int WF_BLOCK = 256;
constexpr int32_t WF_DEAD_BOUNCE = 0x3fffffff; // state.bounce of a finished path: the medium query refuses it immediately

// counter-based sampler of this stage: xxhash32 of (pixel, sample, bounce * 8 + dimension, scene seed) -> [0, 1)
__device__ __forceinline__ float wf_rand(uint32_t px, uint32_t py, uint32_t spp, uint32_t bounce, uint32_t dim, uint32_t scene_seed) {
    return u32_to_unit(xxhash32_4(px | (py << 16), spp, bounce * 8u + dim, scene_seed ^ 0x9e3779b9u));
}

// slab test of the ray o + t d against the cube; returns false when it misses (or lies behind the origin)
__device__ __forceinline__ bool wf_cube(const scn_wf_scene &S, float ox, float oy, float oz, float dx, float dy, float dz, float &t0, float &t1) {
    const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
    float tn = 0.f, tf = INFINITY;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const float inv = 1.0f / d[j];
        float a = (S.cube_center[j] - S.cube_half[j] - o[j]) * inv, b = (S.cube_center[j] + S.cube_half[j] - o[j]) * inv;
        if (d[j] == 0.f) { // parallel to the slab: inside or never
            const bool inside = fabsf(o[j] - S.cube_center[j]) <= S.cube_half[j];
            a = inside ? -INFINITY : INFINITY; b = inside ? INFINITY : -INFINITY;
        }
        tn = fmaxf(tn, fminf(a, b)); tf = fminf(tf, fmaxf(a, b));
    }
    t0 = tn; t1 = tf;
    return tf > tn;
}

struct WfPaths {
    float *ox, *oy, *oz, *dx, *dy, *dz, *tmin, *tmax, *u; uint32_t *px, *py, *spp, *seg;
    uint8_t *first_scatter; float *last_val, *lgx, *lgy, *lgz; int32_t *bounce, *last_gp_id; float *info_t; uint8_t *scheme; float *ctx[7];
    const uint8_t *ok, *exited; const float *t, *gx, *gy, *gz, *weight;
    float *throughput, *radiance; uint8_t *alive;
};

__device__ __forceinline__ void wf_finish(const WfPaths &W, size_t i, float radiance) {
    W.radiance[i] = radiance; W.alive[i] = 0; W.bounce[i] = WF_DEAD_BOUNCE;
}

// PathTracer::traceSample :31-59 — camera sample, first intersection with the scene (the cube), medium state reset
__global__ void __launch_bounds__(WF_BLOCK) k_wf_raygen(const scn_wf_scene S, const WfPaths W, uint32_t tile_first, uint32_t spp_first, uint32_t spp_count, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = (uint32_t)(i % spp_count); const size_t lp = i / spp_count;
    const uint32_t tiles_x = (S.res_x + SCN_WF_TILE - 1) / SCN_WF_TILE;
    const uint32_t tile = tile_first + (uint32_t)(lp / (SCN_WF_TILE * SCN_WF_TILE)), in_tile = (uint32_t)(lp % (SCN_WF_TILE * SCN_WF_TILE));
    const uint32_t x = (tile % tiles_x) * SCN_WF_TILE + in_tile % SCN_WF_TILE, y = (tile / tiles_x) * SCN_WF_TILE + in_tile / SCN_WF_TILE;
    const uint32_t spp = spp_first + s;
    // state.reset() (media/Medium.hpp:70-80) and RayInfo
    W.px[i] = x; W.py[i] = y; W.spp[i] = spp; W.seg[i] = 0u;
    W.first_scatter[i] = 1; W.last_val[i] = 0.f; W.lgx[i] = 0.f; W.lgy[i] = 0.f; W.lgz[i] = 0.f; W.bounce[i] = 0; W.last_gp_id[i] = 0; W.info_t[i] = 0.f;
    if (W.scheme) W.scheme[i] = SCN_SCHEME_UNI;
    if (W.ctx[0]) for (int j = 0; j < 7; ++j) W.ctx[j][i] = 0.f;
    W.throughput[i] = 1.f; W.radiance[i] = 0.f; W.alive[i] = 1;
    W.ox[i] = S.cam_pos[0]; W.oy[i] = S.cam_pos[1]; W.oz[i] = S.cam_pos[2]; W.dx[i] = 0.f; W.dy[i] = 0.f; W.dz[i] = 1.f; W.tmin[i] = 0.f; W.tmax[i] = 0.f; W.u[i] = 0.f;
    if (x >= S.res_x || y >= S.res_y) { wf_finish(W, i, 0.f); return; } // padding of an edge tile
    // Mat4f::lookAt (math/Mat4f.cpp:175-187) and PinholeCamera::sampleDirection (cameras/PinholeCamera.cpp:70-86), box filter
    f3 f = F3(S.cam_look_at[0] - S.cam_pos[0], S.cam_look_at[1] - S.cam_pos[1], S.cam_look_at[2] - S.cam_pos[2]);
    float il = rsqrtf(f.x * f.x + f.y * f.y + f.z * f.z); f = F3(f.x * il, f.y * il, f.z * il);
    const f3 up = F3(S.cam_up[0], S.cam_up[1], S.cam_up[2]);
    f3 r = F3(f.y * up.z - f.z * up.y, f.z * up.x - f.x * up.z, f.x * up.y - f.y * up.x);
    il = rsqrtf(r.x * r.x + r.y * r.y + r.z * r.z); r = F3(r.x * il, r.y * il, r.z * il);
    const f3 uu = F3(r.y * f.z - r.z * f.y, r.z * f.x - r.x * f.z, r.x * f.y - r.y * f.x);
    const float plane_dist = 1.0f / tanf(S.fov_deg * 0.017453292519943295f * 0.5f), ratio = (float)S.res_y / (float)S.res_x, pix = 1.0f / (float)S.res_x;
    const float jx = wf_rand(x, y, spp, 0u, 6u, S.scene_seed) - 0.5f, jy = wf_rand(x, y, spp, 0u, 7u, S.scene_seed) - 0.5f;
    f3 ld = F3(-1.0f + ((float)x + 0.5f + jx) * 2.0f * pix, ratio - ((float)y + 0.5f + jy) * 2.0f * pix, plane_dist);
    il = rsqrtf(ld.x * ld.x + ld.y * ld.y + ld.z * ld.z); ld = F3(ld.x * il, ld.y * il, ld.z * il);
    const f3 d = F3(r.x * ld.x + uu.x * ld.y + f.x * ld.z, r.y * ld.x + uu.y * ld.y + f.y * ld.z, r.z * ld.x + uu.z * ld.y + f.z * ld.z);
    W.dx[i] = d.x; W.dy[i] = d.y; W.dz[i] = d.z;
    float t0, t1;
    if (!wf_cube(S, S.cam_pos[0], S.cam_pos[1], S.cam_pos[2], d.x, d.y, d.z, t0, t1)) { wf_finish(W, i, S.emission); return; } // straight to the environment
    W.tmin[i] = t0; W.tmax[i] = t1; W.u[i] = wf_rand(x, y, spp, 0u, 0u, S.scene_seed);
}

// PathTracer::traceSample :69-169 for one path after its medium query of bounce `bounce`:
//   query refused (false)   -> the path returns what it has gathered (nothing): PathTracer.cpp:69-70
//   exited                  -> the forward surface of the cube passes the ray on to the environment: radiance = throughput * Le
//   scattered in the medium -> BRDFPhaseFunction (:27-44): frame about the normalised gradient, Lambert cosine sample, weight albedo;
//                              Russian roulette (:146-152); next ray from the scatter point to the far wall of the cube
__global__ void __launch_bounds__(WF_BLOCK) k_wf_shade(const scn_wf_scene S, const WfPaths W, int32_t bounce, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n || !W.alive[i]) return;
    if (!W.ok[i]) { wf_finish(W, i, 0.f); return; }
    float thr = W.throughput[i] * W.weight[i];
    if (W.exited[i]) { wf_finish(W, i, thr * S.emission); return; }
    const uint32_t x = W.px[i], y = W.py[i], spp = W.spp[i];
    thr *= S.albedo;
    // Russian roulette uses the bounce count before it is advanced (PathTracer.cpp:146-161)
    if (bounce > 2 && thr < 0.1f) {
        if (wf_rand(x, y, spp, (uint32_t)bounce, 3u, S.scene_seed) < thr) thr = 1.0f; // throughput /= roulettePdf
        else { wf_finish(W, i, 0.f); return; }
    }
    if (bounce + 1 >= S.max_bounces || !(thr > 0.f)) { wf_finish(W, i, 0.f); return; }
    // scatter point and shading frame (math/TangentFrame.hpp:74-99 about the normalised gradient)
    const float t = W.t[i];
    const f3 p = F3(W.ox[i] + t * W.dx[i], W.oy[i] + t * W.dy[i], W.oz[i] + t * W.dz[i]);
    const Frame fr = make_frame(F3(W.gx[i], W.gy[i], W.gz[i]));
    // cosine-weighted hemisphere sample (sampling/SampleWarp.hpp cosineHemisphere): wo.z >= 0 is the side of the normal
    const float u1 = wf_rand(x, y, spp, (uint32_t)bounce, 1u, S.scene_seed), u2 = wf_rand(x, y, spp, (uint32_t)bounce, 2u, S.scene_seed);
    const float phi = 6.283185307179586f * u1, r = sqrtf(u2);
    const f3 wo = F3(cosf(phi) * r, sinf(phi) * r, sqrtf(fmaxf(1.0f - u2, 0.0f)));
    f3 d = to_global(fr, wo);
    const float il = rsqrtf(d.x * d.x + d.y * d.y + d.z * d.z); d = F3(d.x * il, d.y * il, d.z * il);
    float t0, t1;
    if (!wf_cube(S, p.x, p.y, p.z, d.x, d.y, d.z, t0, t1)) { wf_finish(W, i, thr * S.emission); return; } // numerically outside: straight to the environment
    W.ox[i] = p.x; W.oy[i] = p.y; W.oz[i] = p.z; W.dx[i] = d.x; W.dy[i] = d.y; W.dz[i] = d.z; W.tmin[i] = 0.f; W.tmax[i] = t1;
    W.u[i] = wf_rand(x, y, spp, (uint32_t)bounce + 1u, 0u, S.scene_seed);
    W.seg[i] = (uint32_t)bounce + 1u; // state.info.pixelSampleSegment.w() = bounce (PathTracer.cpp:64)
    W.throughput[i] = thr;
}

__global__ void __launch_bounds__(WF_BLOCK) k_wf_accumulate(const float *__restrict__ radiance, float *__restrict__ fb, size_t n_pixels, uint32_t spp_count) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n_pixels) return;
    float s = 0.f;
    for (uint32_t k = 0; k < spp_count; ++k) s += radiance[i * spp_count + k]; // fixed order: deterministic
    fb[i] += s;
}

static int wf_paths(const scn_wf_paths *p, WfPaths &W, bool need_samples) {
    const scn_rays &r = p->rays; const scn_state &s = p->state; const scn_samples &o = p->samples;
    if (!r.ox || !r.oy || !r.oz || !r.dx || !r.dy || !r.dz || !r.tmin || !r.tmax || !r.u || !r.px || !r.py || !r.spp || !r.segment)
        return set_error(SCN_EINVAL, "wavefront: every ray array (including px, py, spp, segment) is required");
    if (!s.first_scatter || !s.last_val || !s.last_gx || !s.last_gy || !s.last_gz || !s.bounce || !s.last_gp_id || !s.info_t)
        return set_error(SCN_EINVAL, "wavefront: state arrays are required");
    if (!p->throughput || !p->radiance || !p->alive) return set_error(SCN_EINVAL, "wavefront: throughput, radiance and alive are required");
    if (need_samples && (!o.ok || !o.exited || !o.t || !o.gx || !o.gy || !o.gz || !o.weight)) return set_error(SCN_EINVAL, "wavefront: sample arrays are required");
    W.ox = (float *)r.ox; W.oy = (float *)r.oy; W.oz = (float *)r.oz; W.dx = (float *)r.dx; W.dy = (float *)r.dy; W.dz = (float *)r.dz;
    W.tmin = (float *)r.tmin; W.tmax = (float *)r.tmax; W.u = (float *)r.u;
    W.px = (uint32_t *)r.px; W.py = (uint32_t *)r.py; W.spp = (uint32_t *)r.spp; W.seg = (uint32_t *)r.segment;
    W.first_scatter = s.first_scatter; W.last_val = s.last_val; W.lgx = s.last_gx; W.lgy = s.last_gy; W.lgz = s.last_gz; W.bounce = s.bounce;
    W.last_gp_id = s.last_gp_id; W.info_t = s.info_t; W.scheme = s.scheme;
    bool all = true; for (int j = 0; j < 7; ++j) all = all && s.ctx[j];
    for (int j = 0; j < 7; ++j) W.ctx[j] = all ? s.ctx[j] : nullptr;
    W.ok = o.ok; W.exited = o.exited; W.t = o.t; W.gx = o.gx; W.gy = o.gy; W.gz = o.gz; W.weight = o.weight;
    W.throughput = p->throughput; W.radiance = p->radiance; W.alive = p->alive;
    return SCN_OK;
}
static int wf_check(const char *what) { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) return set_error(SCN_ECUDA, "%s: %s", what, cudaGetErrorString(e)); return SCN_OK; }

} // namespace scn

using namespace scn;
extern "C" {

uint32_t scn_wf_tile_count(uint32_t res_x, uint32_t res_y) { return ((res_x + SCN_WF_TILE - 1) / SCN_WF_TILE) * ((res_y + SCN_WF_TILE - 1) / SCN_WF_TILE); }

int scn_wf_raygen(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_count, uint32_t spp_first, uint32_t spp_count,
                  const scn_wf_paths *paths, size_t n, void *stream) {
    (void)m;
    if (!scene || !paths) return set_error(SCN_EINVAL, "null argument");
    if (spp_count == 0 || n != (size_t)tile_count * SCN_WF_TILE * SCN_WF_TILE * spp_count) return set_error(SCN_EINVAL, "raygen: n must be tile_count * 256 * spp_count");
    if (tile_first + tile_count > scn_wf_tile_count(scene->res_x, scene->res_y)) return set_error(SCN_EINVAL, "raygen: tile range outside the image");
    if (n == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, false)) return rc;
    k_wf_raygen<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*scene, W, tile_first, spp_first, spp_count, n);
    return wf_check("k_wf_raygen");
}

int scn_wf_shade(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, size_t n, int32_t bounce, uint32_t spp_first, uint32_t spp_count,
                 uint32_t tile_first, void *stream) {
    (void)m; (void)spp_first; (void)spp_count; (void)tile_first;
    if (!scene || !paths) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, true)) return rc;
    k_wf_shade<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*scene, W, bounce, n);
    return wf_check("k_wf_shade");
}

int scn_wf_accumulate(const float *radiance, float *framebuffer, size_t n_pixels, uint32_t spp_count, void *stream) {
    if (!radiance || !framebuffer) return set_error(SCN_EINVAL, "null argument");
    if (n_pixels == 0 || spp_count == 0) return SCN_OK;
    k_wf_accumulate<<<(unsigned)((n_pixels + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(radiance, framebuffer, n_pixels, spp_count);
    return wf_check("k_wf_accumulate");
}

} // extern "C"
