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
    const uint8_t *ok, *exited, *sample_scheme; const float *t, *gx, *gy, *gz, *weight;
    float *throughput, *radiance; uint8_t *alive;
    float *thr_gb[2], *rad_gb[2]; uint8_t *scattered; // nee stage: green / blue planes, "has scattered" flag
};

__device__ __forceinline__ void wf_finish(const WfPaths &W, size_t i, float radiance) {
    W.radiance[i] = radiance; W.alive[i] = 0; W.bounce[i] = WF_DEAD_BOUNCE;
}
// nee stage: radiance accumulates along the path (emission += throughput * ..., PathTracer.cpp:72, TraceBase.cpp:546)
__device__ __forceinline__ void wf_finish_rgb(const WfPaths &W, size_t i, f3 add) {
    W.radiance[i] += add.x; W.rad_gb[0][i] += add.y; W.rad_gb[1][i] += add.z; W.alive[i] = 0; W.bounce[i] = WF_DEAD_BOUNCE;
}

// ---- emitters of the nee stage ------------------------------------------------------------------------------------------------
struct WfQuad { f3 base, e0, e1, n; float area, inv_uv0, inv_uv1; };
// Quad::prepareForRender (primitives/Quad.cpp:296-316)
__device__ __forceinline__ WfQuad wf_quad(const scn_wf_scene &S) {
    WfQuad q; q.base = F3(S.quad_base[0], S.quad_base[1], S.quad_base[2]); q.e0 = F3(S.quad_edge0[0], S.quad_edge0[1], S.quad_edge0[2]); q.e1 = F3(S.quad_edge1[0], S.quad_edge1[1], S.quad_edge1[2]);
    f3 n = F3(q.e1.y * q.e0.z - q.e1.z * q.e0.y, q.e1.z * q.e0.x - q.e1.x * q.e0.z, q.e1.x * q.e0.y - q.e1.y * q.e0.x);
    q.area = sqrtf(n.x * n.x + n.y * n.y + n.z * n.z); q.n = F3(n.x / q.area, n.y / q.area, n.z / q.area);
    q.inv_uv0 = 1.0f / (q.e0.x * q.e0.x + q.e0.y * q.e0.y + q.e0.z * q.e0.z); q.inv_uv1 = 1.0f / (q.e1.x * q.e1.x + q.e1.y * q.e1.y + q.e1.z * q.e1.z);
    return q;
}
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return F3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) { return F3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ f3 norm3(f3 a) { const float il = 1.0f / sqrtf(dot3(a, a)); return F3(a.x * il, a.y * il, a.z * il); }
// Quad::intersect (primitives/Quad.cpp:71-97) over [0, inf): distance and side
__device__ __forceinline__ bool wf_quad_hit(const WfQuad &q, f3 o, f3 d, float &t, bool &back) {
    const float ndw = dot3(d, q.n);
    if (fabsf(ndw) < 1e-6f) return false;
    t = dot3(q.n, sub3(q.base, o)) / ndw;
    if (t < 0.f) return false;
    const f3 v = sub3(F3(o.x + t * d.x, o.y + t * d.y, o.z + t * d.z), q.base);
    const float l0 = dot3(v, q.e0) * q.inv_uv0, l1 = dot3(v, q.e1) * q.inv_uv1;
    if (l0 < 0.f || l0 > 1.f || l1 < 0.f || l1 > 1.f) return false;
    back = ndw >= 0.f;
    return true;
}
// Quad::approximateRadiance (primitives/Quad.cpp:256-279): solid angle of the spherical quadrilateral x emission
__device__ __forceinline__ float wf_quad_approx(const WfQuad &q, f3 p, float emax) {
    const f3 R0 = sub3(q.base, p);
    if (dot3(R0, q.n) >= 0.f) return 0.f;
    const f3 R1 = F3(R0.x + q.e0.x, R0.y + q.e0.y, R0.z + q.e0.z), R2 = F3(R1.x + q.e1.x, R1.y + q.e1.y, R1.z + q.e1.z), R3 = F3(R0.x + q.e1.x, R0.y + q.e1.y, R0.z + q.e1.z);
    const f3 n0 = norm3(cross3(R0, R1)), n1 = norm3(cross3(R1, R2)), n2 = norm3(cross3(R2, R3)), n3 = norm3(cross3(R3, R0));
    const float Q = acosf(dot3(n0, n1)) + acosf(dot3(n1, n2)) + acosf(dot3(n2, n3)) + acosf(dot3(n3, n0));
    return (6.283185307179586f - fabsf(Q)) * emax;
}
// Fresnel::conductorReflectance (bsdfs/Fresnel.hpp:102-123)
__device__ __forceinline__ float wf_conductor(float eta, float k, float c) {
    if (eta == 0.f && k == 0.f) return 1.f;
    const float c2 = c * c, s2 = fmaxf(1.0f - c2, 0.0f), s4 = s2 * s2;
    const float inner = eta * eta - k * k - s2;
    const float ab = sqrtf(fmaxf(inner * inner + 4.0f * eta * eta * k * k, 0.0f));
    const float a = sqrtf(fmaxf((ab + inner) * 0.5f, 0.0f));
    const float Rs = ((ab + c2) - (2.0f * a * c)) / ((ab + c2) + (2.0f * a * c));
    const float Rp = ((c2 * ab + s4) - (2.0f * a * c * s2)) / ((c2 * ab + s4) + (2.0f * a * c * s2));
    return 0.5f * (Rs + Rs * Rp);
}
// albedo x Fresnel of the specular phase bsdfs (bsdfs/MirrorBsdf.cpp:38, bsdfs/ConductorBsdf.cpp:66)
__device__ __forceinline__ f3 wf_specular_weight(const scn_wf_scene &S, float cos_i) {
    f3 w = F3(S.albedo_rgb[0], S.albedo_rgb[1], S.albedo_rgb[2]);
    if (S.phase_type == SCN_WF_PHASE_CONDUCTOR) { w.x *= wf_conductor(S.eta[0], S.k[0], cos_i); w.y *= wf_conductor(S.eta[1], S.k[1], cos_i); w.z *= wf_conductor(S.eta[2], S.k[2], cos_i); }
    return w;
}
__device__ __forceinline__ float power_heuristic(float a, float b) { return (a * a) / (a * a + b * b); }

// PathTracer::traceSample :31-59 — camera sample, first intersection with the scene (the cube), medium state reset
__global__ void __launch_bounds__(WF_BLOCK) k_wf_raygen(const scn_wf_scene S, const WfPaths W, uint32_t tile_first, uint32_t tile_stride, uint32_t spp_first, uint32_t spp_count, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = (uint32_t)(i % spp_count); const size_t lp = i / spp_count;
    const uint32_t tiles_x = (S.res_x + SCN_WF_TILE - 1) / SCN_WF_TILE;
    const uint32_t tile = tile_first + tile_stride * (uint32_t)(lp / (SCN_WF_TILE * SCN_WF_TILE)), in_tile = (uint32_t)(lp % (SCN_WF_TILE * SCN_WF_TILE));
    const uint32_t x = (tile % tiles_x) * SCN_WF_TILE + in_tile % SCN_WF_TILE, y = (tile / tiles_x) * SCN_WF_TILE + in_tile / SCN_WF_TILE;
    const uint32_t spp = spp_first + s;
    // state.reset() (media/Medium.hpp:70-80) and RayInfo
    W.px[i] = x; W.py[i] = y; W.spp[i] = spp; W.seg[i] = 0u;
    W.first_scatter[i] = 1; W.last_val[i] = 0.f; W.lgx[i] = 0.f; W.lgy[i] = 0.f; W.lgz[i] = 0.f; W.bounce[i] = 0; W.last_gp_id[i] = 0; W.info_t[i] = 0.f;
    if (W.scheme) W.scheme[i] = SCN_SCHEME_UNI;
    if (W.ctx[0]) for (int j = 0; j < 7; ++j) W.ctx[j][i] = 0.f;
    W.throughput[i] = 1.f; W.radiance[i] = 0.f; W.alive[i] = 1;
    if (S.nee) { W.thr_gb[0][i] = 1.f; W.thr_gb[1][i] = 1.f; W.rad_gb[0][i] = 0.f; W.rad_gb[1][i] = 0.f; W.scattered[i] = 0; }
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
    const bool hit_cube = wf_cube(S, S.cam_pos[0], S.cam_pos[1], S.cam_pos[2], d.x, d.y, d.z, t0, t1);
    if (S.nee) { // an unscattered path sees the emitters directly (wasSpecular starts true, PathTracer.cpp:58)
        float tq; bool back;
        const bool hit_quad = S.has_quad && wf_quad_hit(wf_quad(S), F3(S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]), d, tq, back);
        if (hit_quad && (!hit_cube || tq < t0)) { wf_finish_rgb(W, i, back ? F3(0, 0, 0) : F3(S.quad_rgb[0], S.quad_rgb[1], S.quad_rgb[2])); return; }
        if (!hit_cube) { wf_finish_rgb(W, i, F3(S.env_rgb[0], S.env_rgb[1], S.env_rgb[2])); return; }
    } else if (!hit_cube) { wf_finish(W, i, S.emission); return; } // straight to the environment
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

// ---- nee stage ----------------------------------------------------------------------------------------------------------------
struct WfShadow {
    float *ox, *oy, *oz, *dx, *dy, *dz, *tmin, *tmax, *u; uint32_t *px, *py, *spp, *seg;
    uint8_t *first_scatter; float *last_val, *lgx, *lgy, *lgz; int32_t *bounce, *last_gp_id; float *info_t;
    float *T; float *contrib[3];
};
__device__ __forceinline__ void wf_shadow_clear(const WfShadow &H, const WfPaths &W, size_t i) {
    // a ray of zero length: the march returns T = 1 without a single evaluation; contrib = 0 makes it irrelevant
    H.ox[i] = 0.f; H.oy[i] = 0.f; H.oz[i] = 0.f; H.dx[i] = 0.f; H.dy[i] = 0.f; H.dz[i] = 1.f; H.tmin[i] = 0.f; H.tmax[i] = 0.f; H.u[i] = 0.f;
    H.px[i] = W.px[i]; H.py[i] = W.py[i]; H.spp[i] = W.spp[i]; H.seg[i] = 0u;
    H.first_scatter[i] = 1; H.last_val[i] = 0.f; H.lgx[i] = 0.f; H.lgy[i] = 0.f; H.lgz[i] = 0.f; H.bounce[i] = 0; H.last_gp_id[i] = 0; H.info_t[i] = 0.f;
    H.contrib[0][i] = 0.f; H.contrib[1][i] = 0.f; H.contrib[2][i] = 0.f;
}
// the shadow ray p + t d, t in [0, far], through the medium inside the cube; MediumState copy with the segment index advanced
// (TraceBase.cpp:543-546) and lastAniso = `aniso` (the sampled gradient, or neeGrad for a specular light sample: :372-376)
__device__ __forceinline__ void wf_shadow_write(const scn_wf_scene &S, const WfShadow &H, const WfPaths &W, size_t i, f3 p, f3 d, float far_t, f3 aniso, float u, f3 contrib) {
    float t0, t1;
    const bool inside = wf_cube(S, p.x, p.y, p.z, d.x, d.y, d.z, t0, t1);
    H.ox[i] = p.x; H.oy[i] = p.y; H.oz[i] = p.z; H.dx[i] = d.x; H.dy[i] = d.y; H.dz[i] = d.z; H.tmin[i] = 0.f; H.tmax[i] = inside ? fminf(t1, far_t) : 0.f; H.u[i] = u;
    H.seg[i] = W.seg[i] + 1u;
    H.first_scatter[i] = W.first_scatter[i]; H.last_val[i] = W.last_val[i]; H.lgx[i] = aniso.x; H.lgy[i] = aniso.y; H.lgz[i] = aniso.z;
    H.bounce[i] = W.bounce[i]; H.last_gp_id[i] = W.last_gp_id[i]; H.info_t[i] = W.info_t[i];
    H.contrib[0][i] = contrib.x; H.contrib[1][i] = contrib.y; H.contrib[2][i] = contrib.z;
}

// volumeEstimateDirect (TraceBase.cpp:506-518) for the paths that scattered at `bounce`: chooseLight (:457-504), volumeLightSample
// (:346-388) into L, volumePhaseSample (:390-420) into Q.  The transmittance along both rays is computed by the caller.
__global__ void __launch_bounds__(WF_BLOCK) k_wf_nee_setup(const __grid_constant__ Params P, const scn_wf_scene S, const WfPaths W, const WfShadow L, const WfShadow Q, int32_t bounce, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n) return;
    wf_shadow_clear(L, W, i); wf_shadow_clear(Q, W, i);
    if (!W.alive[i] || !W.ok[i] || W.exited[i]) return;
    if (!(bounce < S.max_bounces - 1)) return; // TraceBase.cpp:546
    const uint32_t x = W.px[i], y = W.py[i], spp = W.spp[i], nb = (uint32_t)bounce | 0x8000u;
    const float wq = W.weight[i];
    const f3 thr = F3(W.throughput[i] * wq, W.thr_gb[0][i] * wq, W.thr_gb[1][i] * wq);
    const float t = W.t[i];
    const f3 dir = F3(W.dx[i], W.dy[i], W.dz[i]);
    const f3 p = F3(W.ox[i] + t * dir.x, W.oy[i] + t * dir.y, W.oz[i] + t * dir.z);
    const f3 g = F3(W.gx[i], W.gy[i], W.gz[i]);
    // BRDFPhaseFunction::setEventIsectInfo (phasefunctions/BRDFPhaseFunction.cpp:27-44)
    const Frame fr = make_frame(norm3(g));
    const f3 nrm = fr.n;
    const f3 wi = norm3(to_local(fr, F3(-dir.x, -dir.y, -dir.z)));
    const int sch = (int)W.sample_scheme[i];
    const bool specular = S.phase_type != SCN_WF_PHASE_LAMBERT;
    const Info info = {x, y, spp, W.seg[i], S.scene_seed, W.info_t[i]};
    Coeff c; c.vs = 0.f; c.gs = F3(0, 0, 0); c.org = F3(0, 0, 0);
    if (W.ctx[0]) { c.vs = W.ctx[0][i]; c.gs = F3(W.ctx[1][i], W.ctx[2][i], W.ctx[3][i]); c.org = F3(W.ctx[4][i], W.ctx[5][i], W.ctx[6][i]); }
    // ---- chooseLight: lights in primitive order, pdf ~ approximateRadiance
    const WfQuad quad = wf_quad(S);
    const float env_max = fmaxf(S.env_rgb[0], fmaxf(S.env_rgb[1], S.env_rgb[2])), quad_max = fmaxf(S.quad_rgb[0], fmaxf(S.quad_rgb[1], S.quad_rgb[2]));
    const bool env_light = S.env_sample && env_max > 0.f, quad_light = S.has_quad && quad_max > 0.f;
    bool use_quad; float lw;
    if (env_light && quad_light) {
        const float pe = 6.283185307179586f * env_max, pq = wf_quad_approx(quad, p, quad_max), total = pe + pq;
        if (total == 0.f) return;
        const float tsel = wf_rand(x, y, spp, nb, 0u, S.scene_seed) * total;
        const float first = S.quad_first ? pq : pe;
        use_quad = (tsel < first) ? (S.quad_first != 0) : (S.quad_first == 0);
        lw = total / (use_quad ? pq : pe);
    } else if (env_light || quad_light) { use_quad = quad_light; lw = 1.f; }
    else return;
    const f3 Le = use_quad ? F3(S.quad_rgb[0], S.quad_rgb[1], S.quad_rgb[2]) : F3(S.env_rgb[0], S.env_rgb[1], S.env_rgb[2]);
    const f3 aniso = F3(W.lgx[i], W.lgy[i], W.lgz[i]); // state.lastAniso after the query = the sampled gradient

    // ---- volumeLightSample
    if (!(specular && sch == SCN_SCHEME_UNI)) {
        const float u1 = wf_rand(x, y, spp, nb, 1u, S.scene_seed), u2 = wf_rand(x, y, spp, nb, 2u, S.scene_seed);
        bool okl = true; f3 d = F3(0, 0, 1); float dist = INFINITY, pdf_l = 0.f;
        if (use_quad) { // Quad::sampleDirect (primitives/Quad.cpp:172-187)
            if (dot3(quad.n, sub3(p, quad.base)) <= 0.f) okl = false;
            else {
                const f3 qp = F3(quad.base.x + u1 * quad.e0.x + u2 * quad.e1.x, quad.base.y + u1 * quad.e0.y + u2 * quad.e1.y, quad.base.z + u1 * quad.e0.z + u2 * quad.e1.z);
                d = sub3(qp, p); const float r2 = dot3(d, d); dist = sqrtf(r2); d = F3(d.x / dist, d.y / dist, d.z / dist);
                pdf_l = r2 / (-dot3(quad.n, d) * quad.area);
            }
        } else { // InfiniteSphere::sampleDirect, constant emission (primitives/InfiniteSphere.cpp:161-167)
            const float phi = u1 * 6.283185307179586f, z = u2 * 2.0f - 1.0f, r = sqrtf(fmaxf(1.0f - z * z, 0.0f));
            d = F3(cosf(phi) * r, sinf(phi) * r, z); pdf_l = 0.07957747154594767f;
        }
        if (okl) {
            const f3 wo = norm3(to_local(fr, d));
            f3 f = F3(0, 0, 0); float phase_pdf = 0.f; f3 an = aniso;
            if (!specular) { // LambertBsdf::eval / pdf (bsdfs/LambertBsdf.cpp:40-47, 62-69)
                if (wi.z > 0.f && wo.z > 0.f) { const float k = 0.3183098861837907f * wo.z; f = F3(S.albedo_rgb[0] * k, S.albedo_rgb[1] * k, S.albedo_rgb[2] * k); phase_pdf = fabsf(wo.z) * 0.3183098861837907f; }
            } else { // MirrorBsdf / ConductorBsdf::eval, pdf, evalGrad with a sampling scheme other than UNI
                const f3 nl = F3((wi.x + wo.x) * 0.5f, (wi.y + wo.y) * 0.5f, (wi.z + wo.z) * 0.5f);
                const f3 nw = norm3(to_global(fr, nl));
                const float pn = nee_pdf(P, c, dir, nw, p, t, info);
                const f3 w = wf_specular_weight(S, wi.z);
                f = F3(w.x * pn, w.y * pn, w.z * pn); phase_pdf = pn;
                Nee o; nee_shared(P, c, dir, nw, p, info, o); an = F3(o.dist * nw.x, o.dist * nw.y, o.dist * nw.z);
            }
            if (!(f.x == 0.f && f.y == 0.f && f.z == 0.f)) {
                // attenuatedEmission (TraceBase.cpp:156-185): the light must be hit at the expected distance; evalDirect
                bool vis = true; float far_t = INFINITY; f3 e = Le;
                float tq; bool back;
                const bool hq = S.has_quad && wf_quad_hit(quad, p, d, tq, back);
                if (use_quad) { if (!hq || tq * (1.0f + 1e-3f) < dist) vis = false; else { far_t = tq; if (back) e = F3(0, 0, 0); } }
                else if (hq) vis = false; // the quad (null bsdf: no forward lobe) blocks the way to the environment (TraceBase.cpp:92-94)
                if (vis && !(e.x == 0.f && e.y == 0.f && e.z == 0.f)) {
                    float mis = 1.f;
                    if (!(specular && sch == SCN_SCHEME_NEE)) mis = power_heuristic(pdf_l, phase_pdf);
                    const float k = lw * mis / pdf_l;
                    wf_shadow_write(S, L, W, i, p, d, far_t, an, wf_rand(x, y, spp, nb, 3u, S.scene_seed), F3(thr.x * f.x * e.x * k, thr.y * f.y * e.y * k, thr.z * f.z * e.z * k));
                }
            }
        }
    }
    // ---- volumePhaseSample
    if (!(specular && sch == SCN_SCHEME_NEE)) {
        f3 wo, w; float pdf_p; bool oks = true;
        if (!specular) { // LambertBsdf::sample (bsdfs/LambertBsdf.cpp:26-38)
            if (wi.z <= 0.f) oks = false;
            const float u1 = wf_rand(x, y, spp, nb, 5u, S.scene_seed), u2 = wf_rand(x, y, spp, nb, 6u, S.scene_seed);
            const float phi = 6.283185307179586f * u1, r = sqrtf(u2);
            wo = F3(cosf(phi) * r, sinf(phi) * r, sqrtf(fmaxf(1.0f - u2, 0.0f)));
            w = F3(S.albedo_rgb[0], S.albedo_rgb[1], S.albedo_rgb[2]); pdf_p = fabsf(wo.z) * 0.3183098861837907f;
        } else { // MirrorBsdf / ConductorBsdf::sample
            wo = F3(-wi.x, -wi.y, wi.z); w = wf_specular_weight(S, wi.z); pdf_p = 1.f;
            if (sch != SCN_SCHEME_UNI) pdf_p = nee_pdf(P, c, dir, nrm, p, t, info);
        }
        if (oks) {
            const f3 d = norm3(to_global(fr, wo));
            bool vis = true; float far_t = INFINITY, dpdf = 0.07957747154594767f; f3 e = Le;
            float tq; bool back;
            const bool hq = S.has_quad && wf_quad_hit(quad, p, d, tq, back);
            if (use_quad) {
                if (!hq) vis = false;
                else { far_t = tq; if (back) e = F3(0, 0, 0); const float ndw = dot3(quad.n, d), tt = dot3(quad.n, sub3(quad.base, p)) / ndw; dpdf = tt * tt / (fabsf(ndw) * quad.area); } // Quad::directPdf (:216-223)
            } else if (hq) vis = false;
            if (vis && !(e.x == 0.f && e.y == 0.f && e.z == 0.f)) {
                float mis = 1.f;
                if (!specular || sch != SCN_SCHEME_UNI) mis = power_heuristic(pdf_p, dpdf);
                const float k = lw * mis;
                wf_shadow_write(S, Q, W, i, p, d, far_t, aniso, wf_rand(x, y, spp, nb, 4u, S.scene_seed), F3(thr.x * w.x * e.x * k, thr.y * w.y * e.y * k, thr.z * w.z * e.z * k));
            }
        }
    }
}

__global__ void __launch_bounds__(WF_BLOCK) k_wf_nee_gather(const WfPaths W, const WfShadow L, const WfShadow Q, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n) return;
    const float tl = L.T[i], tq = Q.T[i];
    W.radiance[i] += L.contrib[0][i] * tl + Q.contrib[0][i] * tq;
    W.rad_gb[0][i] += L.contrib[1][i] * tl + Q.contrib[1][i] * tq;
    W.rad_gb[1][i] += L.contrib[2][i] * tl + Q.contrib[2][i] * tq;
}

// k_wf_shade with the reference's bookkeeping of direct emitter hits: an emitter met by the path itself counts only while the path
// has not scattered (TraceBase.cpp:592-596, 629-637 with wasSpecular = false after handleVolume); RGB throughput.
// GATHER: k_wf_nee_gather folded in (scn_wf_render): the shadow rays of this bounce were written for live paths only, so the
// gather of a dead path would add 0 and is skipped with it.
template <bool GATHER>
__global__ void __launch_bounds__(WF_BLOCK) k_wf_shade_nee(const scn_wf_scene S, const WfPaths W, const WfShadow L, const WfShadow Q, int32_t bounce, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n || !W.alive[i]) return;
    if (GATHER) {
        const float tl = L.T[i], tq = Q.T[i];
        W.radiance[i] += L.contrib[0][i] * tl + Q.contrib[0][i] * tq;
        W.rad_gb[0][i] += L.contrib[1][i] * tl + Q.contrib[1][i] * tq;
        W.rad_gb[1][i] += L.contrib[2][i] * tl + Q.contrib[2][i] * tq;
    }
    if (!W.ok[i]) { wf_finish_rgb(W, i, F3(0, 0, 0)); return; }
    const float wq = W.weight[i];
    f3 thr = F3(W.throughput[i] * wq, W.thr_gb[0][i] * wq, W.thr_gb[1][i] * wq);
    const float t = W.t[i];
    const f3 dir = F3(W.dx[i], W.dy[i], W.dz[i]);
    const f3 p = F3(W.ox[i] + t * dir.x, W.oy[i] + t * dir.y, W.oz[i] + t * dir.z);
    if (W.exited[i]) {
        f3 e = F3(0, 0, 0);
        if (!W.scattered[i]) {
            float tq; bool back;
            if (S.has_quad && wf_quad_hit(wf_quad(S), p, dir, tq, back)) e = back ? F3(0, 0, 0) : F3(S.quad_rgb[0], S.quad_rgb[1], S.quad_rgb[2]);
            else e = F3(S.env_rgb[0], S.env_rgb[1], S.env_rgb[2]);
        }
        wf_finish_rgb(W, i, F3(thr.x * e.x, thr.y * e.y, thr.z * e.z)); return;
    }
    W.scattered[i] = 1;
    const uint32_t x = W.px[i], y = W.py[i], spp = W.spp[i];
    const Frame fr = make_frame(norm3(F3(W.gx[i], W.gy[i], W.gz[i])));
    const f3 wi = norm3(to_local(fr, F3(-dir.x, -dir.y, -dir.z)));
    f3 wo, w;
    if (S.phase_type == SCN_WF_PHASE_LAMBERT) {
        if (wi.z <= 0.f) { wf_finish_rgb(W, i, F3(0, 0, 0)); return; } // "Wrong normal direction in sampling phase function" (TraceBase.cpp:551-554)
        const float u1 = wf_rand(x, y, spp, (uint32_t)bounce, 1u, S.scene_seed), u2 = wf_rand(x, y, spp, (uint32_t)bounce, 2u, S.scene_seed);
        const float phi = 6.283185307179586f * u1, r = sqrtf(u2);
        wo = F3(cosf(phi) * r, sinf(phi) * r, sqrtf(fmaxf(1.0f - u2, 0.0f)));
        w = F3(S.albedo_rgb[0], S.albedo_rgb[1], S.albedo_rgb[2]);
    } else { wo = F3(-wi.x, -wi.y, wi.z); w = wf_specular_weight(S, wi.z); }
    thr = F3(thr.x * w.x, thr.y * w.y, thr.z * w.z);
    const float tmax = fmaxf(thr.x, fmaxf(thr.y, thr.z));
    if (tmax == 0.f) { wf_finish_rgb(W, i, F3(0, 0, 0)); return; }
    const float rp = fmaxf(fabsf(thr.x), fmaxf(fabsf(thr.y), fabsf(thr.z)));
    if (bounce > 2 && rp < 0.1f) {
        if (wf_rand(x, y, spp, (uint32_t)bounce, 3u, S.scene_seed) < rp) thr = F3(thr.x / rp, thr.y / rp, thr.z / rp);
        else { wf_finish_rgb(W, i, F3(0, 0, 0)); return; }
    }
    if (bounce + 1 >= S.max_bounces) { wf_finish_rgb(W, i, F3(0, 0, 0)); return; }
    const f3 d = norm3(to_global(fr, wo));
    float t0, t1;
    if (!wf_cube(S, p.x, p.y, p.z, d.x, d.y, d.z, t0, t1)) { wf_finish_rgb(W, i, F3(0, 0, 0)); return; }
    W.ox[i] = p.x; W.oy[i] = p.y; W.oz[i] = p.z; W.dx[i] = d.x; W.dy[i] = d.y; W.dz[i] = d.z; W.tmin[i] = 0.f; W.tmax[i] = t1;
    W.u[i] = wf_rand(x, y, spp, (uint32_t)bounce + 1u, 0u, S.scene_seed);
    W.seg[i] = (uint32_t)bounce + 1u;
    W.throughput[i] = thr.x; W.thr_gb[0][i] = thr.y; W.thr_gb[1][i] = thr.z;
}

// PathTraceIntegrator::renderTile (integrators/path_tracer/PathTraceIntegrator.cpp:148-156): a sample with an infinite component, a
// component above 65536 or a NaN is replaced by black before it reaches the framebuffer
__global__ void __launch_bounds__(WF_BLOCK) k_wf_clamp_rgb(float *__restrict__ r, float *__restrict__ g, float *__restrict__ b, size_t n) {
    const size_t i = (size_t)blockIdx.x * WF_BLOCK + threadIdx.x;
    if (i >= n) return;
    const float x = r[i], y = g[i], z = b[i];
    const bool bad = isinf(x) || isinf(y) || isinf(z) || x > 65536.f || y > 65536.f || z > 65536.f || isnan(x) || isnan(y) || isnan(z);
    if (bad) { r[i] = 0.f; g[i] = 0.f; b[i] = 0.f; }
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
    W.ok = o.ok; W.exited = o.exited; W.sample_scheme = o.scheme; W.t = o.t; W.gx = o.gx; W.gy = o.gy; W.gz = o.gz; W.weight = o.weight;
    W.throughput = p->throughput; W.radiance = p->radiance; W.alive = p->alive;
    W.thr_gb[0] = p->throughput_gb[0]; W.thr_gb[1] = p->throughput_gb[1]; W.rad_gb[0] = p->radiance_gb[0]; W.rad_gb[1] = p->radiance_gb[1]; W.scattered = p->scattered;
    return SCN_OK;
}
static int wf_nee_arrays(const scn_wf_paths *p) {
    if (!p->throughput_gb[0] || !p->throughput_gb[1] || !p->radiance_gb[0] || !p->radiance_gb[1] || !p->scattered)
        return set_error(SCN_EINVAL, "wavefront: scene.nee needs throughput_gb, radiance_gb and scattered");
    return SCN_OK;
}
static int wf_shadow(const scn_wf_shadow *h, WfShadow &H) {
    const scn_rays &r = h->rays; const scn_state &s = h->state;
    if (!r.ox || !r.oy || !r.oz || !r.dx || !r.dy || !r.dz || !r.tmin || !r.tmax || !r.u || !r.px || !r.py || !r.spp || !r.segment)
        return set_error(SCN_EINVAL, "shadow rays: every ray array is required");
    if (!s.first_scatter || !s.last_val || !s.last_gx || !s.last_gy || !s.last_gz || !s.bounce || !s.last_gp_id || !s.info_t)
        return set_error(SCN_EINVAL, "shadow rays: state arrays are required");
    if (!h->T || !h->contrib[0] || !h->contrib[1] || !h->contrib[2]) return set_error(SCN_EINVAL, "shadow rays: T and contrib are required");
    H.ox = (float *)r.ox; H.oy = (float *)r.oy; H.oz = (float *)r.oz; H.dx = (float *)r.dx; H.dy = (float *)r.dy; H.dz = (float *)r.dz;
    H.tmin = (float *)r.tmin; H.tmax = (float *)r.tmax; H.u = (float *)r.u;
    H.px = (uint32_t *)r.px; H.py = (uint32_t *)r.py; H.spp = (uint32_t *)r.spp; H.seg = (uint32_t *)r.segment;
    H.first_scatter = s.first_scatter; H.last_val = s.last_val; H.lgx = s.last_gx; H.lgy = s.last_gy; H.lgz = s.last_gz; H.bounce = s.bounce;
    H.last_gp_id = s.last_gp_id; H.info_t = s.info_t;
    H.T = h->T; H.contrib[0] = h->contrib[0]; H.contrib[1] = h->contrib[1]; H.contrib[2] = h->contrib[2];
    return SCN_OK;
}
static int wf_check(const char *what) { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) return set_error(SCN_ECUDA, "%s: %s", what, cudaGetErrorString(e)); return SCN_OK; }

} // namespace scn

using namespace scn;
extern "C" const void *scn_internal_params(const scn_medium *m); // scn_api.cu: the medium's kernel parameters (scn::Params)
extern "C" {

uint32_t scn_wf_tile_count(uint32_t res_x, uint32_t res_y) { return ((res_x + SCN_WF_TILE - 1) / SCN_WF_TILE) * ((res_y + SCN_WF_TILE - 1) / SCN_WF_TILE); }

int scn_wf_raygen_strided(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_stride, uint32_t tile_count, uint32_t spp_first,
                          uint32_t spp_count, const scn_wf_paths *paths, size_t n, void *stream) {
    (void)m;
    if (!scene || !paths) return set_error(SCN_EINVAL, "null argument");
    if (spp_count == 0 || n != (size_t)tile_count * SCN_WF_TILE * SCN_WF_TILE * spp_count) return set_error(SCN_EINVAL, "raygen: n must be tile_count * 256 * spp_count");
    if (tile_stride == 0) return set_error(SCN_EINVAL, "raygen: tile_stride must be >= 1");
    if (tile_count && (uint64_t)tile_first + (uint64_t)tile_stride * (tile_count - 1) >= scn_wf_tile_count(scene->res_x, scene->res_y))
        return set_error(SCN_EINVAL, "raygen: tile range outside the image");
    if (n == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, false)) return rc;
    if (scene->nee) if (int rc = wf_nee_arrays(paths)) return rc;
    k_wf_raygen<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*scene, W, tile_first, tile_stride, spp_first, spp_count, n);
    return wf_check("k_wf_raygen");
}

int scn_wf_raygen(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_count, uint32_t spp_first, uint32_t spp_count,
                  const scn_wf_paths *paths, size_t n, void *stream) {
    return scn_wf_raygen_strided(scene, m, tile_first, 1u, tile_count, spp_first, spp_count, paths, n, stream);
}

int scn_wf_shade(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, size_t n, int32_t bounce, uint32_t spp_first, uint32_t spp_count,
                 uint32_t tile_first, void *stream) {
    (void)m; (void)spp_first; (void)spp_count; (void)tile_first;
    if (!scene || !paths) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, true)) return rc;
    if (scene->nee) {
        if (int rc = wf_nee_arrays(paths)) return rc;
        WfShadow none{};
        k_wf_shade_nee<false><<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*scene, W, none, none, bounce, n);
        return wf_check("k_wf_shade_nee");
    }
    k_wf_shade<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*scene, W, bounce, n);
    return wf_check("k_wf_shade");
}

int scn_wf_nee_setup(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, const scn_wf_shadow *light_rays, const scn_wf_shadow *phase_rays,
                     size_t n, int32_t bounce, void *stream) {
    if (!scene || !m || !paths || !light_rays || !phase_rays) return set_error(SCN_EINVAL, "null argument");
    if (!scene->nee) return set_error(SCN_EINVAL, "scn_wf_nee_setup: scene.nee is 0");
    if (n == 0) return SCN_OK;
    const Params *P = (const Params *)scn_internal_params(m);
    if (scene->phase_type != SCN_WF_PHASE_LAMBERT && !P->sampling_1d)
        return set_error(SCN_EUNSUPPORTED, "mirror / conductor phase functions need a 1D-sampling medium (neePDF / neeGrad are 1D-noise entry points)");
    WfPaths W; if (int rc = wf_paths(paths, W, true)) return rc;
    if (int rc = wf_nee_arrays(paths)) return rc;
    if (!paths->samples.scheme) return set_error(SCN_EINVAL, "scn_wf_nee_setup: samples.scheme is required");
    WfShadow L, Q; if (int rc = wf_shadow(light_rays, L)) return rc; if (int rc = wf_shadow(phase_rays, Q)) return rc;
    k_wf_nee_setup<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(*P, *scene, W, L, Q, bounce, n);
    return wf_check("k_wf_nee_setup");
}

int scn_wf_clamp_samples(const scn_wf_paths *paths, size_t n, void *stream) {
    if (!paths) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    if (!paths->radiance) return set_error(SCN_EINVAL, "wavefront: radiance is required");
    if (int rc = wf_nee_arrays(paths)) return rc;
    k_wf_clamp_rgb<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(paths->radiance, paths->radiance_gb[0], paths->radiance_gb[1], n);
    return wf_check("k_wf_clamp_rgb");
}

int scn_wf_nee_gather(const scn_wf_paths *paths, const scn_wf_shadow *light_rays, const scn_wf_shadow *phase_rays, size_t n, void *stream) {
    if (!paths || !light_rays || !phase_rays) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, false)) return rc;
    if (int rc = wf_nee_arrays(paths)) return rc;
    WfShadow L, Q; if (int rc = wf_shadow(light_rays, L)) return rc; if (int rc = wf_shadow(phase_rays, Q)) return rc;
    k_wf_nee_gather<<<(unsigned)((n + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(W, L, Q, n);
    return wf_check("k_wf_nee_gather");
}

int scn_wf_accumulate(const float *radiance, float *framebuffer, size_t n_pixels, uint32_t spp_count, void *stream) {
    if (!radiance || !framebuffer) return set_error(SCN_EINVAL, "null argument");
    if (n_pixels == 0 || spp_count == 0) return SCN_OK;
    k_wf_accumulate<<<(unsigned)((n_pixels + WF_BLOCK - 1) / WF_BLOCK), WF_BLOCK, 0, (cudaStream_t)stream>>>(radiance, framebuffer, n_pixels, spp_count);
    return wf_check("k_wf_accumulate");
}

// The whole per-rank render loop in one call (what render.py used to drive launch by launch from Python): for every chunk of
// `spp_chunk` samples per pixel  raygen -> max_bounces x { medium query, [nee setup, two transmittance wavefronts], shade (+ gather) }
// -> clamp -> accumulate, all enqueued on `stream` with no host synchronisation.  PathTraceIntegrator::renderTile
// (integrators/path_tracer/PathTraceIntegrator.cpp:135-163) over PathTracer::traceSample (PathTracer.cpp:62-169) for this GPU's tiles.
int scn_wf_render(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_stride, uint32_t tile_count, uint32_t spp_first,
                  uint32_t spp_count, uint32_t spp_chunk, const scn_wf_paths *paths, const scn_wf_shadow *light_rays, const scn_wf_shadow *phase_rays,
                  float *const framebuffer[3], void *stream) {
    if (!scene || !m || !paths || !framebuffer || !framebuffer[0]) return set_error(SCN_EINVAL, "null argument");
    if (spp_chunk == 0) return set_error(SCN_EINVAL, "render: spp_chunk must be >= 1");
    const bool nee = scene->nee != 0;
    if (nee && (!light_rays || !phase_rays || !framebuffer[1] || !framebuffer[2])) return set_error(SCN_EINVAL, "render: scene.nee needs both shadow wavefronts and three framebuffer planes");
    const size_t n_pixels = (size_t)tile_count * SCN_WF_TILE * SCN_WF_TILE;
    if (n_pixels == 0 || spp_count == 0) return SCN_OK;
    WfPaths W; if (int rc = wf_paths(paths, W, true)) return rc;
    WfShadow L{}, Q{};
    if (nee) {
        if (int rc = wf_nee_arrays(paths)) return rc;
        if (int rc = wf_shadow(light_rays, L)) return rc;
        if (int rc = wf_shadow(phase_rays, Q)) return rc;
    }
    cudaStream_t s = (cudaStream_t)stream;
    for (uint32_t done = 0; done < spp_count; done += spp_chunk) {
        const uint32_t c = spp_count - done < spp_chunk ? spp_count - done : spp_chunk;
        const size_t n = n_pixels * c;
        const unsigned grid = (unsigned)((n + WF_BLOCK - 1) / WF_BLOCK);
        if (int rc = scn_wf_raygen_strided(scene, m, tile_first, tile_stride, tile_count, spp_first + done, c, paths, n, stream)) return rc;
        for (int32_t bounce = 0; bounce < scene->max_bounces; ++bounce) {
            if (int rc = scn_sample_distance_batch(m, &paths->rays, &paths->state, &paths->samples, n, stream)) return rc;
            const bool direct = nee && bounce < scene->max_bounces - 1; // volumeEstimateDirect at every scatter of this bounce (TraceBase.cpp:546)
            if (direct) {
                if (int rc = scn_wf_nee_setup(scene, m, paths, light_rays, phase_rays, n, bounce, stream)) return rc;
                if (int rc = scn_transmittance_batch(m, &light_rays->rays, &light_rays->state, light_rays->T, n, stream)) return rc;
                if (int rc = scn_transmittance_batch(m, &phase_rays->rays, &phase_rays->state, phase_rays->T, n, stream)) return rc;
            }
            if (!nee) k_wf_shade<<<grid, WF_BLOCK, 0, s>>>(*scene, W, bounce, n);
            else if (direct) k_wf_shade_nee<true><<<grid, WF_BLOCK, 0, s>>>(*scene, W, L, Q, bounce, n);
            else k_wf_shade_nee<false><<<grid, WF_BLOCK, 0, s>>>(*scene, W, L, Q, bounce, n);
            if (int rc = wf_check("k_wf_shade")) return rc;
        }
        if (nee) if (int rc = scn_wf_clamp_samples(paths, n, stream)) return rc;
        if (int rc = scn_wf_accumulate(paths->radiance, framebuffer[0], n_pixels, c, stream)) return rc;
        if (nee) for (int j = 0; j < 2; ++j) if (int rc = scn_wf_accumulate(paths->radiance_gb[j], framebuffer[1 + j], n_pixels, c, stream)) return rc;
    }
    return SCN_OK;
}

} // extern "C"
