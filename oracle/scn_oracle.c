/* oracle/scn_oracle.c — TEST INFRASTRUCTURE ONLY (see scn_oracle.h for the rules and parity status: PINNED
 * against the compiled reference, oracle/_ref).
 *
 * Scalar C restatement of the reference's sparse-convolution-noise GPIS medium query.  File:line citations
 * are relative to /root/reference/src/core/.  Short names:
 *   SCN  = math/SparseConvolutionNoise.cpp      GPF = math/GPFunctions.cpp / .hpp
 *   SCNM = media/SparseConvolutionNoiseMedium.cpp   GPM = media/GaussianProcessMedium.cpp
 *
 * Build: gcc -std=gnu11 -O2 -msse4.2 -ffp-contract=off (no FMA contraction: the impulse rejection test is a
 * hard discontinuity, so every float op feeding it is rounded separately, exactly like the reference's
 * -msse4.2 build).  The draw order of next3D (z, y, x) is the one g++ gives the reference.
 */
#define _GNU_SOURCE
#include "scn_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;
typedef struct { double x, y, z; } d3;
typedef struct { v3 n, t, b; } frame_t;

#define ANISO_ON_AXIS 1.5f /* ProceduralNonstationaryCovariance::_anisotropyOnAxis, GPF.hpp:2372 */
static __thread char g_err[256];
static __thread uint64_t g_eval_count;
const char *orc_last_error(void) { return g_err; }
uint64_t orc_last_eval_count(void) { return g_eval_count; }
int orc_hw_threads(void) { long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }

/* ------------------------------------------------------------------ small vector helpers (math/Vec.hpp:134-206) */
static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v4 V4(float x, float y, float z, float w) { v4 r = {x, y, z, w}; return r; }
static inline v3 v3_add(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3_sub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3_mul(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 v3_scale(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline v3 v3_div(v3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
static inline float v3_dot(v3 a, v3 b) { float s = a.x * b.x; s += a.y * b.y; s += a.z * b.z; return s; }
static inline float v3_lensq(v3 a) { float r = 0.f; r += a.x * a.x; r += a.y * a.y; r += a.z * a.z; return r; }
static inline float v3_len(v3 a) { return sqrtf(v3_lensq(a)); }
static inline v3 v3_normalized(v3 a) { float inv = 1.0f / v3_len(a); return V3(a.x * inv, a.y * inv, a.z * inv); }
static inline v3 v3_cross(v3 a, v3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline v4 v4_add(v4 a, v4 b) { return V4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
static inline v4 v4_scale(v4 a, float s) { return V4(a.x * s, a.y * s, a.z * s, a.w * s); }
static inline v4 v4_div(v4 a, float s) { return V4(a.x / s, a.y / s, a.z / s, a.w / s); }
static inline float max3f(float a, float b, float c) { float m = a > b ? a : b; return m > c ? m : c; }
static inline double d3_dot(d3 a, d3 b) { double s = a.x * b.x; s += a.y * b.y; s += a.z * b.z; return s; }
static inline double d3_len(d3 a) { double r = 0.; r += a.x * a.x; r += a.y * a.y; r += a.z * a.z; return sqrt(r); }
static inline float sqrf(float a) { return a * a; }

/* ------------------------------------------------------------------ H1: xxhash32 (math/MathUtil.hpp:179-224) */
#define P2 2246822519U
#define P3 3266489917U
#define P4 668265263U
#define P5 374761393U
static inline uint32_t rotl17(uint32_t h) { return (h << 17) | (h >> 15); }
static inline uint32_t xx_final(uint32_t h) { h = P2 * (h ^ (h >> 15)); h = P3 * (h ^ (h >> 13)); return h ^ (h >> 16); }
uint32_t orc_xxhash32_1(uint32_t p) { uint32_t h = p + P5; h = P4 * rotl17(h); return xx_final(h); }
uint32_t orc_xxhash32_2(uint32_t x, uint32_t y) { uint32_t h = y + P5 + x * P3; h = P4 * rotl17(h); return xx_final(h); }
uint32_t orc_xxhash32_3(uint32_t x, uint32_t y, uint32_t z) {
    uint32_t h = z + P5 + x * P3; h = P4 * rotl17(h); h += y * P3; h = P4 * rotl17(h); return xx_final(h);
}
uint32_t orc_xxhash32_4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    uint32_t h = w + P5 + x * P3; h = P4 * rotl17(h); h += y * P3; h = P4 * rotl17(h); h += z * P3; h = P4 * rotl17(h);
    return xx_final(h);
}

/* ------------------------------------------------------------------ H2: PCG32 (sampling/UniformSampler.hpp:41-75, math/BitManip.hpp:47-50) */
typedef struct { uint64_t state; } pcg_t;
static inline uint32_t pcg_nexti(pcg_t *s) {
    uint64_t old = s->state;
    s->state = old * 6364136223846793005ULL + 1ULL; /* _sequence == 0 -> (0 | 1) */
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31u));
}
static inline float pcg_next1d(pcg_t *s) {
    union { uint32_t i; float f; } u;
    u.i = (pcg_nexti(s) >> 9u) | 0x3F800000u;
    return u.f - 1.0f;
}
static inline void pcg_set_state(pcg_t *s, uint64_t st) { s->state = st; pcg_next1d(s); pcg_next1d(s); }
/* next3D = Vec3f(next1D(), next1D(), next1D()) with g++'s right-to-left argument evaluation: z first. */
static inline v3 pcg_next3d(pcg_t *s) { v3 r; r.z = pcg_next1d(s); r.y = pcg_next1d(s); r.x = pcg_next1d(s); return r; }

void orc_pcg_stream(uint64_t state, int n, float *out) { pcg_t s; pcg_set_state(&s, state); for (int i = 0; i < n; ++i) out[i] = pcg_next1d(&s); }
void orc_pcg_next3d(uint64_t state, float *out) {
    pcg_t s; pcg_set_state(&s, state); v3 v = pcg_next3d(&s); out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = pcg_next1d(&s);
}
/* H3: Box-Muller in double (sampling/Gaussian.cpp:21-33); PI is a float constant (math/Angle.hpp:8) */
static inline void rand_normal_2(pcg_t *s, double *z1, double *z2) {
    double u1 = pcg_next1d(s), u2 = pcg_next1d(s);
    const float PI_F = 3.1415926536f;
    double r = sqrt(-2 * log(1. - u1));
    float twopi = 2 * PI_F;
    *z1 = r * cos(twopi * u2);
    *z2 = r * sin(twopi * u2);
}
void orc_box_muller(uint64_t state, int npairs, double *out) {
    pcg_t s; pcg_set_state(&s, state); for (int i = 0; i < npairs; ++i) rand_normal_2(&s, out + 2 * i, out + 2 * i + 1);
}
void orc_cell_impulses(uint32_t ci, uint32_t cj, uint32_t ck, uint32_t seed, uint32_t count, float *out) {
    pcg_t s; pcg_set_state(&s, (uint64_t)(uint32_t)(orc_xxhash32_4(ck, cj, ci, seed) + 1u)); /* SCN:376 */
    for (uint32_t k = 0; k < count; ++k) {
        v3 p = pcg_next3d(&s); float w = pcg_next1d(&s) < 0.5f ? -1.f : 1.f;
        out[4 * k] = p.x; out[4 * k + 1] = p.y; out[4 * k + 2] = p.z; out[4 * k + 3] = w;
    }
}
void orc_multires_level(int k, float *scale, int *additional_seed) {
    float base = 2.5f; float s = powf(base, (float)k); *scale = s; *additional_seed = (int)floorf(logf(s) / logf(base));
}

/* ------------------------------------------------------------------ TangentFrame (math/TangentFrame.hpp:74-99) */
static frame_t frame_make(v3 nin) {
    frame_t f; f.n = v3_normalized(nin);
    float sign = copysignf(1.0f, f.n.z);
    const float a = -1.0f / (sign + f.n.z);
    const float b = f.n.x * f.n.y * a;
    f.t = V3(1.0f + sign * f.n.x * f.n.x * a, sign * b, -sign * f.n.x);
    f.b = V3(b, sign + f.n.y * f.n.y * a, -f.n.y);
    f.t = v3_normalized(f.t); f.b = v3_normalized(f.b);
    return f;
}
static inline v3 frame_to_local(const frame_t *f, v3 p) { return V3(v3_dot(f->t, p), v3_dot(f->b, p), v3_dot(f->n, p)); }
static inline v3 frame_to_global(const frame_t *f, v3 p) { return v3_add(v3_add(v3_scale(f->t, p.x), v3_scale(f->b, p.y)), v3_scale(f->n, p.z)); }

static void mat3_inverse(const float a[9], float r[9]); /* Eigen's fixed-size 3x3 inverse (cofactor form), defined below */
/* Eigen's fixed-size products reduce three terms as x0 + (x1 + x2) */
static inline v3 mat3_mulv_e(const float a[9], v3 b) { return V3(a[0] * b.x + (a[1] * b.y + a[2] * b.z), a[3] * b.x + (a[4] * b.y + a[5] * b.z), a[6] * b.x + (a[7] * b.y + a[8] * b.z)); }
static inline v3 mat3_tmulv_e(const float a[9], v3 b) { return V3(a[0] * b.x + (a[3] * b.y + a[6] * b.z), a[1] * b.x + (a[4] * b.y + a[7] * b.z), a[2] * b.x + (a[5] * b.y + a[8] * b.z)); }
/* ------------------------------------------------------------------ medium */
struct orc_medium {
    scn_medium_desc d;
    uint32_t n_imp;          /* uint(impulseDensity), SCN:380,413 */
    int stationary, multires;
    float l_conv; v3 l_aniso, l_aniso_inv; /* GPF:654-667 */
    float ls_max;            /* sparseConvNoiseMaxLateralScale, GPF:1729-1733 */
    float sigma_s_over_t;
    int absorption_only;
    int kernel;              /* SCN_COV_SQUARED_EXPONENTIAL | SCN_COV_MATERN | SCN_COV_GABOR_ANISO | SCN_COV_GABOR_ISO */
    float k_a, k_f; v3 k_omega; /* Gabor: _a, _f, _omega (GPF:1087-1097) */
    int use_mtx;             /* "useAnisoMtx": full matrices (GPF:668-675), row-major: L = _local_to_world, W = _world_to_local, C = _cov_mtx_inv */
    float L[9], W[9], C[9], C_det;
    int has_aniso;           /* proc_nonstationary "aniso" field: rotation anisotropy of the splat kernel (GPF:1678-1690) */
};

orc_medium *orc_medium_create(const scn_medium_desc *desc) {
    if (!desc) { snprintf(g_err, sizeof g_err, "null desc"); return NULL; }
    orc_medium *m = (orc_medium *)calloc(1, sizeof *m);
    m->d = *desc;
    m->n_imp = (uint32_t)desc->impulse_density;
    m->stationary = desc->cov_type != SCN_COV_PROC_NONSTATIONARY;
    m->kernel = desc->cov_type >= SCN_COV_MATERN ? desc->cov_type : SCN_COV_SQUARED_EXPONENTIAL;
    if (m->kernel != SCN_COV_SQUARED_EXPONENTIAL) {
        /* GPF:870-880, 1087-1097, 1156-1163: no second derivative, no 1D family, transforms never initialised */
        int cond = !desc->single_realization && (desc->correlation_context == SCN_CTX_RENEWAL || desc->correlation_context == SCN_CTX_RENEWAL_PLUS);
        if (desc->sampling_1d || desc->isotropic_3d_sampling || cond) { snprintf(g_err, sizeof g_err, "matern / gabor: world-space 3D sampling without conditioning only"); free(m); return NULL; }
        m->k_a = (float)(1.0 / desc->gabor_a_inv); m->k_f = (float)(1.0 / desc->gabor_f_inv);
        m->k_omega = V3(desc->gabor_omega[0], desc->gabor_omega[1], desc->gabor_omega[2]);
    }
    m->multires = !m->stationary && desc->multires;
    m->has_aniso = !m->stationary && desc->aniso_field.type != SCN_FIELD_NONE;
    if (m->has_aniso && desc->sampling_1d) { snprintf(g_err, sizeof g_err, "proc_nonstationary \"aniso\": 3D sampling only"); free(m); return NULL; }
    m->l_conv = desc->length_scale * sqrtf(2.f) / 2;
    m->l_aniso = V3(m->l_conv * desc->aniso[0], m->l_conv * desc->aniso[1], m->l_conv * desc->aniso[2]);
    m->l_aniso_inv = V3(1.0f / m->l_aniso.x, 1.0f / m->l_aniso.y, 1.0f / m->l_aniso.z);
    if (isinf(m->l_aniso_inv.x) || isnan(m->l_aniso_inv.x)) m->l_aniso_inv.x = 0;
    if (isinf(m->l_aniso_inv.y) || isnan(m->l_aniso_inv.y)) m->l_aniso_inv.y = 0;
    if (isinf(m->l_aniso_inv.z) || isnan(m->l_aniso_inv.z)) m->l_aniso_inv.z = 0;
    m->use_mtx = desc->use_aniso_mtx != 0;
    if (m->use_mtx) { /* GPF:668-675 */
        if (m->kernel != SCN_COV_SQUARED_EXPONENTIAL || m->has_aniso) { snprintf(g_err, sizeof g_err, "useAnisoMtx: squared-exponential kernel without an anisotropy field only"); free(m); return NULL; }
        for (int i = 0; i < 9; ++i) m->L[i] = m->l_conv * desc->aniso_mtx[i];
        mat3_inverse(m->L, m->W);
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m->C[i * 3 + j] = m->W[0 * 3 + i] * m->W[0 * 3 + j] + (m->W[1 * 3 + i] * m->W[1 * 3 + j] + m->W[2 * 3 + i] * m->W[2 * 3 + j]);
        const float *c = m->C; /* Eigen's 3x3 determinant (LU/Determinant.h: bruteforce_det3_helper) */
        m->C_det = c[0] * (c[4] * c[8] - c[5] * c[7]) - c[1] * (c[3] * c[8] - c[5] * c[6]) + c[2] * (c[3] * c[7] - c[4] * c[6]);
    }
    m->ls_max = 1.f;
    if (!m->stationary && desc->ls.type != SCN_FIELD_NONE) {
        const scn_field_desc *f = &desc->ls;
        if (f->type == SCN_FIELD_CONSTANT) { double v = f->val[0] > f->val[1] ? f->val[0] : f->val[1]; v = v > f->val[2] ? v : f->val[2]; m->ls_max = (float)v; }
        else if (f->type == SCN_FIELD_SANDSTONE || f->type == SCN_FIELD_RUST) m->ls_max = 1.f; /* ProceduralNoiseVec::maxVal (GPF:131-137) */
        else if (f->type == SCN_FIELD_REGULAR_GRID) m->ls_max = 1.f; /* the reference's maxVal() of a grid field is 0 (GPF.hpp:474-477): multi-resolution form only, which never reads it */
        else if (f->type == SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT) m->ls_max = (float)((f->max > f->min ? f->max : f->min) * (f->max2 > f->min2 ? f->max2 : f->min2));
        else m->ls_max = (float)(f->max > f->min ? f->max : f->min);
    }
    /* GPM:152-158 */
    float sa = desc->sigma_a * desc->density, ss = desc->sigma_s * desc->density, st = sa + ss;
    m->absorption_only = ss == 0.0f;
    m->sigma_s_over_t = ss / st;
    return m;
}
void orc_medium_destroy(orc_medium *m) { free(m); }

/* ------------------------------------------------------------------ procedural fields (GPF:53-138, GPF.hpp:604-800) */
static inline double lerpd(double a, double b, double r) { return a * (1. - r) + b * r; }
static inline double clampd(double v, double lo, double hi) { double t = v > lo ? v : lo; return t < hi ? t : hi; }
static double field_ramp(const scn_field_desc *f, d3 p, int second) {
    const double c = 1.; /* _const, never read from JSON */
    if (!second) {
        double mn = f->min + c, mx = f->max + c, scale = 1.0 / (f->end - f->start), off = -f->start * scale;
        double coord = f->type == SCN_FIELD_LEFT_RIGHT ? p.x : (f->type == SCN_FIELD_FRONT_BACK ? p.z : p.y);
        return sqrt(exp(lerpd(log(mn * mn), log(mx * mx), clampd(coord * scale + off, 0.0, 1.0))));
    } else {
        double mn = f->min2 + c, mx = f->max2 + c, scale = 1.0 / (f->end2 - f->start2), off = -f->start2 * scale;
        return sqrt(exp(lerpd(log(mn * mn), log(mx * mx), clampd(p.x * scale + off, 0.0, 1.0))));
    }
}
/* RegularGrid<ElemType>::getValue (GPF.hpp:255-306): component c of ncomp interleaved doubles; trilinearInterpolation (:173-188),
 * triquadraticInterpolation (:190-230), lerp = a (1 - t) + b t (math/MathUtil.hpp:91-94) */
static double grid_at(const scn_grid_desc *g, int ncomp, int x, int y, int z, int c) {
    int r = (int)g->res;
    x = x < 0 ? 0 : (x > r - 1 ? r - 1 : x); y = y < 0 ? 0 : (y > r - 1 ? r - 1 : y); z = z < 0 ? 0 : (z > r - 1 ? r - 1 : z);
    return g->values[((size_t)(x * r + y) * r + z) * ncomp + c];
}
static double quadd(double v0, double v1, double v2, double w) { double a = 0.5 * (v0 + v2) - v1, b = 0.5 * (v2 - v0); return w * (w * a + b) + v1; }
static double field_grid(const scn_grid_desc *g, int ncomp, d3 p, int c) {
    double res = (double)g->res;
    double gx = (res * (p.x - g->bounds_min[0])) / (g->bounds_max[0] - g->bounds_min[0]) - 0.5;
    double gy = (res * (p.y - g->bounds_min[1])) / (g->bounds_max[1] - g->bounds_min[1]) - 0.5;
    double gz = (res * (p.z - g->bounds_min[2])) / (g->bounds_max[2] - g->bounds_min[2]) - 0.5;
    int ix = (int)floor(gx), iy = (int)floor(gy), iz = (int)floor(gz);
    double u = gx - (double)ix, v = gy - (double)iy, w = gz - (double)iz;
    if (g->interp == SCN_INTERP_POINT) return grid_at(g, ncomp, ix, iy, iz, c);
    if (g->interp == SCN_INTERP_LINEAR) {
        double z0 = lerpd(lerpd(grid_at(g, ncomp, ix, iy, iz, c), grid_at(g, ncomp, ix + 1, iy, iz, c), u), lerpd(grid_at(g, ncomp, ix, iy + 1, iz, c), grid_at(g, ncomp, ix + 1, iy + 1, iz, c), u), v);
        double z1 = lerpd(lerpd(grid_at(g, ncomp, ix, iy, iz + 1, c), grid_at(g, ncomp, ix + 1, iy, iz + 1, c), u), lerpd(grid_at(g, ncomp, ix, iy + 1, iz + 1, c), grid_at(g, ncomp, ix + 1, iy + 1, iz + 1, c), u), v);
        return lerpd(z0, z1, w);
    }
    double vx[3];
    for (int dx = 0; dx < 3; ++dx) {
        double vy[3];
        for (int dy = 0; dy < 3; ++dy)
            vy[dy] = quadd(grid_at(g, ncomp, ix - 1 + dx, iy - 1 + dy, iz - 1, c), grid_at(g, ncomp, ix - 1 + dx, iy - 1 + dy, iz, c), grid_at(g, ncomp, ix - 1 + dx, iy - 1 + dy, iz + 1, c), w);
        vx[dx] = quadd(vy[0], vy[1], vy[2], v);
    }
    return quadd(vx[0], vx[1], vx[2], u);
}
/* fbm fields: random3 / simplex3d / fbm (math/SdfFunctions.cpp:200-296), float arithmetic, sin in double */
static void fbm_random3(float cx, float cy, float cz, float *r) {
    float dt = cx * 17.0f + cy * 59.4f + cz * 15.0f;
    float j = (float)(4096.0 * sin((double)dt));
    double v = 512.0 * j; r[2] = (float)(v - floor(v));
    j = (float)(j * .125);
    v = 512.0 * j; r[0] = (float)(v - floor(v));
    j = (float)(j * .125);
    v = 512.0 * j; r[1] = (float)(v - floor(v));
    r[0] -= 0.5f; r[1] -= 0.5f; r[2] -= 0.5f;
}
static float dot3f(const float *a, const float *b) { float s = a[0] * b[0]; s += a[1] * b[1]; s += a[2] * b[2]; return s; }
static float fbm_simplex3d(float px, float py, float pz) {
    const float F3c = 0.3333333f, G3c = 0.1666667f;
    float p[3] = {px, py, pz}, k3[3] = {F3c, F3c, F3c}, g3[3] = {G3c, G3c, G3c};
    float ds = dot3f(p, k3), s[3], x[3], e[3], i1[3], i2[3], x1[3], x2[3], x3[3], w[4], d[4], r[3], q[3];
    for (int i = 0; i < 3; ++i) s[i] = floorf(p[i] + ds);
    float dg = dot3f(s, g3);
    for (int i = 0; i < 3; ++i) x[i] = p[i] - s[i] + dg;
    for (int i = 0; i < 3; ++i) e[i] = (x[i] - x[(i + 1) % 3]) < 0.0f ? 0.f : 1.f;
    for (int i = 0; i < 3; ++i) { i1[i] = e[i] * (1.0f - e[(i + 2) % 3]); i2[i] = 1.0f - e[(i + 2) % 3] * (1.0f - e[i]); }
    for (int i = 0; i < 3; ++i) { x1[i] = x[i] - i1[i] + G3c; x2[i] = x[i] - i2[i] + 2.0f * G3c; x3[i] = x[i] - 1.0f + 3.0f * G3c; }
    w[0] = dot3f(x, x); w[1] = dot3f(x1, x1); w[2] = dot3f(x2, x2); w[3] = dot3f(x3, x3);
    for (int i = 0; i < 4; ++i) { float t = 0.6f - w[i]; w[i] = t > 0.0f ? t : 0.0f; }
    fbm_random3(s[0], s[1], s[2], r); d[0] = dot3f(r, x);
    for (int i = 0; i < 3; ++i) q[i] = s[i] + i1[i]; fbm_random3(q[0], q[1], q[2], r); d[1] = dot3f(r, x1);
    for (int i = 0; i < 3; ++i) q[i] = s[i] + i2[i]; fbm_random3(q[0], q[1], q[2], r); d[2] = dot3f(r, x2);
    for (int i = 0; i < 3; ++i) q[i] = s[i] + 1.0f; fbm_random3(q[0], q[1], q[2], r); d[3] = dot3f(r, x3);
    for (int i = 0; i < 4; ++i) { w[i] *= w[i]; w[i] *= w[i]; d[i] *= w[i]; }
    float sum = d[0] * 52.0f; sum += d[1] * 52.0f; sum += d[2] * 52.0f; sum += d[3] * 52.0f;
    return sum;
}
static double fbm_noise(d3 uv, int octaves) {
    float gain = 0.65f, lacunarity = 2.1042f, total, frequency = 0.5f, amplitude = gain;
    uv.x *= 5.0; uv.y *= 5.0; uv.z *= 5.0;
    float qx = (float)uv.x, qy = (float)uv.y, qz = (float)uv.z;
    total = fbm_simplex3d(qx, qy, qz);
    for (int i = 0; i < octaves; ++i) { total += fbm_simplex3d(qx * frequency, qy * frequency, qz * frequency) * amplitude; frequency *= lacunarity; amplitude *= gain; }
    total = (float)((total + 2.0) / 4.0);
    return total;
}
static d3 d3_adds(d3 a, double s) { d3 r = {a.x + s, a.y + s, a.z + s}; return r; }
static d3 d3_muls(d3 a, double s) { d3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static double fbm_sandstone(d3 p, int oct) { d3 ps = d3_muls(p, 0.3); return fbm_noise(d3_adds(ps, fbm_noise(d3_adds(ps, fbm_noise(ps, oct)), oct)), oct); }
static double fbm_rust(d3 p, int oct, int plus) { /* GPF:78-83 (scalar: minus), 112-117 (vector: plus); smoothStep: math/MathUtil.hpp:108-112 */
    d3 ps = d3_muls(p, 2.0);
    double a = fbm_noise(d3_muls(ps, .1), oct) * 0.4, b = fbm_noise(d3_muls(ps, 25.), oct) * 0.1;
    double v = plus ? fbm_noise(d3_adds(ps, a), oct) + b : fbm_noise(d3_adds(ps, a), oct) - b;
    double x = clampd((v - 0.4) / (0.6 - 0.4), 0., 1.);
    return x * x * (3.0 - 2.0 * x);
}
static double field_scalar(const scn_field_desc *f, d3 p) {
    switch (f->type) {
    case SCN_FIELD_SANDSTONE: { double v = fbm_sandstone(p, 2); double cx = sqrt(v * 1.9 * 1.2) - 0.35; return lerpd(f->min, f->max, clampd(cx, 0., 1.)); } /* GPF:70-77 */
    case SCN_FIELD_RUST: return lerpd(f->min, f->max, clampd(fbm_rust(p, 2, 0), 0., 1.));
    case SCN_FIELD_REGULAR_GRID: return field_grid(&f->grid, 1, p, 0); /* RegularGridScalar (GPF.hpp:429-437) */
    case SCN_FIELD_CONSTANT: return f->val[0];
    case SCN_FIELD_BOTTOM_TOP: case SCN_FIELD_LEFT_RIGHT: case SCN_FIELD_FRONT_BACK: return field_ramp(f, p, 0) - 1.;
    case SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT: { float bt = (float)field_ramp(f, p, 0), lr = (float)field_ramp(f, p, 1); return bt * lr - 1. * 1.; }
    default: return 1.;
    }
}
static d3 field_vector(const scn_field_desc *f, d3 p) {
    d3 r;
    if (f->type == SCN_FIELD_CONSTANT) { r.x = f->val[0]; r.y = f->val[1]; r.z = f->val[2]; return r; }
    if (f->type == SCN_FIELD_SANDSTONE) { /* GPF:104-111 */
        double v = fbm_sandstone(p, 10);
        r.x = clampd((sqrt(v * 1.9 * 1.2) - 0.35) * 0.2, 0., 1.); r.y = clampd((sqrt(v * 0.7 * 1.2) - 0.35) * 0.2, 0., 1.); r.z = clampd((sqrt(v * 0.25 * 1.2) - 0.35) * 0.2, 0., 1.); return r;
    }
    if (f->type == SCN_FIELD_RUST) { double v = fbm_rust(p, 10, 1); r.x = lerpd(0.278, 1., v); r.y = lerpd(0.212, 1., v); r.z = lerpd(0.141, 1., v); return r; } /* GPF:112-117 */
    if (f->type == SCN_FIELD_REGULAR_GRID) { r.x = field_grid(&f->grid, 3, p, 0); r.y = field_grid(&f->grid, 3, p, 1); r.z = field_grid(&f->grid, 3, p, 2); return r; } /* RegularGridVector (GPF.hpp:504-512) */
    double v = field_scalar(f, p); r.x = r.y = r.z = v; return r;
}

/* ------------------------------------------------------------------ covariance helpers */
static inline d3 to_d3(v3 p) { d3 r = {p.x, p.y, p.z}; return r; }
/* getKernelScale / sparseConvNoiseLateralScale (GPF:607-609, 1219-1221, 1729-1735) */
static float cov_lateral_scale(const orc_medium *m, v3 p) {
    if (m->stationary || m->d.ls.type == SCN_FIELD_NONE) return 1.0f;
    d3 ls = field_vector(&m->d.ls, to_d3(p));
    return max3f((float)ls.x, (float)ls.y, (float)ls.z);
}
/* nonStationarySplattingKernelScale (GPF.hpp:1493, GPF:1223-1229) */
static float cov_ns_scale(const orc_medium *m, v3 p) {
    if (m->stationary || m->multires) return 1.f;
    return cov_lateral_scale(m, p) / m->ls_max;
}
static float cov_world_scale(const orc_medium *m) { return m->stationary ? 1.f : m->ls_max; } /* GPF.hpp:1494, GPF:1231-1233 */
/* sparseConvNoiseAmplitude (GPF:611-613,709; GPF:1235-1237) */
static float cov_amplitude(const orc_medium *m, v3 p) {
    if (m->stationary) return m->d.sigma;
    double var = m->d.var.type != SCN_FIELD_NONE ? field_scalar(&m->d.var, to_d3(p)) : 1.;
    return (float)(var * m->d.sigma);
}
/* splattingKernelRadius (GPF:696-709, 1245-1249) */
static float cov_radius(const orc_medium *m, int is_identity, float local_scale) {
    if (!m->stationary) { local_scale *= m->multires ? 1.0 : m->ls_max; local_scale *= m->has_aniso ? ANISO_ON_AXIS : 1.f; } /* max(a, 1/a), GPF:1741-1745 */
    float sf = m->d.kernel_scale;
    if (m->kernel == SCN_COV_MATERN) { /* GPF:1020-1025 */
        if (is_identity) return sf;
        float v[3], mf = 0.f;
        for (int i = 0; i < 3; ++i) { v[i] = m->d.length_scale / sqrtf(m->d.aniso[i]); if (isinf(v[i]) || isnan(v[i])) v[i] = 0.f; }
        mf = max3f(v[0], v[1], v[2]);
        return (float)(sf * local_scale * sqrt(2) / 2 * mf);
    }
    if (m->kernel == SCN_COV_GABOR_ANISO) return (float)(sf * sqrt(2) / 2 * 1.0 / m->k_a); /* GPF:1127-1130 */
    if (m->kernel == SCN_COV_GABOR_ISO) return (float)(sf * sqrt(2) / 4 * 1.0 / m->k_a);   /* GPF:1185-1188 */
    if (is_identity) return sf;
    float mf = max3f(m->l_aniso.x, m->l_aniso.y, m->l_aniso.z);
    if (m->use_mtx) mf = max3f((m->L[0] + m->L[1]) + m->L[2], (m->L[3] + m->L[4]) + m->L[5], (m->L[6] + m->L[7]) + m->L[8]); /* col0 + col1 + col2, GPF:704-706 */
    return sf * local_scale * mf;
}
/* Matern / Gabor sparseConvNoiseVariance3D (GPF:1029-1044, 1134-1138, 1192-1196) */
static float alt_variance3d(const orc_medium *m, float density, float R) {
    double per_unit = density / (R * R * R), integral;
    if (m->kernel == SCN_COV_MATERN) {
        float l = m->d.length_scale, v = m->d.matern_v;
        if (v == 0.5f) integral = 2.0 * M_PI * l;
        else if (v == 1.5f) integral = pow(M_PI * l, 3) / (24 * sqrt(3));
        else if (v == 2.5f) integral = M_PI * pow(l, 3) / (5 * sqrt(5));
        else return 0.f;
    } else if (m->kernel == SCN_COV_GABOR_ANISO) integral = pow(1.0 / m->k_a, 3) * (1 + exp(-2.0 * M_PI * sqrf(m->k_f / m->k_a))) / (4 * sqrt(2));
    else integral = 2 * sqrt(2) * M_PI * sqrf(m->k_f) / m->k_a * (1 - exp(-2 * M_PI * m->k_f / sqrf(m->k_a)));
    return (float)(per_unit * integral);
}
/* modified Bessel function of the second kind K_n(x), n = 0, 1 (the reference calls boost::math::cyl_bessel_k): trapezoid rule on
 * K_n(x) = int_0^inf exp(-x cosh t) cosh(n t) dt, which converges geometrically for this integrand */
static double bessel_k(int n, double x) {
    const double h = 0.125;
    double sum = 0.5 * exp(-x);
    for (int i = 1; i < 4096; ++i) {
        double t = i * h, e = x * cosh(t);
        if (e > 745.) break;
        sum += exp(-e) * (n ? cosh(t) : 1.0);
    }
    return sum * h;
}
/* splattingKernel3DVal / 1stGrad of MaternCovariance (GPF:1046-1082), GaborAnisotropicCovariance (GPF:1142-1153) and
 * GaborIsotropicCovariance (GPF:1200-1214); isCov, isIsotropic and the scales are ignored by these classes */
static v4 alt_kernel3d(const orc_medium *m, v3 ab) {
    if (m->kernel == SCN_COV_MATERN) {
        double abLen = v3_len(ab); float l = m->d.length_scale, v = m->d.matern_v, val, gk;
        if (v == 0.5f) { val = (float)(exp(-abLen / l) / abLen); gk = -(float)(exp(-abLen / l) * (1 / pow(abLen, 3) - 1 / (pow(abLen, 2) * l))); }
        else if (v == 1.5f) { double rs = sqrt(3.) * abLen / l; val = (float)bessel_k(0, rs); gk = -(float)(bessel_k(1, rs) * sqrt(3.) / l / abLen); }
        else { val = (float)exp(-sqrt(5.) * abLen / l); gk = -(float)(exp(-sqrt(5.) * abLen / l) * sqrt(5.) / l / abLen); }
        return V4(val, gk * ab.x, gk * ab.y, gk * ab.z);
    }
    float a = m->k_a, f = m->k_f;
    if (m->kernel == SCN_COV_GABOR_ANISO) {
        float od = v3_dot(m->k_omega, ab);
        float val = (float)(exp(-M_PI * sqrf(a) * v3_lensq(ab)) * cos(2.f * M_PI * f * od));
        float A = (float)exp(-M_PI * sqrf(a) * v3_lensq(ab)), B = (float)cos(2.f * M_PI * f * od);
        float g1 = -(float)(A * sin(2.f * M_PI * f * od) * 2.f * M_PI * f), g2 = (float)(B * A * 2.f * M_PI * sqrf(a));
        v3 g = v3_sub(v3_scale(m->k_omega, g1), v3_scale(ab, g2));
        return V4(val, g.x, g.y, g.z);
    }
    float r = v3_len(ab);
    float val = (float)(exp(-M_PI * sqrf(a * r)) * 2 * f / r * sin(2 * M_PI * f * r));
    float gk = (float)(2 * f * exp(-M_PI * sqrf(a) * v3_lensq(ab)) *
                       (-sin(2 * M_PI * f * r) / pow(r, 3) - 2 * M_PI * sqrf(a) * sin(2 * M_PI * f * r) / r + 2 * M_PI * f * cos(2 * M_PI * f * r) / sqrf(r)));
    return V4(val, gk * ab.x, gk * ab.y, gk * ab.z);
}
/* SquaredExponentialCovariance::sparseConvNoiseVariance3D/1D (GPF:741-760) */
static float se_variance3d(const orc_medium *m, float density, float R, int is_identity, float global_scale, float local_scale) {
    double per_unit = density / (R * R * R);
    double det_sqrt = 1.0;
    if (!is_identity) {
        det_sqrt = m->l_aniso.x * m->l_aniso.y * m->l_aniso.z;
        if (m->use_mtx) det_sqrt = 1.0 / sqrt(m->C_det); /* GPF:748-749 */
        det_sqrt *= pow(global_scale, 3);
    }
    det_sqrt *= pow(local_scale, 3);
    double integral = pow(M_PI, 1.5) * det_sqrt;
    return (float)(per_unit * integral);
}
static float se_variance1d(float density, float R, float local_scale) {
    double per_unit = density / R; double integral = sqrt(M_PI) * local_scale; return (float)(per_unit * integral);
}
static float alt_variance3d(const orc_medium *m, float density, float R);
static float cov_variance3d(const orc_medium *m, v3 p, float density, float R, int is_identity, float global_scale) { /* GPF:615-617, 1251-1254 */
    if (m->kernel != SCN_COV_SQUARED_EXPONENTIAL) return alt_variance3d(m, density, R);
    if (m->stationary) return se_variance3d(m, density, R, is_identity, global_scale, 1.0f);
    float ls = m->multires ? 1.0f : cov_ns_scale(m, p);
    return se_variance3d(m, density, R, is_identity, global_scale, ls);
}
static float cov_variance1d(const orc_medium *m, v3 p, float density, float R) { /* GPF:619-621, 1256-1259 */
    return se_variance1d(density, R, m->stationary ? 1.0f : cov_ns_scale(m, p));
}
/* transforms (GPF:631-637, 1261-1284): diagonal world<->iso matrices */
static inline float tr_scale(const orc_medium *m, float s) { return m->stationary ? s : (m->multires ? s : m->ls_max); }
static v3 cov_pos_w2l(const orc_medium *m, v3 v, float s) { s = tr_scale(m, s); return v3_div(m->use_mtx ? mat3_mulv_e(m->W, v) : v3_mul(m->l_aniso_inv, v), s); }
static v3 cov_grad_w2l(const orc_medium *m, v3 v, float s) { s = tr_scale(m, s); return v3_scale(m->use_mtx ? mat3_tmulv_e(m->L, v) : v3_mul(m->l_aniso, v), s); }
static v3 cov_grad_l2w(const orc_medium *m, v3 v, float s) { s = tr_scale(m, s); return v3_div(m->use_mtx ? mat3_tmulv_e(m->W, v) : v3_mul(m->l_aniso_inv, v), s); }
/* getNonstationaryCovSplatCov1D (GPF.hpp:1213, GPF:1716-1727, no anisotropy field) */
static float cov_splat_cov1d(const orc_medium *m, v3 pq, v3 pc) {
    if (m->stationary) return 1.0f;
    float scale = (float)(0.5 * (sqrf(cov_ns_scale(m, pq)) + sqrf(cov_ns_scale(m, pc))));
    return sqrtf(scale);
}

/* Eigen's fixed-size 3x3 inverse (cofactor form) */
static void mat3_inverse(const float a[9], float r[9]);
/* getNonstationaryAniso3D (GPF:1678-1690) with compute_ansio_full (GPF:26-39): G = (S V)^T (S V), V = rotation by angle about z,
 * S = diag(a, 1/a, 1), a = _anisotropyOnAxis = 1.5 (GPF.hpp:2372); row-major.  Returns 0 without an anisotropy field. */
static int cov_aniso3d(const orc_medium *m, v3 p_world, float G[9]) {
    if (!m->has_aniso) return 0;
    float angle = (float)(field_scalar(&m->d.aniso_field, to_d3(p_world)) * (3.1415926536f * 0.5f));
    float c = cosf(angle), sn = sinf(angle);
    float ax[3] = {ANISO_ON_AXIS, 1.f / ANISO_ON_AXIS, 1.f};
    float V[9] = {c, sn, 0.f, -sn, c, 0.f, 0.f, 0.f, 1.f}, M[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M[i * 3 + j] = ax[i] * V[i * 3 + j];
    /* Eigen's unrolled three-term reduction is x0 + (x1 + x2) (redux_novec_unroller splits in halves) */
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) G[i * 3 + j] = M[0 * 3 + i] * M[0 * 3 + j] + (M[1 * 3 + i] * M[1 * 3 + j] + M[2 * 3 + i] * M[2 * 3 + j]);
    return 1;
}
/* getInvCovMtx with an anisotropy matrix (GPF:774-802): invSigma^T aniso_inv invSigma, then the isCov / localScale / 0.5 factors */
static void se_inv_cov_full(const orc_medium *m, const float aniso_inv[9], int is_cov, int is_iso, float gs, float ls, float M[9]) {
    float sg[3] = {1.f, 1.f, 1.f};
    if (!is_iso) { sg[0] = m->l_aniso_inv.x / gs; sg[1] = m->l_aniso_inv.y / gs; sg[2] = m->l_aniso_inv.z / gs; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M[i * 3 + j] = (sg[i] * aniso_inv[i * 3 + j]) * sg[j];
    for (int i = 0; i < 9; ++i) { if (is_cov) M[i] *= 0.5f; M[i] /= sqrf(ls); M[i] *= 0.5f; }
}

/* getInvCovMtx with "useAnisoMtx", not isotropic (GPF:791-801): _cov_mtx_inv / gs^2, then the isCov / localScale / 0.5 factors */
static void se_inv_cov_mtx(const orc_medium *m, int is_cov, float gs, float ls, float M[9]) {
    for (int i = 0; i < 9; ++i) { M[i] = m->C[i] / sqrf(gs); if (is_cov) M[i] *= 0.5f; M[i] /= sqrf(ls); M[i] *= 0.5f; }
}
/* getInvCovMtx diagonal (GPF:774-802) */
static v3 se_inv_cov_diag(const orc_medium *m, int is_cov, int is_iso, float gs, float ls) {
    v3 d = V3(1.f, 1.f, 1.f);
    if (!is_iso) {
        d = v3_mul(m->l_aniso_inv, m->l_aniso_inv);
        d = v3_div(d, sqrf(gs));
    }
    if (is_cov) d = v3_scale(d, 0.5f);
    d = v3_div(d, sqrf(ls));
    d = v3_scale(d, 0.5f);
    return d;
}
/* splattingKernel3D: value + first gradient (GPF:552-563, 804-817) */
static v4 cov_kernel3d(const orc_medium *m, v3 pa, v3 pb, int is_cov, int is_iso, float gs, v3 p_world) {
    float ls = cov_ns_scale(m, p_world);
    v3 ab = v3_sub(pa, pb);
    if (m->kernel != SCN_COV_SQUARED_EXPONENTIAL) return alt_kernel3d(m, ab);
    float G[9];
    int full = cov_aniso3d(m, p_world, G);
    if (full || (m->use_mtx && !is_iso)) { /* GPF:556-558: aniso_inv = inverse, full matrix; GPF:791-795: _cov_mtx_inv / gs^2 */
        float Gi[9], Mf[9];
        if (full) { mat3_inverse(G, Gi); se_inv_cov_full(m, Gi, is_cov, is_iso, gs, ls, Mf); } else se_inv_cov_mtx(m, is_cov, gs, ls, Mf);
        /* dist2_ab = (ab^T M) ab and ab . M.col(j) (GPF.hpp:76-78, GPF:804-817), three-term sums as x0 + (x1 + x2) */
        v3 Mab = V3(ab.x * Mf[0] + (ab.y * Mf[3] + ab.z * Mf[6]), ab.x * Mf[1] + (ab.y * Mf[4] + ab.z * Mf[7]), ab.x * Mf[2] + (ab.y * Mf[5] + ab.z * Mf[8]));
        float f = expf(-(Mab.x * ab.x + (Mab.y * ab.y + Mab.z * ab.z)));
        return V4(f, -2.f * Mab.x * f, -2.f * Mab.y * f, -2.f * Mab.z * f);
    }
    v3 M = se_inv_cov_diag(m, is_cov, is_iso, gs, ls);
    float absq = (ab.x * M.x) * ab.x + ((ab.y * M.y) * ab.y + (ab.z * M.z) * ab.z);
    float f = expf(-absq);
    return V4(f, -2.f * (ab.x * M.x) * f, -2.f * (ab.y * M.y) * f, -2.f * (ab.z * M.z) * f);
}
/* splattingKernel3DGrad: first + second gradient contracted with coeff (GPF:565-577, 810-833) */
static v4 cov_kernel3d_grad(const orc_medium *m, v3 pa, v3 pb, v3 c, int is_cov, int is_iso, float gs, v3 p_world) {
    float ls = cov_ns_scale(m, p_world);
    v3 ab = v3_sub(pa, pb);
    float G[9];
    int full = cov_aniso3d(m, p_world, G);
    if (full || (m->use_mtx && !is_iso)) { /* GPF:565-577, 819-833 with the full matrix */
        float Gi[9], Mf[9];
        if (full) { mat3_inverse(G, Gi); se_inv_cov_full(m, Gi, is_cov, is_iso, gs, ls, Mf); } else se_inv_cov_mtx(m, is_cov, gs, ls, Mf);
        /* dist2_ab = (ab^T M) ab and ab . M.col(j) (GPF.hpp:76-78, GPF:804-817), three-term sums as x0 + (x1 + x2) */
        v3 Mab = V3(ab.x * Mf[0] + (ab.y * Mf[3] + ab.z * Mf[6]), ab.x * Mf[1] + (ab.y * Mf[4] + ab.z * Mf[7]), ab.x * Mf[2] + (ab.y * Mf[5] + ab.z * Mf[8]));
        float f = expf(-(Mab.x * ab.x + (Mab.y * ab.y + Mab.z * ab.z)));
        float d[3] = {-2.f * Mab.x, -2.f * Mab.y, -2.f * Mab.z}, cc[3] = {c.x, c.y, c.z}, out[4] = {0.f, 0.f, 0.f, 0.f};
        for (int i = 0; i < 3; ++i) {
            out[0] += d[i] * f * cc[i];
            for (int j = 0; j < 3; ++j) out[1 + j] += (d[i] * d[j] - 2.f * Mf[i * 3 + j]) * f * cc[i];
        }
        return V4(out[0], out[1], out[2], out[3]);
    }
    v3 M = se_inv_cov_diag(m, is_cov, is_iso, gs, ls);
    float absq = (ab.x * M.x) * ab.x + ((ab.y * M.y) * ab.y + (ab.z * M.z) * ab.z);
    float f = expf(-absq);
    float dx = -2.f * (ab.x * M.x), dy = -2.f * (ab.y * M.y), dz = -2.f * (ab.z * M.z);
    v3 g1 = V3(dx * f, dy * f, dz * f);
    float h00 = (dx * dx - 2.f * M.x) * f, h11 = (dy * dy - 2.f * M.y) * f, h22 = (dz * dz - 2.f * M.z) * f;
    float h01 = (dx * dy - 2.f * 0.f) * f, h02 = (dx * dz - 2.f * 0.f) * f, h12 = (dy * dz - 2.f * 0.f) * f;
    v4 vx = V4(g1.x, h00, h01, h02), vy = V4(g1.y, h01, h11, h12), vz = V4(g1.z, h02, h12, h22);
    return v4_add(v4_add(v4_scale(vx, c.x), v4_scale(vy, c.y)), v4_scale(vz, c.z));
}
/* 1D kernels (GPF:835-865, GPF.hpp:1618-1640, GPF:1290-1332) */
static float se_k1d_val(float ab, float ls) { float den = (float)(2.0 * sqrf(ls)); return expf(-(ab * ab / den)); }
static float se_k1d_grad(float ab, float ls) { float f = se_k1d_val(ab, ls); float den = (float)(2.0 * sqrf(ls)); return (float)(-2.0 * ab / den * f); }
static float se_c1d_val(float ab, float ls) { return se_k1d_val(ab, (float)(ls * sqrt(2))); }
static float se_c1d_grad(float ab, float ls) { return se_k1d_grad(ab, (float)(ls * sqrt(2))); }
static float se_c1d_grad2(float ab, float ls) {
    float f = se_c1d_val(ab, ls); float den = (float)(4.0 * sqrf(ls)); float g1 = (float)(-2.0 * ab / den);
    return (float)((g1 * g1 - 2.0 / den) * f);
}
static float se_c1d_grad2_normal(float ab, float ls) { float f = se_c1d_val(ab, ls); float den = (float)(4.0 * sqrf(ls)); return (float)(-2.0 / den * f); }
/* splattingKernel1D (GPF:579-585) -> (val, grad) */
static void cov_kernel1d(const orc_medium *m, float pq, float pc, v3 pc_world, float *val, float *grad) {
    float ls = cov_ns_scale(m, pc_world);
    ls = (float)(ls * sqrt(1.0)); /* getNonstationaryAniso1D == 1 without an anisotropy field */
    if (m->stationary) ls = 1.0f; /* GPF.hpp:1618-1628 ignores localScale */
    *val = se_k1d_val(pq - pc, ls); *grad = se_k1d_grad(pq - pc, ls);
}
static float cov_c1d_ls(const orc_medium *m, v3 pqw, v3 pcw) { return m->stationary ? 1.0f : cov_splat_cov1d(m, pqw, pcw); }

/* ------------------------------------------------------------------ mean (GPF.hpp:867-1005, GaussianProcess.cpp:379-393) */
/* the material-test knob (math/SdfFunctions.cpp:66-179); expressions keep the reference's literal types (double literals promote) */
static const float SDF_PI = 3.1415926536f;
static float sdf_minf(float a, float b) { return a < b ? a : b; }
static float sdf_maxf(float a, float b) { return a > b ? a : b; }
static float sdf_len2(float x, float y) { float s = x * x; s += y * y; return sqrtf(s); }
static float sdf_len3(float x, float y, float z) { float s = x * x; s += y * y; s += z * z; return sqrtf(s); }
static void sdf_rotx(float deg, const float *b, float *o) { /* Mat4f::rotAxis((1,0,0), deg).transformVector (math/Mat4f.cpp:132-148, Mat4f.hpp:159-166) */
    float angle = deg * (SDF_PI / 180.0f), s = sinf(angle), c = cosf(angle), c1 = 1.0f - c, x = 1.f, y = 0.f, z = 0.f;
    float m[9] = {c + x * x * c1, x * y * c1 - z * s, x * z * c1 + y * s, y * x * c1 + z * s, c + y * y * c1, y * z * c1 - x * s, z * x * c1 - y * s, z * y * c1 + x * s, c + z * z * c1};
    for (int i = 0; i < 3; ++i) o[i] = m[3 * i] * b[0] + m[3 * i + 1] * b[1] + m[3 * i + 2] * b[2];
}
static float sdf_cone(const float *p, float cx, float cy) { float q = sdf_len2(p[0], p[1]); float s = cx * q; s += cy * p[2]; return s; }
static float sdf_capped_cylinder(const float *p, float h, float r) {
    float dx = fabsf(sdf_len2(p[0], p[2])) - h, dy = fabsf(p[1]) - r;
    return sdf_minf(sdf_maxf(dx, dy), 0.0f) + sdf_len2(sdf_maxf(dx, 0.0f), sdf_maxf(dy, 0.0f));
}
static float sdf_tri_prism(const float *p, float hx, float hy) {
    float qx = fabsf(p[0]), qz = fabsf(p[2]);
    return sdf_maxf(qz - hy, sdf_maxf(qx * 0.866025f + p[1] * 0.5f, -p[1]) - hx * 0.5f);
}
static float sdf_lerpf(float a, float b, float r) { return a * (1.0f - r) + b * r; }
static float sdf_smooth_union(float d1, float d2, float k) {
    float h = 0.5f + 0.5f * (d2 - d1) / k; h = sdf_minf(sdf_maxf(h, 0.0f), 1.0f);
    return sdf_lerpf(d2, d1, h) - k * h * (1.0 - h);
}
static float sdf_ssub(float d1, float d2, float k) {
    float h = clampd(0.5 - 0.5 * (d2 + d1) / k, 0.0, 1.0);
    return sdf_lerpf(d2, -d1, h) + k * h * (1.0 - h);
}
static float sdf_base(const float *p) {
    float a[3] = {p[0], p[1] + .9f, p[2]}, b[3] = {p[0], p[1] - .9f, p[2]}, ra[3], rb[3], c[3] = {p[0], p[1], p[2] + -1.f}, rc[3];
    sdf_rotx(-90, a, ra); sdf_rotx(90, b, rb);
    float base = sdf_smooth_union(sdf_cone(ra, (float)(SDF_PI / 3.), (float)(SDF_PI / 3.)), sdf_cone(rb, SDF_PI / 3.f, SDF_PI / 3.f), 0.02);
    base = sdf_maxf(base, sdf_capped_cylinder(p, 1.1f, 0.25f)) * 0.7f;
    base = sdf_maxf(-sdf_capped_cylinder(p, 0.6f, 0.3f), base);
    sdf_rotx(90, c, rc);
    base = sdf_maxf(-sdf_tri_prism(rc, 1.2f, 0.3f), base);
    return base;
}
static float sdf_knob(float x, float y, float z, int outer) {
    const float scale = 0.8f; float k = (float)(1. / scale);
    float p[3] = {x * k, y * k, z * k};
    float sphere = sdf_len3(p[0], p[1], p[2]) - 1.0f, cutout = sdf_len3(p[0], p[1] - 0.5f, p[2] - 0.5f) - 0.7f;
    float e[3] = {p[0], p[1] - 0.2f, p[2] - 0.2f}, te[3];
    sdf_rotx(-45, e, te);
    float etch = sdf_len2(sdf_len2(te[0], te[2]) - 1.0f, te[1]) - 0.05f;
    float inner = sdf_len3(p[0], p[1], p[2]) - 0.75f;
    float d = sdf_ssub(cutout, sphere, 0.1);
    d = outer ? sdf_maxf(d, -inner) : sdf_minf(d, inner);
    d = sdf_maxf(-etch, d);
    float q[3] = {p[0], p[1] - -.775f, p[2]};
    d = sdf_minf(sdf_ssub(sphere, sdf_base(q), 0.1), d);
    return d * scale;
}
/* ProceduralMean::mean (GPF:279-285) over SdfFunctions::eval (math/SdfFunctions.hpp:47-55, SdfFunctions.cpp:148-157, 181-183) or the scalar field "f" */
static double proc_mean(const scn_mean_desc *md, d3 a) {
    float px = (float)a.x, py = (float)a.y, pz = (float)a.z;
    const float *T = md->inv_transform;
    float x = T[0] * px + T[1] * py + T[2] * pz + T[3], y = T[4] * px + T[5] * py + T[6] * pz + T[7], z = T[8] * px + T[9] * py + T[10] * pz + T[11];
    float m;
    switch (md->sdf_func) {
    case SCN_SDF_PLANE: m = y; break;
    case SCN_SDF_TWO_SPHERES: {
        float y1 = y - 10.f, y2 = y - -10.f;
        float a1 = sqrtf(x * x + y1 * y1 + z * z) - 9.5f, a2 = sqrtf(x * x + y2 * y2 + z * z) - 9.5f;
        m = a1 < a2 ? a1 : a2; break;
    }
    case SCN_SDF_KNOB_INNER: {
        const float scale = 0.8f; float k = (float)(1. / scale);
        float qx = x * k, qy = y * k, qz = z * k;
        m = (sqrtf(qx * qx + qy * qy + qz * qz) - 0.75f) * scale; break;
    }
    case SCN_SDF_KNOB: m = sdf_knob(x, y, z, 0); break;
    case SCN_SDF_KNOB_OUTER: m = sdf_knob(x, y, z, 1); break;
    default: { d3 q = {x, y, z}; m = (float)field_scalar(&md->f, q); break; }
    }
    m *= md->scale;
    float r = m + md->offset;
    return md->min > r ? md->min : r;
}
static double mean_eval(const scn_mean_desc *md, d3 a) {
    switch (md->type) {
    case SCN_MEAN_PROCEDURAL: return proc_mean(md, a);
    case SCN_MEAN_HOMOGENEOUS: return md->offset;
    case SCN_MEAN_SPHERICAL: { d3 q = {a.x - md->center[0], a.y - md->center[1], a.z - md->center[2]}; return d3_len(q) - md->radius; }
    case SCN_MEAN_LINEAR: {
        d3 q = {a.x - md->ref[0], a.y - md->ref[1], a.z - md->ref[2]}, dir = {md->dir[0], md->dir[1], md->dir[2]};
        double v = d3_dot(q, dir) * md->scale, mn = (double)md->min; return v > mn ? v : mn;
    }
    default: return 0.;
    }
}
static d3 mean_grad(const scn_mean_desc *md, d3 a) {
    d3 r = {0., 0., 0.};
    switch (md->type) {
    case SCN_MEAN_PROCEDURAL: { /* ProceduralMean::dmean_da (GPF:287-296) */
        double eps = 0.001f;
        d3 ax = a, ay = a, az = a; ax.x += eps; ay.y += eps; az.z += eps;
        double v0 = proc_mean(md, a), v1 = proc_mean(md, ax), v2 = proc_mean(md, ay), v3 = proc_mean(md, az);
        r.x = (v1 - v0) / eps; r.y = (v2 - v0) / eps; r.z = (v3 - v0) / eps; return r;
    }
    case SCN_MEAN_SPHERICAL: { d3 q = {a.x - md->center[0], a.y - md->center[1], a.z - md->center[2]}; double inv = 1.0 / d3_len(q); r.x = q.x * inv; r.y = q.y * inv; r.z = q.z * inv; return r; }
    case SCN_MEAN_LINEAR: {
        d3 q = {a.x - md->ref[0], a.y - md->ref[1], a.z - md->ref[2]}, dir = {md->dir[0], md->dir[1], md->dir[2]};
        if (d3_dot(q, dir) * md->scale < md->min) return r;
        r.x = dir.x * md->scale; r.y = dir.y * md->scale; r.z = dir.z * md->scale; return r;
    }
    default: return r;
    }
}
static void gp_mean_and_id(const orc_medium *m, v3 p, double *mean, int *id) {
    d3 a = to_d3(p);
    *mean = mean_eval(&m->d.mean, a); *id = m->d.gp_id;
    if (m->d.has_mean_additional) { double v = mean_eval(&m->d.mean_additional, a); if (v < *mean) { *mean = v; *id = m->d.id_additional; } }
}
static v3 gp_mean_grad(const orc_medium *m, v3 p, int id) {
    d3 g = id == m->d.gp_id ? mean_grad(&m->d.mean, to_d3(p)) : mean_grad(&m->d.mean_additional, to_d3(p));
    return V3((float)g.x, (float)g.y, (float)g.z);
}

/* ------------------------------------------------------------------ realization (SCN.hpp:23-47, SCN:8-37) */
typedef struct { float value_scale; v3 gradient_scale; v3 ray_origin; } coeff_t;
typedef struct { uint32_t pix[4]; uint32_t scene_seed; float t; } info_t;
typedef struct {
    const orc_medium *m;
    int activate_conditioning, scheme;
    coeff_t c3, c1;
} real_t;

static void real_init(real_t *r, const orc_medium *m) {
    memset(r, 0, sizeof *r);
    r->m = m;
    int ctx = m->d.correlation_context;
    r->activate_conditioning = !m->d.single_realization && (ctx == SCN_CTX_RENEWAL || ctx == SCN_CTX_RENEWAL_PLUS);
    r->scheme = SCN_SCHEME_UNI;
    if (!m->d.single_realization && m->d.sampling_1d) r->scheme = m->d.sampling_scheme_1d;
}
static uint32_t real_seed(const real_t *r, const info_t *info) { /* SCN:40-49 */
    uint32_t seed = r->m->d.seed;
    if (!r->m->d.single_realization) {
        uint32_t w = r->m->d.correlation_context == SCN_CTX_GLOBAL ? 0u : info->pix[3];
        seed += orc_xxhash32_4(info->pix[0], info->pix[1], info->pix[2], w) + orc_xxhash32_1(info->scene_seed);
    }
    return seed;
}
static int real_scheme(const real_t *r, v3 p) { return cov_amplitude(r->m, p) < 1e-6f ? SCN_SCHEME_UNI : r->scheme; } /* SCN:51-58 */

/* kernelScaleLevelRatio (SCN:60-70) */
static v4 real_level_ratio(const real_t *r, v3 p) {
    const float base = 2.5f;
    float ls = cov_lateral_scale(r->m, p);
    float lo = floorf(logf(ls) / logf(base)), hi = lo + 1;
    float s_lo = powf(base, lo), s_hi = powf(base, hi);
    float frac = (ls - s_lo) / (s_hi - s_lo);
    float r_lo = (float)((1.0 - frac) / sqrt(1.0 - 2.0 * frac + 2.0 * frac * frac));
    float r_hi = (float)(frac / sqrt(1.0 - 2.0 * frac + 2.0 * frac * frac));
    return V4(s_lo, s_hi, r_lo, r_hi);
}
static inline int additional_seed(float kss) { return (int)floorf(logf(kss) / logf(2.5f)); } /* SCN:252,272,300,328 */

static __thread uint32_t g_accept_count; /* impulses that passed |p - p_i|^2 < 1 since the last reset (test bookkeeping) */
/* cell3D / noise3D (SCN:362-395) */
static v4 real_cell3d(const real_t *r, v3 p_world, uint32_t ci, uint32_t cj, uint32_t ck, v3 p, uint32_t seed, float R, float kss) {
    pcg_t s; pcg_set_state(&s, (uint64_t)(uint32_t)(orc_xxhash32_4(ck, cj, ci, seed) + 1u));
    uint32_t n = r->m->n_imp;
    v4 sum = V4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t k = 0; k < n; ++k) {
        v3 pi = pcg_next3d(&s);
        float w = pcg_next1d(&s) < 0.5f ? -1.f : 1.f;
        v3 tp = v3_sub(p, pi);
        if (v3_lensq(tp) < 1.0) g_accept_count++;
        if (v3_lensq(tp) < 1.0) sum = v4_add(sum, v4_scale(cov_kernel3d(r->m, v3_scale(p, R), v3_scale(pi, R), 0, r->m->d.isotropic_3d_sampling, kss, p_world), w));
    }
    return sum;
}
static v4 real_noise3d(const real_t *r, v3 p_world, v3 p, uint32_t seed, float R, float kss) {
    v3 pg = v3_div(p, R);
    v3 fl = V3(floorf(pg.x), floorf(pg.y), floorf(pg.z));
    v3 frac = v3_sub(pg, fl);
    int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z;
    v4 sum = V4(0.f, 0.f, 0.f, 0.f);
    for (int dx = -1; dx <= 1; ++dx) for (int dy = -1; dy <= 1; ++dy) for (int dz = -1; dz <= 1; ++dz)
        sum = v4_add(sum, real_cell3d(r, p_world, (uint32_t)(ix + dx), (uint32_t)(iy + dy), (uint32_t)(iz + dz),
                                      v3_sub(frac, V3((float)dx, (float)dy, (float)dz)), seed, R, kss));
    return sum;
}
/* cell1D / noise1D (SCN:397-428) -> (val, grad) */
static void real_noise1d(const real_t *r, v3 p_world, float t, uint32_t seed, float R, float *val, float *grad) {
    float tg = t / R; float fl = floorf(tg); float frac = tg - fl; int i = (int)fl;
    float sv = 0.f, sg = 0.f; uint32_t n = r->m->n_imp;
    for (int dx = -1; dx <= 1; ++dx) {
        float tc = frac - dx; float cv = 0.f, cg = 0.f;
        pcg_t s; pcg_set_state(&s, (uint64_t)(uint32_t)(orc_xxhash32_2((uint32_t)(i + dx), seed) + 1u));
        for (uint32_t k = 0; k < n; ++k) {
            float ti = pcg_next1d(&s); float w = pcg_next1d(&s) < 0.5f ? -1.f : 1.f;
            float tp = tc - ti;
            if (tp * tp < 1.0) { float kv, kg; cov_kernel1d(r->m, R * tc, R * ti, p_world, &kv, &kg); cv += w * kv; cg += w * kg; }
        }
        sv += cv; sg += cg;
    }
    *val = sv; *grad = sg;
}

/* evaluateNoise3DNormalized — world space (SCN:251-260) */
static v4 real_noise3d_world(const real_t *r, v3 p, uint32_t seed, float R, float kss, int conditioning) {
    const orc_medium *m = r->m;
    int add = additional_seed(kss);
    v4 noise = real_noise3d(r, p, p, seed + (uint32_t)add, R, kss);
    float nf = sqrtf(cov_variance3d(m, p, m->d.impulse_density, R, 0, kss));
    noise = v4_div(noise, nf);
    if (r->activate_conditioning && conditioning) {
        v4 a = v4_scale(cov_kernel3d(m, p, r->c3.ray_origin, 1, 0, kss, p), r->c3.value_scale);
        v4 b = cov_kernel3d_grad(m, p, r->c3.ray_origin, r->c3.gradient_scale, 1, 0, kss, p);
        noise = v4_add(noise, v4_add(a, b));
    }
    return noise;
}
/* evaluateNoise3DIsotropicRayNormalized (SCN:291-320); _isotropicRaySpace3DSampling is always true (SCN:17) */
static v4 real_noise3d_isoray(const real_t *r, v3 p, v3 dir, uint32_t seed, float R, float kss, int conditioning) {
    const orc_medium *m = r->m;
    v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
    frame_t fr = frame_make(dir_iso);
    v3 p_iso = cov_pos_w2l(m, p, kss);
    v3 p_ray = frame_to_local(&fr, p_iso);
    int add = additional_seed(kss);
    v4 nr = real_noise3d(r, p, p_ray, seed + (uint32_t)add, R, 1.0f);
    v3 g_iso = frame_to_global(&fr, V3(nr.y, nr.z, nr.w));
    v3 g_w = cov_grad_l2w(m, g_iso, kss);
    v4 nw = V4(nr.x, g_w.x, g_w.y, g_w.z);
    float nf = sqrtf(cov_variance3d(m, p, m->d.impulse_density, R, 1, 1.0f));
    nw = v4_div(nw, nf);
    if (r->activate_conditioning && conditioning) {
        v3 o_iso = cov_pos_w2l(m, r->c3.ray_origin, kss);
        v3 o_ray = frame_to_local(&fr, o_iso);
        v4 a = v4_scale(cov_kernel3d(m, p_ray, o_ray, 1, 1, 1.0f, p), r->c3.value_scale);
        v4 b = cov_kernel3d_grad(m, p_ray, o_ray, r->c3.gradient_scale, 1, 1, 1.0f, p);
        v4 d = v4_add(a, b);
        v3 gd_iso = frame_to_global(&fr, V3(d.y, d.z, d.w));
        v3 gd_w = cov_grad_l2w(m, gd_iso, kss);
        nw = v4_add(nw, V4(d.x, gd_w.x, gd_w.y, gd_w.z));
    }
    return nw;
}
/* evaluateNoise3D (SCN:101-124) */
static v4 real_eval_noise3d(const real_t *r, v3 p, v3 dir, uint32_t seed, int conditioning) {
    const orc_medium *m = r->m;
    if (!m->d.isotropic_3d_sampling) {
        if (!m->multires) return real_noise3d_world(r, p, seed, cov_radius(m, 0, 1.0f), cov_world_scale(m), conditioning);
        v4 info = real_level_ratio(r, p);
        v4 lo = real_noise3d_world(r, p, seed, cov_radius(m, 0, info.x), info.x, conditioning);
        v4 hi = real_noise3d_world(r, p, seed, cov_radius(m, 0, info.y), info.y, conditioning);
        return v4_add(v4_scale(lo, info.z), v4_scale(hi, info.w));
    }
    if (!m->multires) return real_noise3d_isoray(r, p, dir, seed, cov_radius(m, 1, 1.0f), 1.0f, conditioning);
    v4 info = real_level_ratio(r, p);
    v4 lo = real_noise3d_isoray(r, p, dir, seed, cov_radius(m, 1, 1.0f), info.x, conditioning);
    v4 hi = real_noise3d_isoray(r, p, dir, seed, cov_radius(m, 1, 1.0f), info.y, conditioning);
    return v4_add(v4_scale(lo, info.z), v4_scale(hi, info.w));
}
/* evaluateNoise1DNormalized (SCN:322-360) -> (val, gx_cond, gy_cond, gz) */
static v4 real_noise1d_norm(const real_t *r, v3 p, v3 dir, uint32_t seed, float R, float kss, int conditioning, int low_level) {
    const orc_medium *m = r->m;
    v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
    frame_t fr = frame_make(dir_iso);
    v3 p_iso = cov_pos_w2l(m, p, kss);
    v3 p_ray = frame_to_local(&fr, p_iso);
    int add = additional_seed(kss);
    float nv, ng;
    real_noise1d(r, p, p_ray.z, seed + (uint32_t)add, R, &nv, &ng);
    float nf = sqrtf(cov_variance1d(m, p, m->d.impulse_density, R));
    nv /= nf; ng /= nf;
    float gcx = 0.f, gcy = 0.f;
    if (r->activate_conditioning && conditioning) {
        float osf = 1.0f;
        if (m->multires) { v4 oi = real_level_ratio(r, r->c1.ray_origin); osf = low_level ? oi.z : oi.w; }
        v3 o_iso = cov_pos_w2l(m, r->c1.ray_origin, kss);
        v3 o_ray = frame_to_local(&fr, o_iso);
        float ab = p_ray.z - o_ray.z;
        float cls = cov_c1d_ls(m, p, r->c1.ray_origin);
        float vs = r->c1.value_scale;
        float v_val = vs * se_c1d_val(ab, cls), v_grad = vs * se_c1d_grad(ab, cls);
        float gzc = kss * r->c1.gradient_scale.z;
        float g_val = gzc * se_c1d_grad(ab, cls), g_grad = gzc * se_c1d_grad2(ab, cls);
        nv += osf * (v_val + g_val); ng += osf * (v_grad + g_grad);
        if (m->d.correlation_xy) {
            gcx = osf * kss * r->c1.gradient_scale.x * se_c1d_grad2_normal(ab, cls);
            gcy = osf * kss * r->c1.gradient_scale.y * se_c1d_grad2_normal(ab, cls);
        }
    }
    return V4(nv, gcx, gcy, ng);
}
/* evaluateValueNoise1D (SCN:126-135) */
static float real_value_noise1d(const real_t *r, v3 p, v3 dir, uint32_t seed, int conditioning) {
    const orc_medium *m = r->m; float R = cov_radius(m, 1, 1.0f);
    if (!m->multires) return real_noise1d_norm(r, p, dir, seed, R, 1.0f, conditioning, 0).x;
    v4 info = real_level_ratio(r, p);
    float lo = real_noise1d_norm(r, p, dir, seed, R, info.x, conditioning, 1).x;
    float hi = real_noise1d_norm(r, p, dir, seed, R, info.y, conditioning, 0).x;
    return info.z * lo + info.w * hi;
}
/* evaluateGradientNoise1D (SCN:137-239) */
static v3 real_gradient_noise1d(const real_t *r, v3 p, float t_segment, v3 dir, const info_t *info, int conditioning) {
    const orc_medium *m = r->m;
    uint32_t seed = real_seed(r, info);
    v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
    frame_t fr = frame_make(dir_iso);
    float scale_x = cov_splat_cov1d(m, p, p), scale_y = cov_splat_cov1d(m, p, p);
    double xs0 = 1.0 / scale_x, xs1 = 1.0 / scale_y;
    float R = cov_radius(m, 1, 1.0f);
    const float sqrt2f = sqrtf(2.f);
    int plus_xy = m->d.correlation_context == SCN_CTX_RENEWAL_PLUS && m->d.correlation_xy;
    pcg_t s;
    uint32_t xyseed = orc_xxhash32_4(info->pix[0], info->pix[1], info->pix[2], info->pix[3]) + orc_xxhash32_1(info->scene_seed) + 1u;
    if (!m->multires) {
        v4 nb = real_noise1d_norm(r, p, dir, seed, R, 1.0f, conditioning, 0);
        v3 base = V3(nb.y, nb.z, nb.w);
        pcg_set_state(&s, (uint64_t)xyseed);
        double a0, a1; rand_normal_2(&s, &a0, &a1); a0 /= sqrt2f; a1 /= sqrt2f;
        v3 g;
        if (t_segment == 0) { a0 *= xs0; a1 *= xs1; g = v3_add(V3((float)a0, (float)a1, 0.f), base); }
        else if (!plus_xy) {
            double b0, b1; rand_normal_2(&s, &b0, &b1); b0 /= sqrt2f; b1 /= sqrt2f; b0 *= xs0; b1 *= xs1;
            g = v3_add(V3((float)b0, (float)b1, 0.f), base);
        } else {
            float factor = (float)(exp(-t_segment * t_segment / 4.0) * (0.5 - t_segment * t_segment * 0.25));
            double mu0 = factor * 2.0 * a0, mu1 = factor * 2.0 * a1;
            float cov = (float)(0.5 - factor * factor * 2);
            double b0, b1; rand_normal_2(&s, &b0, &b1); float sc = sqrtf(cov);
            b0 = b0 * sc + mu0; b1 = b1 * sc + mu1;
            g = v3_add(V3((float)b0, (float)b1, 0.f), base);
        }
        v3 g_iso = frame_to_global(&fr, g);
        return cov_grad_l2w(m, g_iso, 1.0f);
    }
    v4 si = real_level_ratio(r, p);
    v4 nlo = real_noise1d_norm(r, p, dir, seed, R, si.x, conditioning, 1), nhi = real_noise1d_norm(r, p, dir, seed, R, si.y, conditioning, 0);
    v3 blo = V3(nlo.y, nlo.z, nlo.w), bhi = V3(nhi.y, nhi.z, nhi.w);
    pcg_set_state(&s, (uint64_t)xyseed);
    double l0, l1, h0, h1; rand_normal_2(&s, &l0, &l1); l0 /= sqrt2f; l1 /= sqrt2f; rand_normal_2(&s, &h0, &h1); h0 /= sqrt2f; h1 /= sqrt2f;
    v3 glo, ghi;
    if (t_segment == 0) {
        l0 *= xs0; l1 *= xs1; h0 *= xs0; h1 *= xs1;
        glo = v3_add(V3((float)l0, (float)l1, 0.f), blo); ghi = v3_add(V3((float)h0, (float)h1, 0.f), bhi);
    } else if (!plus_xy) {
        double a0, a1, b0, b1; rand_normal_2(&s, &a0, &a1); a0 /= sqrt2f; a1 /= sqrt2f; rand_normal_2(&s, &b0, &b1); b0 /= sqrt2f; b1 /= sqrt2f;
        a0 *= xs0; a1 *= xs1; b0 *= xs0; b1 *= xs1;
        glo = v3_add(V3((float)a0, (float)a1, 0.f), blo); ghi = v3_add(V3((float)b0, (float)b1, 0.f), bhi);
    } else {
        float factor = (float)(exp(-t_segment * t_segment / 4.0) * (0.5 - t_segment * t_segment * 0.25));
        double ml0 = factor * 2.0 * l0, ml1 = factor * 2.0 * l1, mh0 = factor * 2.0 * h0, mh1 = factor * 2.0 * h1;
        float cov = (float)(0.5 - factor * factor * 2); float sc = sqrtf(cov);
        double a0, a1, b0, b1; rand_normal_2(&s, &a0, &a1); rand_normal_2(&s, &b0, &b1);
        a0 = a0 * sc + ml0; a1 = a1 * sc + ml1; b0 = b0 * sc + mh0; b1 = b1 * sc + mh1;
        glo = v3_add(V3((float)a0, (float)a1, 0.f), blo); ghi = v3_add(V3((float)b0, (float)b1, 0.f), bhi);
    }
    frame_t flo = frame_make(v3_normalized(cov_pos_w2l(m, dir, si.x))), fhi = frame_make(v3_normalized(cov_pos_w2l(m, dir, si.y)));
    v3 gwlo = cov_grad_l2w(m, frame_to_global(&flo, glo), si.x), gwhi = cov_grad_l2w(m, frame_to_global(&fhi, ghi), si.y);
    return v3_add(v3_scale(gwlo, si.z), v3_scale(gwhi, si.w));
}

/* evaluateValue / evaluateGradient (SCN:73-99) */
static float real_value(const real_t *r, v3 p, v3 dir, const info_t *info, int *gp_id) {
    const orc_medium *m = r->m;
    g_eval_count++;
    uint32_t seed = real_seed(r, info);
    float amp = cov_amplitude(m, p);
    float nv = m->d.sampling_1d ? real_value_noise1d(r, p, dir, seed, 1) : real_eval_noise3d(r, p, dir, seed, 1).x;
    double mean; int id; gp_mean_and_id(m, p, &mean, &id);
    if (m->d.surf_vol_phase_separate) id = 1.f < m->d.surf_vol_phase_amp_thresh ? 0 : 1;
    *gp_id = id;
    return (float)(amp * nv + mean);
}
static v3 real_gradient(const real_t *r, v3 p, float t, v3 dir, const info_t *info) {
    const orc_medium *m = r->m;
    g_eval_count++;
    uint32_t seed = real_seed(r, info);
    float amp = cov_amplitude(m, p);
    v3 ng;
    if (m->d.sampling_1d) ng = real_gradient_noise1d(r, p, t, dir, info, 1);
    else { v4 n = real_eval_noise3d(r, p, dir, seed, 1); ng = V3(n.y, n.z, n.w); }
    double mean; int id; gp_mean_and_id(m, p, &mean, &id);
    v3 mg = gp_mean_grad(m, p, id);
    return v3_add(v3_scale(ng, amp), mg);
}

/* Eigen's fixed-size 3x3 inverse (cofactor form), used by sparseConvNoiseOneOverSecondDerivative and neeShared */
static void mat3_inverse(const float a[9], float r[9]) { /* row-major */
#define A(i, j) a[((i) % 3) * 3 + ((j) % 3)]
#define COF(i, j) (A(i + 1, j + 1) * A(i + 2, j + 2) - A(i + 1, j + 2) * A(i + 2, j + 1))
    float c00 = COF(0, 0), c10 = COF(1, 0), c20 = COF(2, 0);
    float det = c00 * a[0] + (c10 * a[3] + c20 * a[6]);
    float inv = 1.0f / det;
    r[0] = c00 * inv; r[1] = c10 * inv; r[2] = c20 * inv;
    r[3] = COF(0, 1) * inv; r[4] = COF(1, 1) * inv; r[5] = COF(2, 1) * inv;
    r[6] = COF(0, 2) * inv; r[7] = COF(1, 2) * inv; r[8] = COF(2, 2) * inv;
#undef COF
#undef A
}
/* sparseConvNoiseOneOverSecondDerivative (GPF:713-739), diagonal case, no anisotropy field */
static v3 cov_one_over_second_derivative_diag(const orc_medium *m, int is_iso) {
    float a[9] = {1.f, 0, 0, 0, 1.f, 0, 0, 0, 1.f};
    if (!is_iso) { a[0] = m->l_aniso_inv.x * m->l_aniso_inv.x; a[4] = m->l_aniso_inv.y * m->l_aniso_inv.y; a[8] = m->l_aniso_inv.z * m->l_aniso_inv.z; }
    for (int i = 0; i < 9; ++i) a[i] *= 0.25f;
    for (int i = 0; i < 9; ++i) a[i] = -2.f * a[i];
    float r[9]; mat3_inverse(a, r);
    return V3(r[0], r[4], r[8]);
}

/* sparseConvNoiseOneOverSecondDerivative with the anisotropy field (GPF:1239-1243, 713-739): (-2 * 0.25 * invSigma^T G^-1 invSigma)^-1 applied to v */
static v3 cov_one_over_second_derivative_apply(const orc_medium *m, v3 p_world, int is_iso, v3 v) {
    float G[9];
    if (m->use_mtx && !is_iso) { /* GPF:728-738 with _cov_mtx_inv */
        float a[9], r[9];
        for (int i = 0; i < 9; ++i) a[i] = -2.f * (m->C[i] * 0.25f);
        mat3_inverse(a, r);
        return mat3_mulv_e(r, v);
    }
    if (!cov_aniso3d(m, p_world, G)) return v3_mul(cov_one_over_second_derivative_diag(m, is_iso), v);
    float Gi[9], a[9], r[9]; mat3_inverse(G, Gi);
    float sg[3] = {1.f, 1.f, 1.f};
    if (!is_iso) { sg[0] = m->l_aniso_inv.x; sg[1] = m->l_aniso_inv.y; sg[2] = m->l_aniso_inv.z; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i * 3 + j] = -2.f * (((sg[i] * Gi[i * 3 + j]) * sg[j]) * 0.25f);
    mat3_inverse(a, r);
    return V3(r[0] * v.x + (r[1] * v.y + r[2] * v.z), r[3] * v.x + (r[4] * v.y + r[5] * v.z), r[6] * v.x + (r[7] * v.y + r[8] * v.z));
}

/* conditioning3D (SCN:441-517); the reference's sanity re-evaluations only print and are left out */
static void real_conditioning3d(real_t *r, v3 p, v3 dir, float target_val, v3 target_grad, const info_t *info) {
    const orc_medium *m = r->m; int ctx = m->d.correlation_context;
    r->c3.value_scale = 0.f; r->c3.gradient_scale = V3(0, 0, 0); r->c3.ray_origin = p;
    v4 ksi = V4(0, 0, 0, 0);
    if (m->multires) ksi = real_level_ratio(r, p);
    uint32_t seed = real_seed(r, info);
    float amp = 0.f; double mean = 0.; int id = 0;
    if (ctx == SCN_CTX_RENEWAL || ctx == SCN_CTX_RENEWAL_PLUS) {
        amp = cov_amplitude(m, p);
        if (amp == 0) return;
        gp_mean_and_id(m, p, &mean, &id);
        float meanf = (float)mean;
        float cur = real_eval_noise3d(r, p, dir, seed, 0).x;
        r->c3.value_scale = (target_val - meanf) / amp - cur;
        if (m->multires) r->c3.value_scale /= ksi.z + ksi.w;
    }
    if (ctx == SCN_CTX_RENEWAL_PLUS) {
        v3 mg = gp_mean_grad(m, p, id);
        v4 n = real_eval_noise3d(r, p, dir, seed, 0);
        v3 cur = V3(n.y, n.z, n.w);
        v3 delta = v3_sub(v3_div(v3_sub(target_grad, mg), amp), cur);
        if (m->d.isotropic_3d_sampling) {
            v3 gs = cov_grad_w2l(m, delta, 1.0f);
            v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
            frame_t fr = frame_make(dir_iso);
            gs = frame_to_local(&fr, gs);
            r->c3.gradient_scale = cov_one_over_second_derivative_apply(m, p, 1, gs);
            if (m->multires) r->c3.gradient_scale = v3_div(r->c3.gradient_scale, ksi.z / ksi.x + ksi.w / ksi.y);
            else r->c3.gradient_scale = v3_scale(r->c3.gradient_scale, sqrf(cov_ns_scale(m, p)));
        } else {
            r->c3.gradient_scale = cov_one_over_second_derivative_apply(m, p, 0, delta);
            if (m->multires) r->c3.gradient_scale = v3_div(r->c3.gradient_scale, ksi.z / (ksi.x * ksi.x) + ksi.w / (ksi.y * ksi.y));
            else r->c3.gradient_scale = v3_scale(r->c3.gradient_scale, sqrf(cov_lateral_scale(m, p)));
        }
    }
}
/* conditioning1D (SCN:519-595) */
static void real_conditioning1d(real_t *r, v3 p, v3 dir, float target_val, v3 target_grad, const info_t *info) {
    const orc_medium *m = r->m; int ctx = m->d.correlation_context;
    r->c1.value_scale = 0.f; r->c1.gradient_scale = V3(0, 0, 0); r->c1.ray_origin = p;
    v4 ksi = V4(0, 0, 0, 0);
    if (m->multires) ksi = real_level_ratio(r, p);
    uint32_t seed = real_seed(r, info);
    float amp = 0.f; double mean = 0.; int id = 0;
    if (ctx == SCN_CTX_RENEWAL || ctx == SCN_CTX_RENEWAL_PLUS) {
        amp = cov_amplitude(m, p);
        if (amp == 0) return;
        gp_mean_and_id(m, p, &mean, &id);
        float meanf = (float)mean;
        float cur = real_value_noise1d(r, p, dir, seed, 0);
        r->c1.value_scale = (target_val - meanf) / amp - cur;
        if (m->multires) r->c1.value_scale /= sqrf(ksi.z) + sqrf(ksi.w);
    }
    if (ctx == SCN_CTX_RENEWAL_PLUS) {
        v3 mg = gp_mean_grad(m, p, id);
        v3 cur = real_gradient_noise1d(r, p, 0.f, dir, info, 0);
        v3 delta = v3_sub(v3_div(v3_sub(target_grad, mg), amp), cur);
        v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
        frame_t fr = frame_make(dir_iso);
        v3 gs = frame_to_local(&fr, cov_grad_w2l(m, delta, 1.0f));
        float sz = cov_splat_cov1d(m, p, p), sx = sz, sy = sz;
        r->c1.gradient_scale = v3_mul(v3_scale(gs, -2.f), V3(sqrf(sx), sqrf(sy), sqrf(sz)));
        if (m->multires) r->c1.gradient_scale = v3_div(r->c1.gradient_scale, sqrf(ksi.z) + sqrf(ksi.w));
        if (!m->d.correlation_xy) { r->c1.gradient_scale.x = 0; r->c1.gradient_scale.y = 0; }
    }
}
static void real_conditioning(real_t *r, v3 p, v3 dir, float tv, v3 tg, const info_t *info) { /* SCN:431-439 */
    if (!r->activate_conditioning) return;
    if (r->m->d.sampling_1d) real_conditioning1d(r, p, dir, tv, tg, info); else real_conditioning3d(r, p, dir, tv, tg, info);
}

/* neeShared / neeGrad / neePDF (SCN:601-743) */
typedef struct { float mtx[9]; v3 gcon, g3, plane_n; float jac, dist; } nee_t;
static v3 mat3_mulv(const float a[9], v3 b) {
    return V3(a[0] * b.x + a[1] * b.y + a[2] * b.z, a[3] * b.x + a[4] * b.y + a[5] * b.z, a[6] * b.x + a[7] * b.y + a[8] * b.z);
}
static void real_nee_shared(const real_t *r, v3 dir, v3 normal, v3 p, const info_t *info, nee_t *o) {
    const orc_medium *m = r->m;
    uint32_t seed = real_seed(r, info);
    float amp = cov_amplitude(m, p);
    double mean; int id; gp_mean_and_id(m, p, &mean, &id);
    v3 mg = gp_mean_grad(m, p, id);
    v3 dir_iso = v3_normalized(cov_pos_w2l(m, dir, 1.0f));
    frame_t fr = frame_make(dir_iso);
    float rc[9] = {fr.t.x, fr.b.x, fr.n.x, fr.t.y, fr.b.y, fr.n.y, fr.t.z, fr.b.z, fr.n.z};
    float sc = m->stationary ? 1.0f : (m->multires ? 1.0f : m->ls_max); /* GPF:1286-1289 */
    v3 w = m->stationary ? m->l_aniso_inv : v3_div(m->l_aniso_inv, sc);
    float wd[3] = {amp * w.x, amp * w.y, amp * w.z};
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) o->mtx[3 * i + j] = wd[i] * rc[3 * i + j];
    if (m->use_mtx) { /* amplitude * _world_to_local_transpose (/ scale) * mtx_ray_coord (SCN:618, GPF:649, 1281-1284) */
        float wt[9];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) wt[i * 3 + j] = amp * (m->stationary ? m->W[j * 3 + i] : m->W[j * 3 + i] / sc);
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) o->mtx[3 * i + j] = wt[i * 3 + 0] * rc[0 * 3 + j] + (wt[i * 3 + 1] * rc[1 * 3 + j] + wt[i * 3 + 2] * rc[2 * 3 + j]);
    }
    float inv[9]; mat3_inverse(o->mtx, inv);
    float R = cov_radius(m, 1, 1.0f);
    if (!m->multires) { v4 n = real_noise1d_norm(r, p, dir, seed, R, 1.0f, 1, 0); o->gcon = V3(n.y, n.z, n.w); }
    else {
        v4 mi = real_level_ratio(r, p);
        v4 lo = real_noise1d_norm(r, p, dir, seed, R, mi.x, 1, 1), hi = real_noise1d_norm(r, p, dir, seed, R, mi.y, 1, 0);
        v3 a = v3_div(v3_scale(V3(lo.y, lo.z, lo.w), mi.z), mi.x), b = v3_div(v3_scale(V3(hi.y, hi.z, hi.w), mi.w), mi.y);
        o->gcon = v3_add(a, b);
    }
    v3 plane_pt = v3_add(mat3_mulv(o->mtx, V3(0.f, 0.f, o->gcon.z)), mg);
    v3 ax = mat3_mulv(o->mtx, V3(1.f, 0.f, 0.f)), ay = mat3_mulv(o->mtx, V3(0.f, 1.f, 0.f));
    v3 cr = v3_cross(ax, ay);
    o->plane_n = v3_normalized(cr);
    o->jac = v3_len(cr);
    o->dist = v3_dot(plane_pt, o->plane_n) / v3_dot(normal, o->plane_n);
    v3 g3w = v3_scale(normal, o->dist);
    o->g3 = mat3_mulv(inv, v3_sub(g3w, mg));
}
static v3 real_nee_grad(const real_t *r, v3 dir, v3 normal, v3 p, const info_t *info) {
    nee_t o; real_nee_shared(r, dir, normal, p, info, &o); return v3_scale(normal, o.dist);
}
static float real_nee_pdf(const real_t *r, v3 dir, v3 normal, v3 p, float t_segment, const info_t *info) {
    const orc_medium *m = r->m;
    nee_t o; real_nee_shared(r, dir, normal, p, info, &o);
    if (o.dist < 0.0) return 0.0f;
    float sx = cov_splat_cov1d(m, p, p), sy = sx;
    int plus_xy = m->d.correlation_context == SCN_CTX_RENEWAL_PLUS && m->d.correlation_xy;
    float sample_x, sample_y, pdf_area;
    pcg_t s;
    uint32_t xyseed = orc_xxhash32_4(info->pix[0], info->pix[1], info->pix[2], info->pix[3]) + orc_xxhash32_1(info->scene_seed) + 1u;
    const float sqrt2f = sqrtf(2.f);
    if (!m->multires) {
        if (!plus_xy) {
            sample_x = (float)((o.g3.x - o.gcon.x) * sqrt(2.0) * sx);
            sample_y = (float)((o.g3.y - o.gcon.y) * sqrt(2.0) * sy);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / (2.0 * M_PI) * 2.0 * sx * sy);
        } else {
            pcg_set_state(&s, (uint64_t)xyseed);
            double a0, a1; rand_normal_2(&s, &a0, &a1); a0 /= sqrt2f; a1 /= sqrt2f;
            float factor = (float)(exp(-t_segment * t_segment / 4.0) * (0.5 - t_segment * t_segment * 0.25));
            double mu0 = factor * 2.0 * a0, mu1 = factor * 2.0 * a1;
            float cov = (float)(0.5 - factor * factor * 2);
            sample_x = (float)((o.g3.x - o.gcon.x - mu0) / sqrtf(cov));
            sample_y = (float)((o.g3.y - o.gcon.y - mu1) / sqrtf(cov));
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / (2.0 * M_PI) / cov);
        }
    } else {
        v4 mi = real_level_ratio(r, p);
        float nss = (float)(1.0 / sqrtf(sqrf(mi.z / mi.x) + sqrf(mi.w / mi.y)));
        sx *= nss; sy *= nss;
        if (!plus_xy) {
            sample_x = (float)((o.g3.x - o.gcon.x) * sqrt(2.0) * sx);
            sample_y = (float)((o.g3.y - o.gcon.y) * sqrt(2.0) * sy);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / (2.0 * M_PI) * 2.0 * sx * sy);
        } else {
            pcg_set_state(&s, (uint64_t)xyseed);
            double a0, a1, b0, b1; rand_normal_2(&s, &a0, &a1); a0 /= sqrt2f; a1 /= sqrt2f; rand_normal_2(&s, &b0, &b1); b0 /= sqrt2f; b1 /= sqrt2f;
            float factor = (float)(exp(-t_segment * t_segment / 4.0) * (0.5 - t_segment * t_segment * 0.25));
            double mu0 = factor * 2.0 * (a0 * mi.z / mi.x + b0 * mi.w / mi.y), mu1 = factor * 2.0 * (a1 * mi.z / mi.x + b1 * mi.w / mi.y);
            float cov = (float)(0.5 - factor * factor * 2);
            sample_x = (float)((o.g3.x - o.gcon.x - mu0) / sqrtf(cov) * nss);
            sample_y = (float)((o.g3.y - o.gcon.y - mu1) / sqrtf(cov) * nss);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / (2.0 * M_PI * cov) * sqrf(nss));
        }
    }
    float cos_l = fabsf(v3_dot(normal, o.plane_n));
    float pdf_normal = pdf_area * (o.dist * o.dist) / cos_l / o.jac;
    v3 nd = V3(-dir.x, -dir.y, -dir.z);
    return (float)(pdf_normal / (4.0 * v3_dot(normal, nd)));
}

/* ------------------------------------------------------------------ medium query (SCNM:93-183, GPM:221-393) */
typedef struct {
    int first_scatter, bounce, last_gp_id, scheme, has_ctx;
    d3 last_aniso; float last_val; info_t info; real_t ctx;
} state_t;

/* intersectGP (SCNM:102-183) */
static int medium_intersect(const orc_medium *m, v3 pos, v3 dir, float near_t, float far_t, float u, state_t *st, double *t_out) {
    real_t noise; real_init(&noise, m);
    d3 rd = to_d3(dir), p = to_d3(pos);
    float step = (far_t - near_t) / (float)m->d.min_step;
    if (m->d.step_size < step) step = m->d.step_size;
    int gp_id = 0;
    double t = near_t;
    if (!st->first_scatter) {
        double tv = st->last_val;
        real_conditioning(&noise, V3((float)p.x, (float)p.y, (float)p.z), dir, (float)tv,
                          V3((float)st->last_aniso.x, (float)st->last_aniso.y, (float)st->last_aniso.z), &st->info);
    }
#define AT(tt) V3((float)(p.x + (tt) * rd.x), (float)(p.y + (tt) * rd.y), (float)(p.z + (tt) * rd.z))
    double f0 = real_value(&noise, AT(t), dir, &st->info, &gp_id);
    int sign0 = f0 < 0 ? -1 : 1;
    double pf = f0;
    t = near_t + step * u; /* float arithmetic, then widened (SCNM:129) */
    int nstep = 0;
    while (t < far_t) {
        nstep++;
        double fc = real_value(&noise, AT(t), dir, &st->info, &gp_id);
        int signc = fc < 0 ? -1 : 1;
        if (!st->first_scatter && nstep == 1) sign0 = signc;
        else if (signc != sign0) {
            double intp = pf / (pf - fc), t_test, t_prev;
            t_prev = (t - step) * (1.0 - intp) + t * intp;
            for (;;) {
                t_test = (t - step) * (1.0 - intp) + t * intp;
                double ft = real_value(&noise, AT(t_test), dir, &st->info, &gp_id);
                int sign_t = ft < 0 ? -1 : 1;
                if (sign_t == sign0) break;
                intp *= 0.9;
                if (intp <= 0.01) { t_prev = t_test = 0; break; }
                t_prev = t_test;
            }
            t = t_prev;
            st->last_gp_id = gp_id;
            st->scheme = real_scheme(&noise, AT(t));
            st->ctx = noise; st->has_ctx = 1;
            st->last_val = 0.0f;
            *t_out = t;
            return 1;
        }
        pf = fc;
        t += step;
    }
    t = far_t;
    st->scheme = SCN_SCHEME_UNI;
    st->ctx = noise; st->has_ctx = 1;
    st->last_val = real_value(&noise, AT(t), dir, &st->info, &gp_id);
    *t_out = t;
    return 0;
#undef AT
}
/* sampleGradient (SCNM:93-100) */
static d3 medium_sample_gradient(v3 dir, d3 ip, float t, state_t *st) {
    v3 g = real_gradient(&st->ctx, V3((float)ip.x, (float)ip.y, (float)ip.z), t, dir, &st->info);
    return to_d3(g);
}
typedef struct { int exited; float t, cont_t; d3 aniso; int gp_id, scheme; float weight; } sample_t;

static float medium_transmittance(const orc_medium *m, v3 pos, v3 dir, float near_t, float far_t, float u, state_t *st);

/* sampleDistance (GPM:221-341) */
static int medium_sample_distance(const orc_medium *m, v3 pos, v3 dir, float near_t, float far_t, float u, state_t *st, sample_t *sm) {
    double start_t = near_t;
    if (!isfinite(far_t)) far_t = (float)(start_t + 2000);
    float max_t = far_t;
    if (st->bounce >= m->d.max_bounces) return 0;
    if (max_t == 0.f) {
        /* GPM:239-248 returns before gpId / info.t / advance() are touched */
        sm->t = max_t; sm->weight = 1.f; sm->exited = 1; sm->scheme = SCN_SCHEME_UNI; sm->gp_id = 0;
        return 1;
    } else if (m->absorption_only) {
        /* GPM:250-258: weight = transmittance() evaluated on the ORIGINAL ray (infinite farT is refused first) */
        if (max_t == INFINITY) return 0;
        sm->t = max_t; sm->weight = medium_transmittance(m, pos, dir, near_t, far_t, u, st); sm->exited = 1; sm->scheme = SCN_SCHEME_UNI;
    } else {
        double t = max_t;
        d3 ro = to_d3(pos), rd = to_d3(dir);
        { double inv = 1.0 / d3_len(rd); rd.x *= inv; rd.y *= inv; rd.z *= inv; }
        do {
            float nt = (float)start_t;
            sm->exited = !medium_intersect(m, pos, dir, nt, far_t, u, st, &t);
            if (t < max_t) {
                d3 ip = {ro.x + rd.x * t, ro.y + rd.y * t, ro.z + rd.z * t};
                d3 g = medium_sample_gradient(dir, ip, (float)t, st);
                st->last_aniso = sm->aniso = g; st->first_scatter = 0;
                if (!isfinite((g.x + g.y + g.z) * (1.0 / 3.0))) { sm->aniso.x = 1.; sm->aniso.y = 0.; sm->aniso.z = 0.; return 0; }
            }
            start_t = t;
        } while (t < max_t && sm->exited);
        if (!sm->exited) {
            d3 dd = to_d3(dir);
            if (d3_dot(sm->aniso, dd) > 0) return 0;
            double l2 = 0.; l2 += sm->aniso.x * sm->aniso.x; l2 += sm->aniso.y * sm->aniso.y; l2 += sm->aniso.z * sm->aniso.z;
            if (l2 < 0.0000001f) { sm->aniso.x = 1.; sm->aniso.y = 0.; sm->aniso.z = 0.; return 0; }
            sm->weight = 1.f;
        } else {
            d3 ip = {ro.x + rd.x * t, ro.y + rd.y * t, ro.z + rd.z * t};
            sm->aniso = medium_sample_gradient(dir, ip, (float)t, st);
            sm->weight = 1.f;
        }
        sm->t = (float)t < max_t ? (float)t : max_t;
        sm->cont_t = (float)t;
        sm->weight *= m->sigma_s_over_t;
        st->last_aniso = sm->aniso;
        sm->scheme = st->scheme;
        st->first_scatter = 0; st->scheme = SCN_SCHEME_UNI; st->bounce++; /* state.advance() */
    }
    sm->gp_id = st->last_gp_id;
    st->info.t += sm->t;
    return 1;
}
/* transmittance (GPM:343-393) */
static float medium_transmittance(const orc_medium *m, v3 pos, v3 dir, float near_t, float far_t, float u, state_t *st) {
    double start_t = near_t;
    if (!isfinite(far_t)) far_t = (float)(start_t + 2000);
    float max_t = far_t;
    double t = max_t;
    d3 ro = to_d3(pos), rd = to_d3(dir);
    { double inv = 1.0 / d3_len(rd); rd.x *= inv; rd.y *= inv; rd.z *= inv; }
    int exited;
    do {
        float nt = (float)start_t;
        exited = !medium_intersect(m, pos, dir, nt, far_t, u, st, &t);
        if (t < max_t) {
            d3 ip = {ro.x + rd.x * t, ro.y + rd.y * t, ro.z + rd.z * t};
            d3 g = medium_sample_gradient(dir, ip, (float)t, st);
            st->last_aniso = g; st->first_scatter = 0;
            if (!isfinite((g.x + g.y + g.z) * (1.0 / 3.0))) return 0.f;
        }
        start_t = t;
    } while (t < max_t && exited);
    return exited ? 1.f : 0.f;
}

/* ------------------------------------------------------------------ threading + batch entry points */
typedef void (*range_fn)(void *ctx, size_t b, size_t e);
typedef struct { range_fn fn; void *ctx; size_t b, e; uint64_t evals; } job_t;
static void *job_main(void *arg) { job_t *j = (job_t *)arg; g_eval_count = 0; j->fn(j->ctx, j->b, j->e); j->evals = g_eval_count; return NULL; }
static void parallel_for(size_t n, int nthreads, range_fn fn, void *ctx) {
    if (nthreads <= 1 || n < 2) { g_eval_count = 0; fn(ctx, 0, n); return; }
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256]; job_t jobs[256]; int used = 0;
    size_t chunk = (n + (size_t)nthreads - 1) / (size_t)nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t b = chunk * (size_t)t, e = b + chunk; if (b > n) b = n; if (e > n) e = n;
        if (b < e) { jobs[used].fn = fn; jobs[used].ctx = ctx; jobs[used].b = b; jobs[used].e = e; pthread_create(&th[used], NULL, job_main, &jobs[used]); used++; }
    }
    uint64_t total = 0;
    for (int t = 0; t < used; ++t) { pthread_join(th[t], NULL); total += jobs[t].evals; }
    g_eval_count = total;
}
static info_t make_info(const uint32_t *pix, size_t i, uint32_t scene_seed, const float *info_t_arr) {
    info_t in; if (pix) { in.pix[0] = pix[4 * i]; in.pix[1] = pix[4 * i + 1]; in.pix[2] = pix[4 * i + 2]; in.pix[3] = pix[4 * i + 3]; } else memset(in.pix, 0, sizeof in.pix);
    in.scene_seed = scene_seed; in.t = info_t_arr ? info_t_arr[i] : 0.f; return in;
}
static void set_coeffs(real_t *r, const float *c) {
    r->c3.value_scale = c[0]; r->c3.gradient_scale = V3(c[1], c[2], c[3]); r->c3.ray_origin = V3(c[4], c[5], c[6]); r->c1 = r->c3;
}
static void get_coeffs(const real_t *r, float *c) {
    const coeff_t *k = r->m->d.sampling_1d ? &r->c1 : &r->c3;
    c[0] = k->value_scale; c[1] = k->gradient_scale.x; c[2] = k->gradient_scale.y; c[3] = k->gradient_scale.z;
    c[4] = k->ray_origin.x; c[5] = k->ray_origin.y; c[6] = k->ray_origin.z;
}
#define P3AT(a, i) V3((a)[3 * (i)], (a)[3 * (i) + 1], (a)[3 * (i) + 2])

typedef struct { orc_medium *m; const float *p, *d; uint32_t seed; float *out; } noise_args;
static void noise3d_range(void *c, size_t b, size_t e) {
    noise_args *a = (noise_args *)c; real_t r; real_init(&r, a->m);
    for (size_t i = b; i < e; ++i) { v4 v = real_eval_noise3d(&r, P3AT(a->p, i), P3AT(a->d, i), a->seed, 0); a->out[4 * i] = v.x; a->out[4 * i + 1] = v.y; a->out[4 * i + 2] = v.z; a->out[4 * i + 3] = v.w; }
}
int orc_eval_noise3d(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int nthreads) {
    noise_args a = {m, p, d, seed, out4}; parallel_for(n, nthreads, noise3d_range, &a); return 0;
}
static void accept_range(void *c, size_t b, size_t e) {
    noise_args *a = (noise_args *)c; real_t r; real_init(&r, a->m); uint32_t *out = (uint32_t *)a->out;
    for (size_t i = b; i < e; ++i) { g_accept_count = 0; (void)real_eval_noise3d(&r, P3AT(a->p, i), P3AT(a->d, i), a->seed, 0); out[i] = g_accept_count; }
}
int orc_accepted_counts(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, uint32_t *out, int nthreads) {
    noise_args a = {m, p, d, seed, (float *)out}; parallel_for(n, nthreads, accept_range, &a); return 0;
}
static void noise1d_range(void *c, size_t b, size_t e) {
    noise_args *a = (noise_args *)c; real_t r; real_init(&r, a->m); float R = cov_radius(a->m, 1, 1.0f);
    for (size_t i = b; i < e; ++i) { v4 v = real_noise1d_norm(&r, P3AT(a->p, i), P3AT(a->d, i), a->seed, R, 1.0f, 0, 0); a->out[4 * i] = v.x; a->out[4 * i + 1] = v.y; a->out[4 * i + 2] = v.z; a->out[4 * i + 3] = v.w; }
}
int orc_eval_noise1d(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int nthreads) {
    noise_args a = {m, p, d, seed, out4}; parallel_for(n, nthreads, noise1d_range, &a); return 0;
}

typedef struct { orc_medium *m; const float *p, *d, *t; const uint32_t *pix; uint32_t scene_seed; const float *info_t, *coeffs; float *val, *grad; int *gpid; } eval_args;
static void eval_range(void *c, size_t b, size_t e) {
    eval_args *a = (eval_args *)c; real_t r; real_init(&r, a->m);
    for (size_t i = b; i < e; ++i) {
        if (a->coeffs) set_coeffs(&r, a->coeffs + 7 * i);
        info_t in = make_info(a->pix, i, a->scene_seed, a->info_t);
        float ti = a->t ? a->t[i] : 0.f; int id = 0;
        if (a->val) { a->val[i] = real_value(&r, P3AT(a->p, i), P3AT(a->d, i), &in, &id); if (a->gpid) a->gpid[i] = id; }
        if (a->grad) { v3 g = real_gradient(&r, P3AT(a->p, i), ti, P3AT(a->d, i), &in); a->grad[3 * i] = g.x; a->grad[3 * i + 1] = g.y; a->grad[3 * i + 2] = g.z; }
    }
}
int orc_eval(orc_medium *m, size_t n, const float *p, const float *d, const float *t, const uint32_t *pix, uint32_t sceneSeed,
             const float *info_t_arr, const float *coeffs, float *out_val, float *out_grad, int *out_gpid, int nthreads) {
    eval_args a = {m, p, d, t, pix, sceneSeed, info_t_arr, coeffs, out_val, out_grad, out_gpid}; parallel_for(n, nthreads, eval_range, &a); return 0;
}

typedef struct { orc_medium *m; const float *p, *d, *tv, *tg; const uint32_t *pix; uint32_t scene_seed; const float *info_t; float *out; } cond_args;
static void cond_range(void *c, size_t b, size_t e) {
    cond_args *a = (cond_args *)c;
    for (size_t i = b; i < e; ++i) {
        real_t r; real_init(&r, a->m); info_t in = make_info(a->pix, i, a->scene_seed, a->info_t);
        real_conditioning(&r, P3AT(a->p, i), P3AT(a->d, i), a->tv[i], P3AT(a->tg, i), &in);
        get_coeffs(&r, a->out + 7 * i);
    }
}
int orc_conditioning(orc_medium *m, size_t n, const float *p, const float *d, const float *target_val, const float *target_grad,
                     const uint32_t *pix, uint32_t sceneSeed, const float *info_t_arr, float *out_coeffs, int nthreads) {
    cond_args a = {m, p, d, target_val, target_grad, pix, sceneSeed, info_t_arr, out_coeffs}; parallel_for(n, nthreads, cond_range, &a); return 0;
}

typedef struct {
    orc_medium *m; const float *o, *d, *tmin, *tmax, *u; const uint32_t *pix; uint32_t scene_seed;
    uint8_t *fs; float *lv; double *lg; int *bounce, *lgp; float *info_t;
    uint8_t *ok, *exited; float *t_out, *cont_t; double *grad; int *gp_id; uint8_t *scheme; float *coeffs, *weight;
    const uint8_t *cfs; const float *clv; const double *clg; const int *cbounce, *clgp; const float *cinfo_t; float *T_out;
} sd_args;
static void sd_range(void *c, size_t b, size_t e) {
    sd_args *a = (sd_args *)c;
    for (size_t i = b; i < e; ++i) {
        state_t st; memset(&st, 0, sizeof st);
        st.first_scatter = a->fs[i] != 0; st.last_val = a->lv[i];
        st.last_aniso.x = a->lg[3 * i]; st.last_aniso.y = a->lg[3 * i + 1]; st.last_aniso.z = a->lg[3 * i + 2];
        st.bounce = a->bounce[i]; st.last_gp_id = a->lgp[i]; st.scheme = SCN_SCHEME_UNI;
        st.info = make_info(a->pix, i, a->scene_seed, a->info_t);
        sample_t sm; memset(&sm, 0, sizeof sm);
        int res = medium_sample_distance(a->m, P3AT(a->o, i), P3AT(a->d, i), a->tmin[i], a->tmax[i], a->u[i], &st, &sm);
        a->ok[i] = (uint8_t)res; a->exited[i] = (uint8_t)sm.exited; a->t_out[i] = sm.t; a->cont_t[i] = sm.cont_t;
        a->grad[3 * i] = sm.aniso.x; a->grad[3 * i + 1] = sm.aniso.y; a->grad[3 * i + 2] = sm.aniso.z;
        a->gp_id[i] = res ? sm.gp_id : st.last_gp_id; a->scheme[i] = (uint8_t)sm.scheme;
        if (a->weight) { a->weight[3 * i] = a->weight[3 * i + 1] = a->weight[3 * i + 2] = sm.weight; }
        for (int k = 0; k < 7; ++k) a->coeffs[7 * i + k] = 0.f;
        if (st.has_ctx) get_coeffs(&st.ctx, a->coeffs + 7 * i);
        a->fs[i] = (uint8_t)st.first_scatter; a->lv[i] = st.last_val;
        a->lg[3 * i] = st.last_aniso.x; a->lg[3 * i + 1] = st.last_aniso.y; a->lg[3 * i + 2] = st.last_aniso.z;
        a->bounce[i] = st.bounce; a->lgp[i] = st.last_gp_id; a->info_t[i] = st.info.t;
    }
}
int orc_sample_distance(orc_medium *m, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                        const uint32_t *pix, uint32_t sceneSeed,
                        uint8_t *first_scatter, float *last_val, double *last_grad, int *bounce, int *last_gp_id, float *info_t_arr,
                        uint8_t *ok, uint8_t *exited, float *t_out, float *cont_t, double *grad, int *gp_id, uint8_t *scheme,
                        float *coeffs, float *weight, int nthreads) {
    sd_args a; memset(&a, 0, sizeof a);
    a.m = m; a.o = o; a.d = d; a.tmin = tmin; a.tmax = tmax; a.u = u; a.pix = pix; a.scene_seed = sceneSeed;
    a.fs = first_scatter; a.lv = last_val; a.lg = last_grad; a.bounce = bounce; a.lgp = last_gp_id; a.info_t = info_t_arr;
    a.ok = ok; a.exited = exited; a.t_out = t_out; a.cont_t = cont_t; a.grad = grad; a.gp_id = gp_id; a.scheme = scheme; a.coeffs = coeffs; a.weight = weight;
    parallel_for(n, nthreads, sd_range, &a); return 0;
}
static void tr_range(void *c, size_t b, size_t e) {
    sd_args *a = (sd_args *)c;
    for (size_t i = b; i < e; ++i) {
        state_t st; memset(&st, 0, sizeof st);
        st.first_scatter = a->cfs[i] != 0; st.last_val = a->clv[i];
        st.last_aniso.x = a->clg[3 * i]; st.last_aniso.y = a->clg[3 * i + 1]; st.last_aniso.z = a->clg[3 * i + 2];
        st.bounce = a->cbounce[i]; st.last_gp_id = a->clgp[i]; st.scheme = SCN_SCHEME_UNI;
        st.info = make_info(a->pix, i, a->scene_seed, a->cinfo_t);
        a->T_out[i] = medium_transmittance(a->m, P3AT(a->o, i), P3AT(a->d, i), a->tmin[i], a->tmax[i], a->u[i], &st);
    }
}
int orc_transmittance(orc_medium *m, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                      const uint32_t *pix, uint32_t sceneSeed,
                      const uint8_t *first_scatter, const float *last_val, const double *last_grad, const int *bounce,
                      const int *last_gp_id, const float *info_t_arr, float *T_out, int nthreads) {
    sd_args a; memset(&a, 0, sizeof a);
    a.m = m; a.o = o; a.d = d; a.tmin = tmin; a.tmax = tmax; a.u = u; a.pix = pix; a.scene_seed = sceneSeed;
    a.cfs = first_scatter; a.clv = last_val; a.clg = last_grad; a.cbounce = bounce; a.clgp = last_gp_id; a.cinfo_t = info_t_arr; a.T_out = T_out;
    parallel_for(n, nthreads, tr_range, &a); return 0;
}

typedef struct { orc_medium *m; const float *dir, *nrm, *p, *ts; const uint32_t *pix; uint32_t scene_seed; const float *info_t, *coeffs; float *pdf, *grad; } nee_args;
static void nee_range(void *c, size_t b, size_t e) {
    nee_args *a = (nee_args *)c; real_t r; real_init(&r, a->m);
    for (size_t i = b; i < e; ++i) {
        if (a->coeffs) set_coeffs(&r, a->coeffs + 7 * i);
        info_t in = make_info(a->pix, i, a->scene_seed, a->info_t);
        if (a->pdf) a->pdf[i] = real_nee_pdf(&r, P3AT(a->dir, i), P3AT(a->nrm, i), P3AT(a->p, i), a->ts[i], &in);
        if (a->grad) { v3 g = real_nee_grad(&r, P3AT(a->dir, i), P3AT(a->nrm, i), P3AT(a->p, i), &in); a->grad[3 * i] = g.x; a->grad[3 * i + 1] = g.y; a->grad[3 * i + 2] = g.z; }
    }
}
int orc_nee(orc_medium *m, size_t n, const float *ray_dir, const float *normal, const float *p, const float *t_segment,
            const uint32_t *pix, uint32_t sceneSeed, const float *info_t_arr, const float *coeffs,
            float *out_pdf, float *out_grad, int nthreads) {
    nee_args a = {m, ray_dir, normal, p, t_segment, pix, sceneSeed, info_t_arr, coeffs, out_pdf, out_grad}; parallel_for(n, nthreads, nee_range, &a); return 0;
}
