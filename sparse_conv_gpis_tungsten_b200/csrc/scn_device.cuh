// scn_device.cuh — device functions of the GPIS medium query (hash, PCG, kernels, noise, conditioning, NEE).
//
// Numerics contract (DESIGN.md "Bit-exactness"): everything that feeds an INTEGER or a hard DECISION of the
// reference — cell indices, hashes, PCG streams, the impulse rejection test |p - p_i|^2 < 1, march control in
// double — uses separately rounded IEEE operations (__fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn, __dmul_rn/...)
// in the reference's association order, because the reference is built without FMA (-msse4.2).  After the
// test, kernel evaluation and accumulation are ordinary FP32 FMA/SFU code (1e-4 relative tolerance).
//
// File:line citations are relative to the reference's src/core/ (SCN = math/SparseConvolutionNoise.cpp,
// GPF = math/GPFunctions.cpp).  The same header compiles for the host (tests/hostsim) through SCN_HD so that
// control flow can be debugged without a GPU; the product library only ever launches the __global__ kernels.
#ifndef SCN_DEVICE_CUH_
#define SCN_DEVICE_CUH_

#include <math.h>
#include <stdint.h>

#include "scn_internal.h"

#if defined(__CUDACC__)
#define SCN_HD __host__ __device__ __forceinline__
// Large bodies (the impulse loops and everything that calls them more than once) are real functions on the device:
// forcing them inline multiplies the 27-cell loop nest dozens of times and takes ptxas tens of minutes.
#define SCN_FN inline __host__ __device__ __noinline__
#else
#define SCN_HD inline
#define SCN_FN inline
#endif

namespace scn {

struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };
struct dbl3 { double x, y, z; };
struct Frame { f3 n, t, b; };
struct Coeff { float vs; f3 gs; f3 org; };                 // SparseConvConditioningCoefficients3D/1D
struct Info { uint32_t px, py, spp, seg, scene_seed; float t; }; // RayInfo

// ---------------------------------------------------------------- bounds checks of the tuned kernels (debug build only)
// compute-sanitizer is closed on this pool (profiles/r02_compute_sanitizer_closed_on_pool.txt), so the checks it would make on the cached
// march kernels are written out: -DSCN_DEBUG_BOUNDS (tools/debug_bounds.py builds that library next to the product one) counts every
// violated index / capacity assertion in a device counter and poisons the shared-memory caches with NaN at kernel start, so that a read of a
// slot that was never written shows up as a changed result.  The product build compiles all of it out.
#if defined(SCN_DEBUG_BOUNDS) && defined(__CUDACC__)
static __device__ unsigned long long g_scn_bounds_fail[8];
#define SCN_CHECK(cond, code) do { if (!(cond)) atomicAdd(&g_scn_bounds_fail[code], 1ULL); } while (0)
#define SCN_SMEM_FILL __int_as_float(0x7fc00000)
#else
#define SCN_CHECK(cond, code) ((void)0)
#define SCN_SMEM_FILL 0.f
#endif

// ---------------------------------------------------------------- exactly rounded primitives
#if defined(__CUDA_ARCH__)
SCN_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
SCN_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
SCN_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
SCN_HD float xdiv(float a, float b) { return __fdiv_rn(a, b); }
SCN_HD float xsqrt(float a) { return __fsqrt_rn(a); }
SCN_HD double ymul(double a, double b) { return __dmul_rn(a, b); }
SCN_HD double yadd(double a, double b) { return __dadd_rn(a, b); }
SCN_HD double ysub(double a, double b) { return __dsub_rn(a, b); }
SCN_HD double ydiv(double a, double b) { return __ddiv_rn(a, b); }
SCN_HD double ysqrt(double a) { return __dsqrt_rn(a); }
SCN_HD float fast_exp(float x) { return __expf(x); }
#else
SCN_HD float xmul(float a, float b) { return a * b; }
SCN_HD float xadd(float a, float b) { return a + b; }
SCN_HD float xsub(float a, float b) { return a - b; }
SCN_HD float xdiv(float a, float b) { return a / b; }
SCN_HD float xsqrt(float a) { return sqrtf(a); }
SCN_HD double ymul(double a, double b) { return a * b; }
SCN_HD double yadd(double a, double b) { return a + b; }
SCN_HD double ysub(double a, double b) { return a - b; }
SCN_HD double ydiv(double a, double b) { return a / b; }
SCN_HD double ysqrt(double a) { return sqrt(a); }
SCN_HD float fast_exp(float x) { return expf(x); }
#endif

SCN_HD f3 F3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
SCN_HD f4 F4(float x, float y, float z, float w) { f4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
// math/Vec.hpp:134-206 in the reference's association order
SCN_HD float xdot(f3 a, f3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
SCN_HD float xlensq(f3 a) { return xadd(xadd(xmul(a.x, a.x), xmul(a.y, a.y)), xmul(a.z, a.z)); }
SCN_HD f3 xnormalized(f3 a) { float inv = xdiv(1.0f, xsqrt(xlensq(a))); return F3(xmul(a.x, inv), xmul(a.y, inv), xmul(a.z, inv)); }
SCN_HD f3 xsub3(f3 a, f3 b) { return F3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
SCN_HD f3 xadd3(f3 a, f3 b) { return F3(xadd(a.x, b.x), xadd(a.y, b.y), xadd(a.z, b.z)); }
SCN_HD f3 xscale3(f3 a, float s) { return F3(xmul(a.x, s), xmul(a.y, s), xmul(a.z, s)); }
SCN_HD f3 xdiv3(f3 a, float s) { return F3(xdiv(a.x, s), xdiv(a.y, s), xdiv(a.z, s)); }
SCN_HD f3 xmul3(f3 a, f3 b) { return F3(xmul(a.x, b.x), xmul(a.y, b.y), xmul(a.z, b.z)); }
SCN_HD f4 add4(f4 a, f4 b) { return F4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
SCN_HD f4 scale4(f4 a, float s) { return F4(a.x * s, a.y * s, a.z * s, a.w * s); }
SCN_HD float sqr(float a) { return a * a; }
SCN_HD float max3(float a, float b, float c) { float m = a > b ? a : b; return m > c ? m : c; }

// ---------------------------------------------------------------- H1 xxhash32 (math/MathUtil.hpp:179-224)
constexpr uint32_t XP2 = 2246822519U, XP3 = 3266489917U, XP4 = 668265263U, XP5 = 374761393U;
SCN_HD uint32_t rotl17(uint32_t h) { return (h << 17) | (h >> 15); }
SCN_HD uint32_t xx_final(uint32_t h) { h = XP2 * (h ^ (h >> 15)); h = XP3 * (h ^ (h >> 13)); return h ^ (h >> 16); }
SCN_HD uint32_t xxhash32_1(uint32_t p) { uint32_t h = p + XP5; h = XP4 * rotl17(h); return xx_final(h); }
SCN_HD uint32_t xxhash32_2(uint32_t x, uint32_t y) { uint32_t h = y + XP5 + x * XP3; h = XP4 * rotl17(h); return xx_final(h); }
SCN_HD uint32_t xxhash32_4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    uint32_t h = w + XP5 + x * XP3; h = XP4 * rotl17(h); h += y * XP3; h = XP4 * rotl17(h); h += z * XP3; h = XP4 * rotl17(h);
    return xx_final(h);
}

// ---------------------------------------------------------------- H2 PCG32 XSH-RR (sampling/UniformSampler.hpp:41-75)
constexpr uint64_t PCG_MULT = 6364136223846793005ULL;
SCN_HD uint32_t pcg_output(uint64_t old) {
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31u));
}
SCN_HD float u32_to_unit(uint32_t i) { // math/BitManip.hpp:47-50
#if defined(__CUDA_ARCH__)
    return __uint_as_float((i >> 9u) | 0x3F800000u) - 1.0f;
#else
    union { uint32_t i; float f; } u; u.i = (i >> 9u) | 0x3F800000u; return u.f - 1.0f;
#endif
}
struct Pcg {
    uint64_t s;
    SCN_HD float next() { uint64_t old = s; s = old * PCG_MULT + 1ULL; return u32_to_unit(pcg_output(old)); }
    // set_state(h + 1u): the +1 wraps in 32 bits before widening (SCN:376,409); then two draws are discarded.
    SCN_HD void seed32(uint32_t h) { s = (uint64_t)(uint32_t)(h + 1u); s = s * PCG_MULT + 1ULL; s = s * PCG_MULT + 1ULL; }
};

// ---------------------------------------------------------------- TangentFrame (math/TangentFrame.hpp:74-99)
SCN_HD Frame make_frame(f3 nin) {
    Frame f; f.n = xnormalized(nin);
    float sign = copysignf(1.0f, f.n.z);
    float a = xdiv(-1.0f, xadd(sign, f.n.z));
    float b = xmul(xmul(f.n.x, f.n.y), a);
    f.t = F3(xadd(1.0f, xmul(xmul(xmul(sign, f.n.x), f.n.x), a)), xmul(sign, b), xmul(-sign, f.n.x));
    f.b = F3(b, xadd(sign, xmul(xmul(f.n.y, f.n.y), a)), -f.n.y);
    f.t = xnormalized(f.t); f.b = xnormalized(f.b);
    return f;
}
SCN_HD f3 to_local(const Frame &f, f3 p) { return F3(xdot(f.t, p), xdot(f.b, p), xdot(f.n, p)); }
SCN_HD f3 to_global(const Frame &f, f3 p) {
    return F3(f.t.x * p.x + f.b.x * p.y + f.n.x * p.z, f.t.y * p.x + f.b.y * p.y + f.n.y * p.z, f.t.z * p.x + f.b.z * p.y + f.n.z * p.z);
}

// ---------------------------------------------------------------- procedural fields (GPF:53-138) — double like the reference
// FieldK holds what the reference recomputes per call from the JSON numbers (log(min^2), log(max^2), 1/(end-start),
// -start*scale): evaluated once on the host with the host libm.
SCN_HD double field_ramp(const FieldK &k, double c, int second) {
    double r = yadd(ymul(c, k.scale[second]), k.off[second]); r = r > 0.0 ? r : 0.0; r = r < 1.0 ? r : 1.0;
    return sqrt(exp(yadd(ymul(k.la[second], ysub(1.0, r)), ymul(k.lb[second], r))));
}
// RegularGrid<ElemType>::getValue (math/GPFunctions.hpp:255-306), component `c` of `ncomp` interleaved ones, in double like the reference:
// p_grid = res (p - min) / diagonal - 0.5, clamped integer coordinates, element (x res + y) res + z; point, trilinear (lerp = a (1 - t) + b t,
// math/MathUtil.hpp:91-94, x innermost then y then z: GPFunctions.hpp:173-188) or triquadratic (z, then y, then x: :190-230) interpolation.
SCN_HD double grid_at(const FieldK &k, int x, int y, int z, int c) {
    const int r = k.gres;
    x = x < 0 ? 0 : (x > r - 1 ? r - 1 : x); y = y < 0 ? 0 : (y > r - 1 ? r - 1 : y); z = z < 0 ? 0 : (z > r - 1 ? r - 1 : z);
    return k.grid[((size_t)(x * r + y) * r + z) * k.gcomp + c];
}
SCN_HD double lerp_d(double a, double b, double t) { return yadd(ymul(a, ysub(1.0, t)), ymul(b, t)); }
SCN_HD double quad_d(double v0, double v1, double v2, double w) { // _interpolate (GPFunctions.hpp:193-202)
    const double a = ysub(ymul(0.5, yadd(v0, v2)), v1), b = ymul(0.5, ysub(v2, v0));
    return yadd(ymul(w, yadd(ymul(w, a), b)), v1);
}
SCN_FN double field_grid(const FieldK &k, dbl3 p, int c) {
    const double res = (double)k.gres;
    const double gx = ysub(ydiv(ymul(res, ysub(p.x, k.gmin[0])), k.gdiag[0]), 0.5), gy = ysub(ydiv(ymul(res, ysub(p.y, k.gmin[1])), k.gdiag[1]), 0.5),
                 gz = ysub(ydiv(ymul(res, ysub(p.z, k.gmin[2])), k.gdiag[2]), 0.5);
    const double fx = floor(gx), fy = floor(gy), fz = floor(gz);
    const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    const double u = ysub(gx, (double)ix), v = ysub(gy, (double)iy), w = ysub(gz, (double)iz);
    if (k.ginterp == SCN_INTERP_POINT) return grid_at(k, ix, iy, iz, c);
    if (k.ginterp == SCN_INTERP_LINEAR) {
        const double z0 = lerp_d(lerp_d(grid_at(k, ix, iy, iz, c), grid_at(k, ix + 1, iy, iz, c), u), lerp_d(grid_at(k, ix, iy + 1, iz, c), grid_at(k, ix + 1, iy + 1, iz, c), u), v);
        const double z1 = lerp_d(lerp_d(grid_at(k, ix, iy, iz + 1, c), grid_at(k, ix + 1, iy, iz + 1, c), u), lerp_d(grid_at(k, ix, iy + 1, iz + 1, c), grid_at(k, ix + 1, iy + 1, iz + 1, c), u), v);
        return lerp_d(z0, z1, w);
    }
    double vx[3];
    for (int dx = 0; dx < 3; ++dx) {
        double vy[3];
        for (int dy = 0; dy < 3; ++dy)
            vy[dy] = quad_d(grid_at(k, ix - 1 + dx, iy - 1 + dy, iz - 1, c), grid_at(k, ix - 1 + dx, iy - 1 + dy, iz, c), grid_at(k, ix - 1 + dx, iy - 1 + dy, iz + 1, c), w);
        vx[dx] = quad_d(vy[0], vy[1], vy[2], v);
    }
    return quad_d(vx[0], vx[1], vx[2], u);
}
// ---- fbm fields (math/SdfFunctions.cpp:200-296): simplex noise over a sin() hash.  Float arithmetic in the reference's order; the hash
// takes sin in DOUBLE of the float dot product and rounds 4096 sin(.) to float (random3, :201-210), which the device's double sin
// reproduces (both are within an ulp of the exact value; the rounding to float hides that except ~1e-8 of the time).
SCN_HD f3 fbm_random3(f3 c) {
    const float dt = xadd(xadd(xmul(c.x, 17.0f), xmul(c.y, 59.4f)), xmul(c.z, 15.0f));
    float j = (float)ymul(4096.0, sin((double)dt));
    f3 r;
    double v = ymul(512.0, (double)j); r.z = (float)ysub(v, floor(v));
    j = (float)ymul((double)j, .125);
    v = ymul(512.0, (double)j); r.x = (float)ysub(v, floor(v));
    j = (float)ymul((double)j, .125);
    v = ymul(512.0, (double)j); r.y = (float)ysub(v, floor(v));
    return F3(xsub(r.x, 0.5f), xsub(r.y, 0.5f), xsub(r.z, 0.5f));
}
SCN_FN float fbm_simplex3d(f3 p) {
    const float F3c = 0.3333333f, G3c = 0.1666667f;
    const float ds = xadd(xadd(xmul(p.x, F3c), xmul(p.y, F3c)), xmul(p.z, F3c));
    const f3 s = F3(floorf(xadd(p.x, ds)), floorf(xadd(p.y, ds)), floorf(xadd(p.z, ds)));
    const float dg = xadd(xadd(xmul(s.x, G3c), xmul(s.y, G3c)), xmul(s.z, G3c));
    const f3 x = F3(xadd(xsub(p.x, s.x), dg), xadd(xsub(p.y, s.y), dg), xadd(xsub(p.z, s.z), dg));
    const f3 e = F3(xsub(x.x, x.y) < 0.f ? 0.f : 1.f, xsub(x.y, x.z) < 0.f ? 0.f : 1.f, xsub(x.z, x.x) < 0.f ? 0.f : 1.f);
    const f3 i1 = F3(xmul(e.x, xsub(1.0f, e.z)), xmul(e.y, xsub(1.0f, e.x)), xmul(e.z, xsub(1.0f, e.y)));
    const f3 i2 = F3(xsub(1.0f, xmul(e.z, xsub(1.0f, e.x))), xsub(1.0f, xmul(e.x, xsub(1.0f, e.y))), xsub(1.0f, xmul(e.y, xsub(1.0f, e.z))));
    const float g2 = xmul(2.0f, G3c), g3 = xmul(3.0f, G3c);
    const f3 x1 = F3(xadd(xsub(x.x, i1.x), G3c), xadd(xsub(x.y, i1.y), G3c), xadd(xsub(x.z, i1.z), G3c));
    const f3 x2 = F3(xadd(xsub(x.x, i2.x), g2), xadd(xsub(x.y, i2.y), g2), xadd(xsub(x.z, i2.z), g2));
    const f3 x3 = F3(xadd(xsub(x.x, 1.0f), g3), xadd(xsub(x.y, 1.0f), g3), xadd(xsub(x.z, 1.0f), g3));
    float w0 = xsub(0.6f, xdot(x, x)), w1 = xsub(0.6f, xdot(x1, x1)), w2 = xsub(0.6f, xdot(x2, x2)), w3 = xsub(0.6f, xdot(x3, x3));
    w0 = w0 > 0.f ? w0 : 0.f; w1 = w1 > 0.f ? w1 : 0.f; w2 = w2 > 0.f ? w2 : 0.f; w3 = w3 > 0.f ? w3 : 0.f;
    float d0 = xdot(fbm_random3(s), x), d1 = xdot(fbm_random3(F3(xadd(s.x, i1.x), xadd(s.y, i1.y), xadd(s.z, i1.z))), x1),
          d2 = xdot(fbm_random3(F3(xadd(s.x, i2.x), xadd(s.y, i2.y), xadd(s.z, i2.z))), x2), d3 = xdot(fbm_random3(F3(xadd(s.x, 1.0f), xadd(s.y, 1.0f), xadd(s.z, 1.0f))), x3);
    w0 = xmul(w0, w0); w1 = xmul(w1, w1); w2 = xmul(w2, w2); w3 = xmul(w3, w3);
    w0 = xmul(w0, w0); w1 = xmul(w1, w1); w2 = xmul(w2, w2); w3 = xmul(w3, w3);
    d0 = xmul(d0, w0); d1 = xmul(d1, w1); d2 = xmul(d2, w2); d3 = xmul(d3, w3);
    return xadd(xadd(xadd(xmul(d0, 52.0f), xmul(d1, 52.0f)), xmul(d2, 52.0f)), xmul(d3, 52.0f));
}
SCN_FN double fbm_noise(dbl3 uv, int octaves) { // fbm (math/SdfFunctions.cpp:276-296)
    const float gain = 0.65f, lacunarity = 2.1042f;
    float frequency = 0.5f, amplitude = gain;
    uv.x = ymul(uv.x, 5.0); uv.y = ymul(uv.y, 5.0); uv.z = ymul(uv.z, 5.0);
    const f3 q = F3((float)uv.x, (float)uv.y, (float)uv.z);
    float total = fbm_simplex3d(q);
    for (int i = 0; i < octaves; ++i) {
        total = xadd(total, xmul(fbm_simplex3d(xscale3(q, frequency)), amplitude));
        frequency = xmul(frequency, lacunarity); amplitude = xmul(amplitude, gain);
    }
    total = (float)ydiv(yadd((double)total, 2.0), 4.0);
    return (double)total;
}
SCN_HD dbl3 dadd_s(dbl3 a, double s) { dbl3 r; r.x = yadd(a.x, s); r.y = yadd(a.y, s); r.z = yadd(a.z, s); return r; }
SCN_HD dbl3 dmul_s(dbl3 a, double s) { dbl3 r; r.x = ymul(a.x, s); r.y = ymul(a.y, s); r.z = ymul(a.z, s); return r; }
SCN_HD double clamp01_d(double v) { const double t = v > 0. ? v : 0.; return t < 1. ? t : 1.; } // min(max(v, 0), 1) with Tungsten's a > b ? a : b
// the fbm value both Sandstone branches share (GPF:72-73 with 2 octaves, :106-107 with 10), and Rust's smoothStep argument (:80, :114)
SCN_FN double fbm_sandstone(dbl3 p, int oct) { const dbl3 ps = dmul_s(p, 0.3); return fbm_noise(dadd_s(ps, fbm_noise(dadd_s(ps, fbm_noise(ps, oct)), oct)), oct); }
SCN_FN double fbm_rust(dbl3 p, int oct, double sign) {
    const dbl3 ps = dmul_s(p, 2.0);
    const double a = ymul(fbm_noise(dmul_s(ps, .1), oct), 0.4), b = ymul(fbm_noise(dmul_s(ps, 25.), oct), 0.1);
    const double v = yadd(fbm_noise(dadd_s(ps, a), oct), sign > 0 ? b : -b); // scalar field: minus (GPF:80); vector field: plus (GPF:114)
    const double x = clamp01_d(ydiv(ysub(v, 0.4), ysub(0.6, 0.4)));
    return ymul(ymul(x, x), ysub(3.0, ymul(2.0, x))); // smoothStep (math/MathUtil.hpp:108-112)
}
// FULL = false (the iso-ray instantiations of k_march for media whose fields are constants / ramps): the grid and fbm fields are compiled
// out, because a reachable call chain into them costs the two-level kernel its register allocation (C5: 92 B of spills, 32.6 -> 29.3 M queries/s).
SCN_HD bool field_is_heavy(const FieldK &k) { return k.type == SCN_FIELD_REGULAR_GRID || k.type == SCN_FIELD_SANDSTONE || k.type == SCN_FIELD_RUST; }
SCN_FN double field_scalar_heavy(const FieldK &k, dbl3 p) {
    switch (k.type) {
    case SCN_FIELD_SANDSTONE: { // ProceduralNoise::operator() (GPF:70-77)
        const double f = fbm_sandstone(p, 2);
        const double cx = ysub(sqrt(ymul(ymul(f, 1.9), 1.2)), 0.35), t = clamp01_d(cx);
        return yadd(ymul(k.fmin, ysub(1.0, t)), ymul(k.fmax, t));
    }
    case SCN_FIELD_RUST: { const double t = clamp01_d(fbm_rust(p, 2, -1.0)); return yadd(ymul(k.fmin, ysub(1.0, t)), ymul(k.fmax, t)); } // GPF:78-83
    default: return field_grid(k, p, 0);
    }
}
template <bool FULL = true>
SCN_FN double field_scalar(const FieldK &k, dbl3 p) {
    if (FULL && field_is_heavy(k)) return field_scalar_heavy(k, p);
    switch (k.type) {
    case SCN_FIELD_CONSTANT: return k.val[0];
    case SCN_FIELD_BOTTOM_TOP: return field_ramp(k, p.y, 0) - 1.;
    case SCN_FIELD_LEFT_RIGHT: return field_ramp(k, p.x, 0) - 1.;
    case SCN_FIELD_FRONT_BACK: return field_ramp(k, p.z, 0) - 1.;
    case SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT: { float bt = (float)field_ramp(k, p.y, 0), lr = (float)field_ramp(k, p.x, 1); return (double)xmul(bt, lr) - 1.; }
    default: return 1.;
    }
}
SCN_HD dbl3 to_dbl3(f3 p) { dbl3 r; r.x = p.x; r.y = p.y; r.z = p.z; return r; }

// ---------------------------------------------------------------- covariance helpers (see oracle/scn_oracle.c for the line map)
SCN_FN float cov_lateral_scale_heavy(const Params &P, f3 p) {
    if (P.ls.type == SCN_FIELD_SANDSTONE) { // ProceduralNoiseVec::operator() (GPF:104-111): clamp(col * 0.2, 0, 1), 10 octaves; largest component
        const double f = fbm_sandstone(to_dbl3(p), 10);
        const double c0 = clamp01_d(ymul(ysub(sqrt(ymul(ymul(f, 1.9), 1.2)), 0.35), 0.2)), c1 = clamp01_d(ymul(ysub(sqrt(ymul(ymul(f, 0.7), 1.2)), 0.35), 0.2)),
                     c2 = clamp01_d(ymul(ysub(sqrt(ymul(ymul(f, 0.25), 1.2)), 0.35), 0.2));
        return max3((float)c0, (float)c1, (float)c2);
    }
    if (P.ls.type == SCN_FIELD_RUST) { // GPF:112-117: lerp((0.278, 0.212, 0.141), 1, f)
        const double f = fbm_rust(to_dbl3(p), 10, 1.0), g = ysub(1.0, f);
        return max3((float)yadd(ymul(0.278, g), ymul(1.0, f)), (float)yadd(ymul(0.212, g), ymul(1.0, f)), (float)yadd(ymul(0.141, g), ymul(1.0, f)));
    }
    // regular_grid: getKernelScale (GPF:1729-1735): Vec3f of the interpolated Vec3d, largest component
    const dbl3 q = to_dbl3(p);
    return max3((float)field_grid(P.ls, q, 0), (float)field_grid(P.ls, q, 1), (float)field_grid(P.ls, q, 2));
}
template <bool FULL = true>
SCN_HD float cov_lateral_scale(const Params &P, f3 p) { // sparseConvNoiseLateralScale
    if (P.stationary || P.ls.type == SCN_FIELD_NONE) return 1.0f;
    if (P.ls.type == SCN_FIELD_CONSTANT) return max3((float)P.ls.val[0], (float)P.ls.val[1], (float)P.ls.val[2]);
    if (FULL && field_is_heavy(P.ls)) return cov_lateral_scale_heavy(P, p);
    return (float)field_scalar<false>(P.ls, to_dbl3(p));
}
template <bool FULL = true>
SCN_HD float cov_ns_scale(const Params &P, f3 p) { // nonStationarySplattingKernelScale
    if (P.stationary || P.multires) return 1.f;
    return xdiv(cov_lateral_scale<FULL>(P, p), P.ls_max);
}
template <bool FULL = true>
SCN_HD float cov_amplitude(const Params &P, f3 p) { // sparseConvNoiseAmplitude
    if (P.stationary) return P.sigma;
    double var = P.var.type != SCN_FIELD_NONE ? field_scalar<FULL>(P.var, to_dbl3(p)) : 1.;
    return (float)(var * (double)P.sigma);
}
SCN_HD float tr_scale(const Params &P, float s) { return P.stationary ? s : (P.multires ? s : P.ls_max); }
SCN_HD f3 l_inv(const Params &P) { return F3(P.l_aniso_inv[0], P.l_aniso_inv[1], P.l_aniso_inv[2]); }
SCN_HD f3 l_fwd(const Params &P) { return F3(P.l_aniso[0], P.l_aniso[1], P.l_aniso[2]); }
// "useAnisoMtx": full 3x3 transforms (GPF:668-675); Eigen's fixed-size products reduce three terms as x0 + (x1 + x2), separately rounded
SCN_HD f3 mtx_mulv(const float a[9], f3 b) {
    return F3(xadd(xmul(a[0], b.x), xadd(xmul(a[1], b.y), xmul(a[2], b.z))), xadd(xmul(a[3], b.x), xadd(xmul(a[4], b.y), xmul(a[5], b.z))),
              xadd(xmul(a[6], b.x), xadd(xmul(a[7], b.y), xmul(a[8], b.z))));
}
SCN_HD f3 mtx_tmulv(const float a[9], f3 b) {
    return F3(xadd(xmul(a[0], b.x), xadd(xmul(a[3], b.y), xmul(a[6], b.z))), xadd(xmul(a[1], b.x), xadd(xmul(a[4], b.y), xmul(a[7], b.z))),
              xadd(xmul(a[2], b.x), xadd(xmul(a[5], b.y), xmul(a[8], b.z))));
}
SCN_HD f3 cov_pos_w2l(const Params &P, f3 v, float s) { return xdiv3(P.use_mtx ? mtx_mulv(P.mW, v) : xmul3(l_inv(P), v), tr_scale(P, s)); }     // GPF:631-633
SCN_HD f3 cov_grad_w2l(const Params &P, f3 v, float s) { return xscale3(P.use_mtx ? mtx_tmulv(P.mL, v) : xmul3(l_fwd(P), v), tr_scale(P, s)); } // GPF:639
SCN_HD f3 cov_grad_l2w(const Params &P, f3 v, float s) { return xdiv3(P.use_mtx ? mtx_tmulv(P.mW, v) : xmul3(l_inv(P), v), tr_scale(P, s)); }    // GPF:641
SCN_HD float cov_splat_cov1d(const Params &P, f3 pq, f3 pc) { // getNonstationaryCovSplatCov1D
    if (P.stationary) return 1.0f;
    float s = (float)(0.5 * (double)(sqr(cov_ns_scale(P, pq)) + sqr(cov_ns_scale(P, pc))));
    return sqrtf(s);
}
// variance normalisation (GPF:741-760) for the point-dependent (brute-force non-stationary) case
SCN_HD float se_variance3d(const Params &P, float R, int is_identity, float gs, float ls) {
    double per_unit = (double)xdiv(P.density, xmul(xmul(R, R), R));
    double det_sqrt = 1.0;
    if (!is_identity) { det_sqrt = P.use_mtx ? P.mC_dsi : (double)xmul(xmul(P.l_aniso[0], P.l_aniso[1]), P.l_aniso[2]); det_sqrt *= (double)gs * (double)gs * (double)gs; }
    det_sqrt *= (double)ls * (double)ls * (double)ls;
    return (float)(per_unit * (5.568327996831707 * det_sqrt)); // pow(M_PI, 1.5)
}
SCN_HD float se_variance1d(const Params &P, float R, float ls) {
    return (float)((double)xdiv(P.density, R) * (1.7724538509055159 * (double)ls)); // sqrt(M_PI)
}
// getInvCovMtx diagonal (GPF:774-802)
SCN_HD f3 se_inv_cov_diag(const Params &P, int is_cov, int is_iso, float gs, float ls) {
    f3 d = F3(1.f, 1.f, 1.f);
    if (!is_iso) { float g2 = gs * gs; d = F3(P.l_aniso_inv[0] * P.l_aniso_inv[0] / g2, P.l_aniso_inv[1] * P.l_aniso_inv[1] / g2, P.l_aniso_inv[2] * P.l_aniso_inv[2] / g2); }
    float k = (is_cov ? 0.5f : 1.0f), l2 = ls * ls;
    d = F3(d.x * k / l2 * 0.5f, d.y * k / l2 * 0.5f, d.z * k / l2 * 0.5f);
    return d;
}
// splattingKernel3D (GPF:552-563,804-817): value + first gradient for ab = pa - pb and the diagonal M
SCN_HD f4 se_kernel3d(f3 ab, f3 M) {
    float ax = ab.x * M.x, ay = ab.y * M.y, az = ab.z * M.z;
    float f = fast_exp(-(ax * ab.x + (ay * ab.y + az * ab.z)));
    float m2f = -2.f * f;
    return F4(f, ax * m2f, ay * m2f, az * m2f);
}
// splattingKernel3DGrad (GPF:565-577,810-833) contracted with coeff c
SCN_HD f4 se_kernel3d_grad(f3 ab, f3 M, f3 c) {
    float ax = ab.x * M.x, ay = ab.y * M.y, az = ab.z * M.z;
    float f = fast_exp(-(ax * ab.x + (ay * ab.y + az * ab.z)));
    float dx = -2.f * ax, dy = -2.f * ay, dz = -2.f * az;
    float h00 = (dx * dx - 2.f * M.x) * f, h11 = (dy * dy - 2.f * M.y) * f, h22 = (dz * dz - 2.f * M.z) * f;
    float h01 = dx * dy * f, h02 = dx * dz * f, h12 = dy * dz * f;
    return F4((dx * f) * c.x + (dy * f) * c.y + (dz * f) * c.z,
              h00 * c.x + h01 * c.y + h02 * c.z, h01 * c.x + h11 * c.y + h12 * c.z, h02 * c.x + h12 * c.y + h22 * c.z);
}
// ---------------------------------------------------------------- rotation-anisotropic non-stationary kernel (GPF:1678-1690)
// 3x3 inverse in cofactor form (Eigen's fixed-size path), row-major
SCN_HD void mat3_inverse(const float a[9], float r[9]) {
#define SCN_A(i, j) a[((i) % 3) * 3 + ((j) % 3)]
#define SCN_COF(i, j) (SCN_A(i + 1, j + 1) * SCN_A(i + 2, j + 2) - SCN_A(i + 1, j + 2) * SCN_A(i + 2, j + 1))
    float c00 = SCN_COF(0, 0), c10 = SCN_COF(1, 0), c20 = SCN_COF(2, 0);
    float inv = 1.0f / (c00 * a[0] + (c10 * a[3] + c20 * a[6]));
    r[0] = c00 * inv; r[1] = c10 * inv; r[2] = c20 * inv;
    r[3] = SCN_COF(0, 1) * inv; r[4] = SCN_COF(1, 1) * inv; r[5] = SCN_COF(2, 1) * inv;
    r[6] = SCN_COF(0, 2) * inv; r[7] = SCN_COF(1, 2) * inv; r[8] = SCN_COF(2, 2) * inv;
#undef SCN_COF
#undef SCN_A
}
// getNonstationaryAniso3D with compute_ansio_full (GPF:26-39): G = (S V)^T (S V), V = rotation by field(p) * pi/2 about z,
// S = diag(1.5, 1/1.5, 1) (_anisotropyOnAxis, GPF.hpp:2372); row-major
SCN_FN void cov_aniso3d(const Params &P, f3 p_world, float G[9]) {
    const float angle = (float)(field_scalar(P.aniso_f, to_dbl3(p_world)) * (double)(3.1415926536f * 0.5f));
    const float c = cosf(angle), sn = sinf(angle);
    const float ax[3] = {1.5f, 1.f / 1.5f, 1.f};
    const float V[9] = {c, sn, 0.f, -sn, c, 0.f, 0.f, 0.f, 1.f};
    float M[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) M[i * 3 + j] = ax[i] * V[i * 3 + j];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) G[i * 3 + j] = M[i] * M[j] + (M[3 + i] * M[3 + j] + M[6 + i] * M[6 + j]); // Eigen's x0 + (x1 + x2)
}
// The splat kernel's inverse covariance at a query point: diagonal (d) or, with the anisotropy field, the full matrix
// invSigma^T G^-1 invSigma with the isCov / localScale / 0.5 factors of getInvCovMtx (GPF:774-802)
struct KMat { bool full; f3 d; float m[9]; };
SCN_FN KMat kmat_make(const Params &P, f3 p_world, int is_cov, int is_iso, float gs, float ls) {
    KMat K; K.full = P.has_aniso != 0 || (P.use_mtx && !is_iso); K.d = F3(0.f, 0.f, 0.f);
    if (!K.full) { K.d = se_inv_cov_diag(P, is_cov, is_iso, gs, ls); return K; }
    if (!P.has_aniso) { // "useAnisoMtx" (GPF:791-801): _cov_mtx_inv / gs^2, then the isCov / localScale / 0.5 factors
        const float k = (is_cov ? 0.5f : 1.0f), l2 = ls * ls, g2 = gs * gs;
        for (int i = 0; i < 9; ++i) K.m[i] = P.mC[i] / g2 * k / l2 * 0.5f;
        return K;
    }
    float G[9], Gi[9]; cov_aniso3d(P, p_world, G); mat3_inverse(G, Gi);
    float sg[3] = {1.f, 1.f, 1.f};
    if (!is_iso) { sg[0] = P.l_aniso_inv[0] / gs; sg[1] = P.l_aniso_inv[1] / gs; sg[2] = P.l_aniso_inv[2] / gs; }
    const float k = (is_cov ? 0.5f : 1.0f), l2 = ls * ls;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) K.m[i * 3 + j] = (sg[i] * Gi[i * 3 + j]) * sg[j] * k / l2 * 0.5f;
    return K;
}
SCN_HD f4 se_kernel3d(f3 ab, const KMat &K) {
    if (!K.full) return se_kernel3d(ab, K.d);
    const f3 Mab = F3(ab.x * K.m[0] + (ab.y * K.m[3] + ab.z * K.m[6]), ab.x * K.m[1] + (ab.y * K.m[4] + ab.z * K.m[7]), ab.x * K.m[2] + (ab.y * K.m[5] + ab.z * K.m[8]));
    const float f = fast_exp(-(ab.x * Mab.x + (ab.y * Mab.y + ab.z * Mab.z))), m2f = -2.f * f;
    return F4(f, Mab.x * m2f, Mab.y * m2f, Mab.z * m2f);
}
SCN_HD f4 se_kernel3d_grad(f3 ab, const KMat &K, f3 c) {
    if (!K.full) return se_kernel3d_grad(ab, K.d, c);
    const f3 Mab = F3(ab.x * K.m[0] + (ab.y * K.m[3] + ab.z * K.m[6]), ab.x * K.m[1] + (ab.y * K.m[4] + ab.z * K.m[7]), ab.x * K.m[2] + (ab.y * K.m[5] + ab.z * K.m[8]));
    const float f = fast_exp(-(ab.x * Mab.x + (ab.y * Mab.y + ab.z * Mab.z)));
    const float d[3] = {-2.f * Mab.x, -2.f * Mab.y, -2.f * Mab.z}, cc[3] = {c.x, c.y, c.z};
    float o[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < 3; ++i) {
        o[0] += d[i] * f * cc[i];
        for (int j = 0; j < 3; ++j) o[1 + j] += (d[i] * d[j] - 2.f * K.m[i * 3 + j]) * f * cc[i];
    }
    return F4(o[0], o[1], o[2], o[3]);
}
// ---------------------------------------------------------------- alternative splat kernels (GPF:1046-1082, 1142-1153, 1200-1214)
// Modified Bessel functions K0, K1 (the reference calls boost::math::cyl_bessel_k in MaternCovariance, v = 1.5): polynomial
// approximations of Abramowitz & Stegun 9.8.1-9.8.8, |relative error| < 3e-7, evaluated in double.
SCN_HD double bessel_i0_small(double x) { // |x| <= 3.75
    double t = x / 3.75; t *= t;
    return 1.0 + t * (3.5156229 + t * (3.0899424 + t * (1.2067492 + t * (0.2659732 + t * (0.0360768 + t * 0.0045813)))));
}
SCN_HD double bessel_i1_small(double x) {
    double t = x / 3.75; t *= t;
    return x * (0.5 + t * (0.87890594 + t * (0.51498869 + t * (0.15084934 + t * (0.02658733 + t * (0.00301532 + t * 0.00032411))))));
}
SCN_HD double bessel_k0(double x) {
    if (x <= 2.0) {
        double y = x * x / 4.0;
        return -log(x / 2.0) * bessel_i0_small(x) + (-0.57721566 + y * (0.42278420 + y * (0.23069756 + y * (0.03488590 + y * (0.00262698 + y * (0.00010750 + y * 0.0000074))))));
    }
    double y = 2.0 / x;
    return exp(-x) / sqrt(x) * (1.25331414 + y * (-0.07832358 + y * (0.02189568 + y * (-0.01062446 + y * (0.00587872 + y * (-0.00251540 + y * 0.00053208))))));
}
SCN_HD double bessel_k1(double x) {
    if (x <= 2.0) {
        double y = x * x / 4.0;
        return log(x / 2.0) * bessel_i1_small(x) + (1.0 / x) * (1.0 + y * (0.15443144 + y * (-0.67278579 + y * (-0.18156897 + y * (-0.01919402 + y * (-0.00110404 + y * -0.00004686))))));
    }
    double y = 2.0 / x;
    return exp(-x) / sqrt(x) * (1.25331414 + y * (0.23498619 + y * (-0.03655620 + y * (0.01504268 + y * (-0.00780353 + y * (0.00325614 + y * -0.00068245))))));
}
// splattingKernel3D for "matern" / "gabor_aniso" / "gabor_iso": value + first gradient for ab = pa - pb.  The reference evaluates
// these in double and rounds each factor to float where it casts; isCov / isIsotropic / scales are ignored by these classes.
SCN_FN f4 alt_kernel3d(const Params &P, f3 ab) {
    const float lsq = xadd(xadd(xmul(ab.x, ab.x), xmul(ab.y, ab.y)), xmul(ab.z, ab.z));
    const float rf = xsqrt(lsq);
    if (P.kernel == SCN_COV_MATERN) { // GPF:1046-1082
        const double r = (double)rf, l = (double)P.k_l;
        float val, gk;
        if (P.k_v == 0.5f) {
            double e = exp(-r / l);
            val = (float)(e / r);
            gk = -(float)(e * (1.0 / (r * r * r) - 1.0 / (r * r * l))); // sic: the reference's sign between the two terms
        } else if (P.k_v == 1.5f) {
            double rs = 1.7320508075688772 * r / l;
            val = (float)bessel_k0(rs);
            gk = -(float)(bessel_k1(rs) * 1.7320508075688772 / l / r);
        } else {
            double e = exp(-2.23606797749979 * r / l);
            val = (float)e;
            gk = -(float)(e * 2.23606797749979 / l / r);
        }
        return F4(val, gk * ab.x, gk * ab.y, gk * ab.z);
    }
    const double PI_D = 3.14159265358979323846;
    const float a = P.k_a, f = P.k_f;
    if (P.kernel == SCN_COV_GABOR_ANISO) { // GPF:1142-1153
        const float od = xadd(xadd(xmul(P.k_omega[0], ab.x), xmul(P.k_omega[1], ab.y)), xmul(P.k_omega[2], ab.z));
        const double ph = 2.0 * PI_D * (double)f * (double)od;
        const double env = exp(-PI_D * (double)(a * a) * (double)lsq);
        const float A = (float)env, B = (float)cos(ph);
        const float val = (float)(env * cos(ph));
        const float g1 = -(float)((double)A * sin(ph) * 2.0 * PI_D * (double)f);
        const float g2 = (float)((double)B * (double)A * 2.0 * PI_D * (double)(a * a));
        return F4(val, g1 * P.k_omega[0] - g2 * ab.x, g1 * P.k_omega[1] - g2 * ab.y, g1 * P.k_omega[2] - g2 * ab.z);
    }
    // gabor_iso, GPF:1200-1214
    const double r = (double)rf;
    const double ph = 2.0 * PI_D * (double)f * r;
    const float ar = a * rf;
    const float val = (float)(exp(-PI_D * (double)(ar * ar)) * 2.0 * (double)f / r * sin(ph));
    const double a2 = (double)(a * a);
    const float gk = (float)(2.0 * (double)f * exp(-PI_D * a2 * (double)lsq) *
                             (-sin(ph) / (r * r * r) - 2.0 * PI_D * a2 * sin(ph) / r + 2.0 * PI_D * (double)f * cos(ph) / (double)(rf * rf)));
    return F4(val, gk * ab.x, gk * ab.y, gk * ab.z);
}
// 1D kernels (GPF:835-865)
SCN_HD float se_k1d_val(float ab, float ls) { float den = 2.0f * ls * ls; return fast_exp(-(ab * ab / den)); }
SCN_HD float se_k1d_grad(float ab, float ls) { float den = 2.0f * ls * ls; return -2.0f * ab / den * se_k1d_val(ab, ls); }
SCN_HD float se_c1d_val(float ab, float ls) { return se_k1d_val(ab, ls * 1.41421356237f); }
SCN_HD float se_c1d_grad(float ab, float ls) { return se_k1d_grad(ab, ls * 1.41421356237f); }
SCN_HD float se_c1d_grad2(float ab, float ls) { float den = 4.0f * ls * ls; float g1 = -2.0f * ab / den; return (g1 * g1 - 2.0f / den) * se_c1d_val(ab, ls); }
SCN_HD float se_c1d_grad2_normal(float ab, float ls) { float den = 4.0f * ls * ls; return -2.0f / den * se_c1d_val(ab, ls); }

// ---------------------------------------------------------------- mean (GPF.hpp:867-1005, GaussianProcess.cpp:379-393) in double
// ProceduralMean::mean (math/GPFunctions.cpp:279-285): the point goes to float, through the inverse config transform
// (Mat4f::transformPoint, math/Mat4f.hpp:168-175), into the SDF function (math/SdfFunctions.cpp) or the scalar field; float scale / offset / floor.
// ---- the material-test knob (math/SdfFunctions.cpp:66-179), float arithmetic in the reference's order; double where its literals promote
SCN_HD float sdf_min(float a, float b) { return a < b ? a : b; } // math/MathUtil.hpp:14-29
SCN_HD float sdf_max(float a, float b) { return a > b ? a : b; }
SCN_HD float sdf_len2(float x, float y) { return xsqrt(xadd(xmul(x, x), xmul(y, y))); }
SCN_HD float sdf_len3(f3 v) { return xsqrt(xlensq(v)); }
SCN_HD f3 sdf_rotx(const float *r, f3 b) { // Mat4f::rotAxis((1,0,0), .).transformVector (math/Mat4f.hpp:159-166): rows (a11,0,0), (0,c,-s), (0,s,c)
    return F3(xadd(xadd(xmul(r[0], b.x), xmul(0.f, b.y)), xmul(0.f, b.z)), xadd(xadd(xmul(0.f, b.x), xmul(r[1], b.y)), xmul(-r[2], b.z)),
              xadd(xadd(xmul(0.f, b.x), xmul(r[2], b.y)), xmul(r[1], b.z)));
}
SCN_HD float sdf_cone(f3 p, float c) { return xadd(xmul(c, sdf_len2(p.x, p.y)), xmul(c, p.z)); }
SCN_HD float sdf_capped_cylinder(f3 p, float h, float r) {
    const float dx = xsub(fabsf(sdf_len2(p.x, p.z)), h), dy = xsub(fabsf(p.y), r);
    return xadd(sdf_min(sdf_max(dx, dy), 0.0f), sdf_len2(sdf_max(dx, 0.0f), sdf_max(dy, 0.0f)));
}
SCN_HD float sdf_tri_prism(f3 p, float hx, float hy) {
    const f3 q = F3(fabsf(p.x), fabsf(p.y), fabsf(p.z));
    return sdf_max(xsub(q.z, hy), xsub(sdf_max(xadd(xmul(q.x, 0.866025f), xmul(p.y, 0.5f)), -p.y), xmul(hx, 0.5f)));
}
SCN_HD float sdf_smooth_union(float d1, float d2, float k) { // opSmoothUnion: lerp(d2, d1, h) - k h (1.0 - h), the last factor in double
    float h = xadd(0.5f, xdiv(xmul(0.5f, xsub(d2, d1)), k)); h = sdf_min(sdf_max(h, 0.0f), 1.0f);
    const float l = xadd(xmul(d2, xsub(1.0f, h)), xmul(d1, h));
    return (float)ysub((double)l, ymul((double)xmul(k, h), ysub(1.0, (double)h)));
}
SCN_HD float sdf_ssub(float d1, float d2, float k) { // ssub: h in double, lerp(d2, -d1, h) + k h (1.0 - h)
    double hd = ysub(0.5, ydiv(ymul(0.5, (double)xadd(d2, d1)), (double)k)); hd = hd > 0.0 ? hd : 0.0; hd = hd < 1.0 ? hd : 1.0;
    const float h = (float)hd;
    const float l = xadd(xmul(d2, xsub(1.0f, h)), xmul(-d1, h));
    return (float)yadd((double)l, ymul((double)xmul(k, h), ysub(1.0, (double)h)));
}
SCN_FN float sdf_base(const float *kn, f3 p) { // sdBase
    float base = sdf_smooth_union(sdf_cone(sdf_rotx(kn + 0, F3(p.x, xadd(p.y, .9f), p.z)), kn[9]), sdf_cone(sdf_rotx(kn + 3, F3(p.x, xsub(p.y, .9f), p.z)), kn[10]), 0.02f);
    base = xmul(sdf_max(base, sdf_capped_cylinder(p, 1.1f, 0.25f)), 0.7f);
    base = sdf_max(-sdf_capped_cylinder(p, 0.6f, 0.3f), base);
    base = sdf_max(-sdf_tri_prism(sdf_rotx(kn + 3, F3(p.x, p.y, xadd(p.z, -1.f))), 1.2f, 0.3f), base);
    return base;
}
SCN_FN float sdf_knob(const float *kn, f3 p0, bool outer) { // SdfFunctions::knob / knob_outer over sdKnob / sdKnobOuter
    const f3 p = xscale3(p0, 1.25f); // p *= 1. / 0.8f
    const float sphere = xsub(sdf_len3(p), 1.0f), cutout = xsub(sdf_len3(F3(p.x, xsub(p.y, 0.5f), xsub(p.z, 0.5f))), 0.7f);
    const f3 te = sdf_rotx(kn + 6, F3(p.x, xsub(p.y, 0.2f), xsub(p.z, 0.2f)));
    const float etch = xsub(sdf_len2(xsub(sdf_len2(te.x, te.z), 1.0f), te.y), 0.05f); // sdTorus(., (1, 0.05))
    const float inner = xsub(sdf_len3(p), 0.75f);
    float d = sdf_ssub(cutout, sphere, 0.1f);
    d = outer ? sdf_max(d, -inner) : sdf_min(d, inner);
    d = sdf_max(-etch, d);
    d = sdf_min(sdf_ssub(sphere, sdf_base(kn, F3(p.x, xsub(p.y, -.775f), p.z)), 0.1f), d);
    return xmul(d, 0.8f);
}
SCN_FN double proc_mean(const Params &P, const scn_mean_desc &md, const FieldK &f, dbl3 a) {
    const float px = (float)a.x, py = (float)a.y, pz = (float)a.z;
    const float *T = md.inv_transform;
    const f3 p = F3(xadd(xadd(xadd(xmul(T[0], px), xmul(T[1], py)), xmul(T[2], pz)), T[3]), xadd(xadd(xadd(xmul(T[4], px), xmul(T[5], py)), xmul(T[6], pz)), T[7]),
                    xadd(xadd(xadd(xmul(T[8], px), xmul(T[9], py)), xmul(T[10], pz)), T[11]));
    float m;
    switch (md.sdf_func) {
    case SCN_SDF_PLANE: m = p.y; break;
    case SCN_SDF_TWO_SPHERES: { // min(|p - (0, 10, 0)| - 9.5, |p - (0, -10, 0)| - 9.5)
        const float a1 = xsub(xsqrt(xlensq(F3(p.x, xsub(p.y, 10.f), p.z))), 9.5f), a2 = xsub(xsqrt(xlensq(F3(p.x, xsub(p.y, -10.f), p.z))), 9.5f);
        m = a1 < a2 ? a1 : a2; break;
    }
    case SCN_SDF_KNOB_INNER: { // p *= 1. / 0.8f (1.25f after the conversion to float); (|p| - 0.75) * 0.8f
        const f3 q = xscale3(p, 1.25f);
        m = xmul(xsub(xsqrt(xlensq(q)), 0.75f), 0.8f); break;
    }
    case SCN_SDF_KNOB: m = sdf_knob(P.knob_rot, p, false); break;
    case SCN_SDF_KNOB_OUTER: m = sdf_knob(P.knob_rot, p, true); break;
    default: m = (float)field_scalar(f, to_dbl3(p)); break;
    }
    m = xmul(m, md.scale);
    const float r = xadd(m, md.offset);
    return (double)(md.min > r ? md.min : r);
}
// PROC = false: the cached march kernels (scn_march.cuh, scn_march1d.cuh) are compiled without the procedural mean — a reachable call
// into the procedural-field code costs them their spill-free register allocation (406 B of spills in the stationary 3D kernel, -15 %) —
// and media with such a mean are served by the generic kernels (march_block_size / march1d_supported).
template <bool PROC = true>
SCN_HD double mean_eval(const Params &P, const scn_mean_desc &md, const FieldK &mf, dbl3 a) {
    switch (md.type) {
    case SCN_MEAN_PROCEDURAL: return PROC ? proc_mean(P, md, mf, a) : 0.;
    case SCN_MEAN_HOMOGENEOUS: return (double)md.offset;
    case SCN_MEAN_SPHERICAL: {
        double qx = ysub(a.x, md.center[0]), qy = ysub(a.y, md.center[1]), qz = ysub(a.z, md.center[2]);
        return ysub(ysqrt(yadd(yadd(ymul(qx, qx), ymul(qy, qy)), ymul(qz, qz))), (double)md.radius);
    }
    case SCN_MEAN_LINEAR: {
        double qx = ysub(a.x, md.ref[0]), qy = ysub(a.y, md.ref[1]), qz = ysub(a.z, md.ref[2]);
        double v = ymul(yadd(yadd(ymul(qx, md.dir[0]), ymul(qy, md.dir[1])), ymul(qz, md.dir[2])), (double)md.scale), mn = (double)md.min;
        return v > mn ? v : mn;
    }
    default: return 0.;
    }
}
template <bool PROC = true>
SCN_HD f3 mean_grad(const Params &P, const scn_mean_desc &md, const FieldK &mf, dbl3 a) {
    switch (md.type) {
    case SCN_MEAN_PROCEDURAL: if (PROC) { // ProceduralMean::dmean_da (math/GPFunctions.cpp:287-296): forward differences, eps = double(0.001f)
        const double eps = (double)0.001f;
        dbl3 ax = a, ay = a, az = a; ax.x = yadd(a.x, eps); ay.y = yadd(a.y, eps); az.z = yadd(a.z, eps);
        const double v0 = proc_mean(P, md, mf, a), v1 = proc_mean(P, md, mf, ax), v2 = proc_mean(P, md, mf, ay), v3 = proc_mean(P, md, mf, az);
        return F3((float)ydiv(ysub(v1, v0), eps), (float)ydiv(ysub(v2, v0), eps), (float)ydiv(ysub(v3, v0), eps));
    } else return F3(0.f, 0.f, 0.f);
    case SCN_MEAN_SPHERICAL: {
        double qx = ysub(a.x, md.center[0]), qy = ysub(a.y, md.center[1]), qz = ysub(a.z, md.center[2]);
        double inv = ydiv(1.0, ysqrt(yadd(yadd(ymul(qx, qx), ymul(qy, qy)), ymul(qz, qz))));
        return F3((float)ymul(qx, inv), (float)ymul(qy, inv), (float)ymul(qz, inv));
    }
    case SCN_MEAN_LINEAR: {
        double qx = ysub(a.x, md.ref[0]), qy = ysub(a.y, md.ref[1]), qz = ysub(a.z, md.ref[2]);
        double v = ymul(yadd(yadd(ymul(qx, md.dir[0]), ymul(qy, md.dir[1])), ymul(qz, md.dir[2])), (double)md.scale);
        if (v < (double)md.min) return F3(0.f, 0.f, 0.f);
        return F3((float)ymul(md.dir[0], (double)md.scale), (float)ymul(md.dir[1], (double)md.scale), (float)ymul(md.dir[2], (double)md.scale));
    }
    default: return F3(0.f, 0.f, 0.f);
    }
}
template <bool PROC = true>
SCN_HD void gp_mean_and_id(const Params &P, f3 p, double &mean, int &id) {
    dbl3 a = to_dbl3(p);
    mean = mean_eval<PROC>(P, P.mean, P.mean_f, a); id = P.gp_id;
    if (P.has_mean_add) { double v = mean_eval<PROC>(P, P.mean_add, P.mean_add_f, a); if (v < mean) { mean = v; id = P.id_add; } }
}
template <bool PROC = true>
SCN_HD f3 gp_mean_grad(const Params &P, f3 p, int id) { return id == P.gp_id ? mean_grad<PROC>(P, P.mean, P.mean_f, to_dbl3(p)) : mean_grad<PROC>(P, P.mean_add, P.mean_add_f, to_dbl3(p)); }
SCN_HD bool has_procedural_mean(const Params &P) { return P.mean.type == SCN_MEAN_PROCEDURAL || (P.has_mean_add && P.mean_add.type == SCN_MEAN_PROCEDURAL); }

// ---------------------------------------------------------------- seeds and multi-resolution levels
SCN_HD uint32_t compute_seed(const Params &P, const Info &in) { // SCN:40-49
    uint32_t seed = P.global_seed;
    if (!P.single_realization) seed += xxhash32_4(in.px, in.py, in.spp, P.ctx == SCN_CTX_GLOBAL ? 0u : in.seg) + xxhash32_1(in.scene_seed);
    return seed;
}
struct Levels { float s_lo, s_hi, r_lo, r_hi; int k_lo; };
// kernelScaleLevelRatio (SCN:60-70).  The level index is floor(logf(ls)/logf(2.5f)) on the CPU; the device takes
// log in double and rounds to float (== the correctly rounded logf in all but double-rounding corner cases), and
// reads 2.5^k from the host-computed table so that the grid scaling matches the reference's powf bit for bit.
template <bool FULL = true>
SCN_FN Levels level_ratio(const Params &P, f3 p) {
    float ls = cov_lateral_scale<FULL>(P, p);
    float lg = (float)log((double)ls);
    float lo = floorf(xdiv(lg, P.log_base)); // logf(2.5f) from the host libm
    int k = (int)lo;
    int idx = k - LVL_MIN; idx = idx < 0 ? 0 : (idx > LVL_COUNT - 2 ? LVL_COUNT - 2 : idx);
    Levels L; L.k_lo = idx; L.s_lo = P.lvl_scale[idx]; L.s_hi = P.lvl_scale[idx + 1];
    float frac = xdiv(xsub(ls, L.s_lo), xsub(L.s_hi, L.s_lo));
    double fd = (double)frac, den = sqrt(1.0 - 2.0 * fd + 2.0 * fd * fd);
    L.r_lo = (float)((1.0 - fd) / den); L.r_hi = (float)(fd / den);
    return L;
}

// ---------------------------------------------------------------- cell3D / noise3D (SCN:362-395)
// Impulse k of a cell: position draws z, y, x (g++ evaluates next3D's arguments right to left), then the sign
// draw.  `M` is the diagonal of the kernel's inverse covariance in the units of R*p (GPF:774-802).
SCN_FN f4 noise3d(const Params &P, f3 p, uint32_t seed, float R, const KMat &M) {
    f3 pg = xdiv3(p, R);
    f3 fl = F3(floorf(pg.x), floorf(pg.y), floorf(pg.z));
    f3 frac = xsub3(pg, fl);
    int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z;
    const uint32_t n = P.n_imp;
    f4 sum = F4(0.f, 0.f, 0.f, 0.f);
    for (int dx = -1; dx <= 1; ++dx) {
        float cx = xsub(frac.x, (float)dx);
        for (int dy = -1; dy <= 1; ++dy) {
            float cy = xsub(frac.y, (float)dy);
            for (int dz = -1; dz <= 1; ++dz) {
                float cz = xsub(frac.z, (float)dz);
                Pcg g; g.seed32(xxhash32_4((uint32_t)(iz + dz), (uint32_t)(iy + dy), (uint32_t)(ix + dx), seed));
                f4 cs = F4(0.f, 0.f, 0.f, 0.f);
                for (uint32_t k = 0; k < n; ++k) {
                    float pz = g.next(), py = g.next(), px = g.next(), ws = g.next();
                    float tx = xsub(cx, px), ty = xsub(cy, py), tz = xsub(cz, pz);
                    float l2 = xadd(xadd(xmul(tx, tx), xmul(ty, ty)), xmul(tz, tz));
                    if (l2 < 1.0f) {
                        // kernel(R*p, R*p_i): ab = R*p - R*p_i (SCN:390)
                        f3 ab = F3(xsub(xmul(R, cx), xmul(R, px)), xsub(xmul(R, cy), xmul(R, py)), xsub(xmul(R, cz), xmul(R, pz)));
                        f4 kv = P.kernel == SCN_COV_SQUARED_EXPONENTIAL ? se_kernel3d(ab, M) : alt_kernel3d(P, ab);
                        float w = ws < 0.5f ? -1.f : 1.f;
                        cs.x += w * kv.x; cs.y += w * kv.y; cs.z += w * kv.z; cs.w += w * kv.w;
                    }
                }
                sum = add4(sum, cs);
            }
        }
    }
    return sum;
}
// cell1D / noise1D (SCN:397-428) -> (val, grad); ls = kernel local scale at the query point
SCN_FN void noise1d(const Params &P, float t, uint32_t seed, float R, float ls, float &val, float &grad) {
    float tg = xdiv(t, R); float fl = floorf(tg); float frac = xsub(tg, fl); int i = (int)fl;
    float sv = 0.f, sg = 0.f; const uint32_t n = P.n_imp;
    for (int dx = -1; dx <= 1; ++dx) {
        float tc = xsub(frac, (float)dx); float cv = 0.f, cg = 0.f;
        Pcg g; g.seed32(xxhash32_2((uint32_t)(i + dx), seed));
        for (uint32_t k = 0; k < n; ++k) {
            float ti = g.next(), ws = g.next();
            float tp = xsub(tc, ti);
            if (xmul(tp, tp) < 1.0f) {
                float ab = xsub(xmul(R, tc), xmul(R, ti));
                float w = ws < 0.5f ? -1.f : 1.f;
                float kv = se_k1d_val(ab, ls);
                cv += w * kv; cg += w * (-2.0f * ab / (2.0f * ls * ls) * kv);
            }
        }
        sv += cv; sg += cg;
    }
    val = sv; grad = sg;
}

// ---------------------------------------------------------------- realization
struct Real { Coeff c; }; // conditioning coefficients travel with the ray (registers)

// evaluateNoise3DNormalized — world space (SCN:251-260)
SCN_FN f4 noise3d_world(const Params &P, const Coeff &c, f3 p, uint32_t seed, float R, float kss, int add_seed, float nf, bool conditioning) {
    float ls = cov_ns_scale(P, p);
    KMat M = kmat_make(P, p, 0, 0, kss, ls);
    f4 n = noise3d(P, p, seed + (uint32_t)add_seed, R, M);
    n = F4(n.x / nf, n.y / nf, n.z / nf, n.w / nf);
    if (P.activate_conditioning && conditioning) {
        KMat Mc = kmat_make(P, p, 1, 0, kss, ls);
        f3 ab = F3(p.x - c.org.x, p.y - c.org.y, p.z - c.org.z);
        n = add4(n, add4(scale4(se_kernel3d(ab, Mc), c.vs), se_kernel3d_grad(ab, Mc, c.gs)));
    }
    return n;
}
// evaluateNoise3DIsotropicRayNormalized (SCN:291-320)
SCN_FN f4 noise3d_isoray(const Params &P, const Coeff &c, const Frame &fr, f3 p, uint32_t seed, float kss, int add_seed, bool conditioning) {
    f3 p_iso = cov_pos_w2l(P, p, kss);
    f3 p_ray = to_local(fr, p_iso);
    float ls = cov_ns_scale(P, p);
    KMat M = kmat_make(P, p, 0, 1, 1.0f, ls);
    f4 nr = noise3d(P, p_ray, seed + (uint32_t)add_seed, P.R_iso, M);
    f3 g_iso = to_global(fr, F3(nr.y, nr.z, nr.w));
    f3 gw = cov_grad_l2w(P, g_iso, kss);
    float nf = (P.stationary || P.multires) ? P.nf_iso : sqrtf(se_variance3d(P, P.R_iso, 1, 1.0f, ls));
    f4 n = F4(nr.x / nf, gw.x / nf, gw.y / nf, gw.z / nf);
    if (P.activate_conditioning && conditioning) {
        f3 o_ray = to_local(fr, cov_pos_w2l(P, c.org, kss));
        KMat Mc = kmat_make(P, p, 1, 1, 1.0f, ls);
        f3 ab = F3(p_ray.x - o_ray.x, p_ray.y - o_ray.y, p_ray.z - o_ray.z);
        f4 d = add4(scale4(se_kernel3d(ab, Mc), c.vs), se_kernel3d_grad(ab, Mc, c.gs));
        f3 gdw = cov_grad_l2w(P, to_global(fr, F3(d.y, d.z, d.w)), kss);
        n = add4(n, F4(d.x, gdw.x, gdw.y, gdw.z));
    }
    return n;
}
SCN_HD Frame ray_frame(const Params &P, f3 dir) { return make_frame(xnormalized(cov_pos_w2l(P, dir, 1.0f))); }

// evaluateNoise3D (SCN:101-124).  `fr` = ray_frame(dir) (only read in isotropic mode).
SCN_FN f4 eval_noise3d(const Params &P, const Coeff &c, const Frame &fr, f3 p, uint32_t seed, bool conditioning) {
    if (!P.iso3d) {
        if (!P.multires) {
            float nf = P.stationary ? P.nf_world : sqrtf(se_variance3d(P, P.R_world, 0, P.world_scale, cov_ns_scale(P, p)));
            return noise3d_world(P, c, p, seed, P.R_world, P.world_scale, P.add_seed_world, nf, conditioning);
        }
        Levels L = level_ratio(P, p);
        f4 lo = noise3d_world(P, c, p, seed, P.lvl_R_world[L.k_lo], L.s_lo, P.lvl_add_seed[L.k_lo], P.lvl_nf_world[L.k_lo], conditioning);
        f4 hi = noise3d_world(P, c, p, seed, P.lvl_R_world[L.k_lo + 1], L.s_hi, P.lvl_add_seed[L.k_lo + 1], P.lvl_nf_world[L.k_lo + 1], conditioning);
        return add4(scale4(lo, L.r_lo), scale4(hi, L.r_hi));
    }
    if (!P.multires) return noise3d_isoray(P, c, fr, p, seed, 1.0f, 0, conditioning);
    Levels L = level_ratio(P, p);
    f4 lo = noise3d_isoray(P, c, fr, p, seed, L.s_lo, P.lvl_add_seed[L.k_lo], conditioning);
    f4 hi = noise3d_isoray(P, c, fr, p, seed, L.s_hi, P.lvl_add_seed[L.k_lo + 1], conditioning);
    return add4(scale4(lo, L.r_lo), scale4(hi, L.r_hi));
}

// evaluateNoise1DNormalized (SCN:322-360) -> (val, gx_cond, gy_cond, gz)
SCN_FN f4 noise1d_norm(const Params &P, const Coeff &c, const Frame &fr, f3 p, uint32_t seed, float kss, int add_seed, bool conditioning, bool low_level) {
    f3 p_ray = to_local(fr, cov_pos_w2l(P, p, kss));
    float lsq = cov_ns_scale(P, p);
    float nv, ng;
    noise1d(P, p_ray.z, seed + (uint32_t)add_seed, P.R_iso, P.stationary ? 1.0f : lsq, nv, ng);
    float nf = (P.stationary || P.multires) ? P.nf_1d : sqrtf(se_variance1d(P, P.R_iso, lsq));
    nv /= nf; ng /= nf;
    float gcx = 0.f, gcy = 0.f;
    if (P.activate_conditioning && conditioning) {
        float osf = 1.0f;
        if (P.multires) { Levels oi = level_ratio(P, c.org); osf = low_level ? oi.r_lo : oi.r_hi; }
        f3 o_ray = to_local(fr, cov_pos_w2l(P, c.org, kss));
        float ab = p_ray.z - o_ray.z;
        float cls = P.stationary ? 1.0f : cov_splat_cov1d(P, p, c.org);
        float gzc = kss * c.gs.z;
        nv += osf * (c.vs * se_c1d_val(ab, cls) + gzc * se_c1d_grad(ab, cls));
        ng += osf * (c.vs * se_c1d_grad(ab, cls) + gzc * se_c1d_grad2(ab, cls));
        if (P.corr_xy) {
            float k2 = se_c1d_grad2_normal(ab, cls);
            gcx = osf * kss * c.gs.x * k2; gcy = osf * kss * c.gs.y * k2;
        }
    }
    return F4(nv, gcx, gcy, ng);
}
// evaluateValueNoise1D (SCN:126-135)
SCN_FN float value_noise1d(const Params &P, const Coeff &c, const Frame &fr, f3 p, uint32_t seed, bool conditioning) {
    if (!P.multires) return noise1d_norm(P, c, fr, p, seed, 1.0f, 0, conditioning, false).x;
    Levels L = level_ratio(P, p);
    float lo = noise1d_norm(P, c, fr, p, seed, L.s_lo, P.lvl_add_seed[L.k_lo], conditioning, true).x;
    float hi = noise1d_norm(P, c, fr, p, seed, L.s_hi, P.lvl_add_seed[L.k_lo + 1], conditioning, false).x;
    return L.r_lo * lo + L.r_hi * hi;
}
// Box-Muller pair in double (sampling/Gaussian.cpp:21-33); 2*PI is the float constant 6.2831855f
SCN_HD void rand_normal_2(Pcg &g, double &z1, double &z2) {
    double u1 = (double)g.next(), u2 = (double)g.next();
    double r = sqrt(-2.0 * log(1.0 - u1));
    double a = (double)6.28318548202514648f * u2;
    z1 = r * cos(a); z2 = r * sin(a);
}
// evaluateGradientNoise1D (SCN:137-239)
SCN_FN f3 gradient_noise1d(const Params &P, const Coeff &c, const Frame &fr, f3 p, float t_segment, f3 dir, const Info &in, bool conditioning) {
    uint32_t seed = compute_seed(P, in);
    float sxy = cov_splat_cov1d(P, p, p);
    double xs = 1.0 / (double)sxy;
    const double sqrt2 = (double)1.41421354f;
    bool plus_xy = P.ctx == SCN_CTX_RENEWAL_PLUS && P.corr_xy;
    Pcg g; g.seed32(xxhash32_4(in.px, in.py, in.spp, in.seg) + xxhash32_1(in.scene_seed));
    if (!P.multires) {
        f4 nb = noise1d_norm(P, c, fr, p, seed, 1.0f, 0, conditioning, false);
        double a0, a1; rand_normal_2(g, a0, a1); a0 /= sqrt2; a1 /= sqrt2;
        double x0, x1;
        if (t_segment == 0.f) { x0 = a0 * xs; x1 = a1 * xs; }
        else if (!plus_xy) { double b0, b1; rand_normal_2(g, b0, b1); x0 = b0 / sqrt2 * xs; x1 = b1 / sqrt2 * xs; }
        else {
            float factor = (float)(exp((double)(-t_segment * t_segment) / 4.0) * (0.5 - (double)(t_segment * t_segment) * 0.25));
            float cov = (float)(0.5 - (double)(factor * factor * 2.f));
            double b0, b1; rand_normal_2(g, b0, b1); double sc = (double)sqrtf(cov);
            x0 = b0 * sc + (double)factor * 2.0 * a0; x1 = b1 * sc + (double)factor * 2.0 * a1;
        }
        f3 gi = to_global(fr, F3((float)x0 + nb.y, (float)x1 + nb.z, 0.f + nb.w));
        return cov_grad_l2w(P, gi, 1.0f);
    }
    Levels L = level_ratio(P, p);
    f4 nlo = noise1d_norm(P, c, fr, p, seed, L.s_lo, P.lvl_add_seed[L.k_lo], conditioning, true);
    f4 nhi = noise1d_norm(P, c, fr, p, seed, L.s_hi, P.lvl_add_seed[L.k_lo + 1], conditioning, false);
    double l0, l1, h0, h1; rand_normal_2(g, l0, l1); l0 /= sqrt2; l1 /= sqrt2; rand_normal_2(g, h0, h1); h0 /= sqrt2; h1 /= sqrt2;
    double xl0, xl1, xh0, xh1;
    if (t_segment == 0.f) { xl0 = l0 * xs; xl1 = l1 * xs; xh0 = h0 * xs; xh1 = h1 * xs; }
    else if (!plus_xy) {
        double a0, a1, b0, b1; rand_normal_2(g, a0, a1); rand_normal_2(g, b0, b1);
        xl0 = a0 / sqrt2 * xs; xl1 = a1 / sqrt2 * xs; xh0 = b0 / sqrt2 * xs; xh1 = b1 / sqrt2 * xs;
    } else {
        float factor = (float)(exp((double)(-t_segment * t_segment) / 4.0) * (0.5 - (double)(t_segment * t_segment) * 0.25));
        float cov = (float)(0.5 - (double)(factor * factor * 2.f)); double sc = (double)sqrtf(cov), f2 = (double)factor * 2.0;
        double a0, a1, b0, b1; rand_normal_2(g, a0, a1); rand_normal_2(g, b0, b1);
        xl0 = a0 * sc + f2 * l0; xl1 = a1 * sc + f2 * l1; xh0 = b0 * sc + f2 * h0; xh1 = b1 * sc + f2 * h1;
    }
    // the two levels share the frame: transformPosDirWorldtoLocal(dir, s).normalized() is scale invariant up to rounding
    Frame flo = make_frame(xnormalized(cov_pos_w2l(P, dir, L.s_lo))), fhi = make_frame(xnormalized(cov_pos_w2l(P, dir, L.s_hi)));
    f3 gwlo = cov_grad_l2w(P, to_global(flo, F3((float)xl0 + nlo.y, (float)xl1 + nlo.z, nlo.w)), L.s_lo);
    f3 gwhi = cov_grad_l2w(P, to_global(fhi, F3((float)xh0 + nhi.y, (float)xh1 + nhi.z, nhi.w)), L.s_hi);
    return F3(L.r_lo * gwlo.x + L.r_hi * gwhi.x, L.r_lo * gwlo.y + L.r_hi * gwhi.y, L.r_lo * gwlo.z + L.r_hi * gwhi.z);
}

// evaluateValue (SCN:73-89)
SCN_FN float eval_value(const Params &P, const Coeff &c, const Frame &fr, f3 p, uint32_t seed, int &gp_id) {
    float amp = cov_amplitude(P, p);
    float nv = P.sampling_1d ? value_noise1d(P, c, fr, p, seed, true) : eval_noise3d(P, c, fr, p, seed, true).x;
    double mean; int id; gp_mean_and_id(P, p, mean, id);
    if (P.surf_vol_sep) id = 1.f < P.surf_vol_thresh ? 0 : 1;
    gp_id = id;
    return (float)yadd((double)xmul(amp, nv), mean);
}
// evaluateGradient (SCN:92-99)
SCN_FN f3 eval_gradient(const Params &P, const Coeff &c, const Frame &fr, f3 p, float t, f3 dir, const Info &in, uint32_t seed) {
    float amp = cov_amplitude(P, p);
    f3 ng;
    if (P.sampling_1d) ng = gradient_noise1d(P, c, fr, p, t, dir, in, true);
    else { f4 n = eval_noise3d(P, c, fr, p, seed, true); ng = F3(n.y, n.z, n.w); }
    double mean; int id; gp_mean_and_id(P, p, mean, id);
    f3 mg = gp_mean_grad(P, p, id);
    return F3(amp * ng.x + mg.x, amp * ng.y + mg.y, amp * ng.z + mg.z);
}


// sparseConvNoiseOneOverSecondDerivative with the anisotropy field (GPF:1239-1243, 713-739): (-2 * 0.25 * invSigma^T G^-1 invSigma)^-1 v
SCN_FN f3 one_over_second_derivative_aniso(const Params &P, f3 p_world, int is_iso, f3 v) {
    float G[9], Gi[9], a[9], r[9]; cov_aniso3d(P, p_world, G); mat3_inverse(G, Gi);
    float sg[3] = {1.f, 1.f, 1.f};
    if (!is_iso) { sg[0] = P.l_aniso_inv[0]; sg[1] = P.l_aniso_inv[1]; sg[2] = P.l_aniso_inv[2]; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i * 3 + j] = -2.f * (((sg[i] * Gi[i * 3 + j]) * sg[j]) * 0.25f);
    mat3_inverse(a, r);
    return F3(r[0] * v.x + (r[1] * v.y + r[2] * v.z), r[3] * v.x + (r[4] * v.y + r[5] * v.z), r[6] * v.x + (r[7] * v.y + r[8] * v.z));
}

// conditioning3D / conditioning1D (SCN:431-595).  The reference's residual re-evaluations only print.
SCN_FN void conditioning(const Params &P, Coeff &c, const Frame &fr, f3 p, f3 dir, float target_val, f3 target_grad, const Info &in) {
    if (!P.activate_conditioning) return;
    c.vs = 0.f; c.gs = F3(0.f, 0.f, 0.f); c.org = p;
    Levels L; L.s_lo = L.s_hi = L.r_lo = L.r_hi = 0.f; L.k_lo = 0;
    if (P.multires) L = level_ratio(P, p);
    uint32_t seed = compute_seed(P, in);
    float amp = cov_amplitude(P, p);
    if (amp == 0.f) return;
    double mean; int id; gp_mean_and_id(P, p, mean, id);
    float meanf = (float)mean;
    if (!P.sampling_1d) {
        f4 n0 = eval_noise3d(P, c, fr, p, seed, false);
        c.vs = (target_val - meanf) / amp - n0.x;
        if (P.multires) c.vs /= L.r_lo + L.r_hi;
        if (P.ctx == SCN_CTX_RENEWAL_PLUS) {
            f3 mg = gp_mean_grad(P, p, id);
            f3 delta = F3((target_grad.x - mg.x) / amp - n0.y, (target_grad.y - mg.y) / amp - n0.z, (target_grad.z - mg.z) / amp - n0.w);
            if (P.iso3d) {
                f3 gs = to_local(fr, cov_grad_w2l(P, delta, 1.0f));
                // (-2 * 0.25 * I)^-1 = -2 I (GPF:713-739)
                c.gs = F3(-2.f * gs.x, -2.f * gs.y, -2.f * gs.z);
                if (P.has_aniso) c.gs = one_over_second_derivative_aniso(P, p, 1, gs);
                if (P.multires) { float dv = L.r_lo / L.s_lo + L.r_hi / L.s_hi; c.gs = F3(c.gs.x / dv, c.gs.y / dv, c.gs.z / dv); }
                else { float k = sqr(cov_ns_scale(P, p)); c.gs = F3(c.gs.x * k, c.gs.y * k, c.gs.z * k); }
            } else {
                float a[9] = {-2.f * (P.l_aniso_inv[0] * P.l_aniso_inv[0] * 0.25f), 0, 0, 0, -2.f * (P.l_aniso_inv[1] * P.l_aniso_inv[1] * 0.25f), 0, 0, 0,
                              -2.f * (P.l_aniso_inv[2] * P.l_aniso_inv[2] * 0.25f)};
                float r[9]; mat3_inverse(a, r);
                c.gs = F3(r[0] * delta.x, r[4] * delta.y, r[8] * delta.z);
                if (P.has_aniso) c.gs = one_over_second_derivative_aniso(P, p, 0, delta);
                else if (P.use_mtx) { // GPF:728-738 with _cov_mtx_inv
                    float am[9], rm[9];
                    for (int i = 0; i < 9; ++i) am[i] = -2.f * (P.mC[i] * 0.25f);
                    mat3_inverse(am, rm);
                    c.gs = F3(rm[0] * delta.x + (rm[1] * delta.y + rm[2] * delta.z), rm[3] * delta.x + (rm[4] * delta.y + rm[5] * delta.z), rm[6] * delta.x + (rm[7] * delta.y + rm[8] * delta.z));
                }
                if (P.multires) { float dv = L.r_lo / (L.s_lo * L.s_lo) + L.r_hi / (L.s_hi * L.s_hi); c.gs = F3(c.gs.x / dv, c.gs.y / dv, c.gs.z / dv); }
                else { float k = sqr(cov_lateral_scale(P, p)); c.gs = F3(c.gs.x * k, c.gs.y * k, c.gs.z * k); }
            }
        }
        return;
    }
    float cur = value_noise1d(P, c, fr, p, seed, false);
    c.vs = (target_val - meanf) / amp - cur;
    if (P.multires) c.vs /= sqr(L.r_lo) + sqr(L.r_hi);
    if (P.ctx == SCN_CTX_RENEWAL_PLUS) {
        f3 mg = gp_mean_grad(P, p, id);
        // evaluateGradientNoise1D(p, info.t, 0, ...) with conditioning=false: c.vs is set but unused there
        f3 cg = gradient_noise1d(P, c, fr, p, 0.f, dir, in, false);
        f3 delta = F3((target_grad.x - mg.x) / amp - cg.x, (target_grad.y - mg.y) / amp - cg.y, (target_grad.z - mg.z) / amp - cg.z);
        f3 gs = to_local(fr, cov_grad_w2l(P, delta, 1.0f));
        float s2 = sqr(cov_splat_cov1d(P, p, p));
        f3 out = F3(-2.f * gs.x * s2, -2.f * gs.y * s2, -2.f * gs.z * s2);
        if (P.multires) { float dv = sqr(L.r_lo) + sqr(L.r_hi); out = F3(out.x / dv, out.y / dv, out.z / dv); }
        if (!P.corr_xy) { out.x = 0.f; out.y = 0.f; }
        c.gs = out;
    }
}

// neeShared / neeGrad / neePDF (SCN:601-743)
struct Nee { f3 gcon, g3, plane_n; float jac, dist; };
SCN_HD f3 mat3_mulv(const float a[9], f3 b) {
    return F3(a[0] * b.x + a[1] * b.y + a[2] * b.z, a[3] * b.x + a[4] * b.y + a[5] * b.z, a[6] * b.x + a[7] * b.y + a[8] * b.z);
}
SCN_FN void nee_shared(const Params &P, const Coeff &c, f3 dir, f3 normal, f3 p, const Info &in, Nee &o) {
    uint32_t seed = compute_seed(P, in);
    float amp = cov_amplitude(P, p);
    double mean; int id; gp_mean_and_id(P, p, mean, id);
    f3 mg = gp_mean_grad(P, p, id);
    Frame fr = ray_frame(P, dir);
    float sc = (P.stationary || P.multires) ? 1.0f : P.ls_max;
    float wd[3] = {amp * (P.stationary ? P.l_aniso_inv[0] : P.l_aniso_inv[0] / sc), amp * (P.stationary ? P.l_aniso_inv[1] : P.l_aniso_inv[1] / sc),
                   amp * (P.stationary ? P.l_aniso_inv[2] : P.l_aniso_inv[2] / sc)};
    float m[9] = {wd[0] * fr.t.x, wd[0] * fr.b.x, wd[0] * fr.n.x, wd[1] * fr.t.y, wd[1] * fr.b.y, wd[1] * fr.n.y, wd[2] * fr.t.z, wd[2] * fr.b.z, wd[2] * fr.n.z};
    if (P.use_mtx) { // amplitude * _world_to_local_transpose (/ scale) * mtx_ray_coord (SCN:618, GPF:649, 1281-1284)
        const float rc[9] = {fr.t.x, fr.b.x, fr.n.x, fr.t.y, fr.b.y, fr.n.y, fr.t.z, fr.b.z, fr.n.z};
        float wt[9];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) wt[i * 3 + j] = amp * (P.stationary ? P.mW[j * 3 + i] : P.mW[j * 3 + i] / sc);
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m[3 * i + j] = wt[i * 3] * rc[j] + (wt[i * 3 + 1] * rc[3 + j] + wt[i * 3 + 2] * rc[6 + j]);
    }
    float inv[9]; mat3_inverse(m, inv);
    if (!P.multires) { f4 n = noise1d_norm(P, c, fr, p, seed, 1.0f, 0, true, false); o.gcon = F3(n.y, n.z, n.w); }
    else {
        Levels L = level_ratio(P, p);
        f4 lo = noise1d_norm(P, c, fr, p, seed, L.s_lo, P.lvl_add_seed[L.k_lo], true, true), hi = noise1d_norm(P, c, fr, p, seed, L.s_hi, P.lvl_add_seed[L.k_lo + 1], true, false);
        o.gcon = F3(lo.y * L.r_lo / L.s_lo + hi.y * L.r_hi / L.s_hi, lo.z * L.r_lo / L.s_lo + hi.z * L.r_hi / L.s_hi, lo.w * L.r_lo / L.s_lo + hi.w * L.r_hi / L.s_hi);
    }
    f3 pp = mat3_mulv(m, F3(0.f, 0.f, o.gcon.z)); pp = F3(pp.x + mg.x, pp.y + mg.y, pp.z + mg.z);
    f3 ax = F3(m[0], m[3], m[6]), ay = F3(m[1], m[4], m[7]);
    f3 cr = F3(ax.y * ay.z - ax.z * ay.y, ax.z * ay.x - ax.x * ay.z, ax.x * ay.y - ax.y * ay.x);
    float len = sqrtf(cr.x * cr.x + cr.y * cr.y + cr.z * cr.z), il = 1.0f / len;
    o.plane_n = F3(cr.x * il, cr.y * il, cr.z * il); o.jac = len;
    o.dist = (pp.x * o.plane_n.x + pp.y * o.plane_n.y + pp.z * o.plane_n.z) / (normal.x * o.plane_n.x + normal.y * o.plane_n.y + normal.z * o.plane_n.z);
    o.g3 = mat3_mulv(inv, F3(o.dist * normal.x - mg.x, o.dist * normal.y - mg.y, o.dist * normal.z - mg.z));
}
SCN_FN float nee_pdf(const Params &P, const Coeff &c, f3 dir, f3 normal, f3 p, float ts, const Info &in) {
    Nee o; nee_shared(P, c, dir, normal, p, in, o);
    if (o.dist < 0.0f) return 0.0f;
    float sx = cov_splat_cov1d(P, p, p), sy = sx;
    bool plus_xy = P.ctx == SCN_CTX_RENEWAL_PLUS && P.corr_xy;
    double sample_x, sample_y, pdf_area;
    const double sqrt2 = (double)1.41421354f, TWO_PI = 2.0 * 3.14159265358979323846;
    Pcg g; g.seed32(xxhash32_4(in.px, in.py, in.spp, in.seg) + xxhash32_1(in.scene_seed));
    float factor = (float)(exp((double)(-ts * ts) / 4.0) * (0.5 - (double)(ts * ts) * 0.25));
    float cov = (float)(0.5 - (double)(factor * factor * 2.f));
    if (!P.multires) {
        if (!plus_xy) {
            sample_x = (float)((double)(o.g3.x - o.gcon.x) * 1.4142135623730951 * (double)sx); sample_y = (float)((double)(o.g3.y - o.gcon.y) * 1.4142135623730951 * (double)sy);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / TWO_PI * 2.0 * (double)sx * (double)sy);
        } else {
            double a0, a1; rand_normal_2(g, a0, a1); a0 /= sqrt2; a1 /= sqrt2;
            sample_x = (float)(((double)(o.g3.x - o.gcon.x) - (double)factor * 2.0 * a0) / (double)sqrtf(cov));
            sample_y = (float)(((double)(o.g3.y - o.gcon.y) - (double)factor * 2.0 * a1) / (double)sqrtf(cov));
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / TWO_PI / (double)cov);
        }
    } else {
        Levels L = level_ratio(P, p);
        float nss = 1.0f / sqrtf(sqr(L.r_lo / L.s_lo) + sqr(L.r_hi / L.s_hi));
        sx *= nss; sy *= nss;
        if (!plus_xy) {
            sample_x = (float)((double)(o.g3.x - o.gcon.x) * 1.4142135623730951 * (double)sx); sample_y = (float)((double)(o.g3.y - o.gcon.y) * 1.4142135623730951 * (double)sy);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / TWO_PI * 2.0 * (double)sx * (double)sy);
        } else {
            double a0, a1, b0, b1; rand_normal_2(g, a0, a1); a0 /= sqrt2; a1 /= sqrt2; rand_normal_2(g, b0, b1); b0 /= sqrt2; b1 /= sqrt2;
            double mu0 = (double)factor * 2.0 * (a0 * (double)L.r_lo / (double)L.s_lo + b0 * (double)L.r_hi / (double)L.s_hi);
            double mu1 = (double)factor * 2.0 * (a1 * (double)L.r_lo / (double)L.s_lo + b1 * (double)L.r_hi / (double)L.s_hi);
            sample_x = (float)(((double)(o.g3.x - o.gcon.x) - mu0) / (double)sqrtf(cov) * (double)nss);
            sample_y = (float)(((double)(o.g3.y - o.gcon.y) - mu1) / (double)sqrtf(cov) * (double)nss);
            pdf_area = (float)(exp(-sample_x * sample_x / 2.0) * exp(-sample_y * sample_y / 2.0) / (TWO_PI * (double)cov) * (double)sqr(nss));
        }
    }
    float cos_l = fabsf(normal.x * o.plane_n.x + normal.y * o.plane_n.y + normal.z * o.plane_n.z);
    float pdf_normal = (float)pdf_area * (o.dist * o.dist) / cos_l / o.jac;
    float ndw = -(normal.x * dir.x + normal.y * dir.y + normal.z * dir.z);
    return (float)((double)pdf_normal / (4.0 * (double)ndw));
}

// ---------------------------------------------------------------- medium query (SCNM:93-183, GPM:221-393)
struct RayState {
    bool first_scatter; int bounce, last_gp_id, scheme; f3 last_aniso; float last_val; Info info; Coeff c;
};
struct Sample { bool ok, exited; float t, cont_t; f3 aniso; int gp_id, scheme; float weight; uint32_t n_evals; };

SCN_HD f3 ray_at(dbl3 o, dbl3 d, double t) { return F3((float)yadd(o.x, ymul(t, d.x)), (float)yadd(o.y, ymul(t, d.y)), (float)yadd(o.z, ymul(t, d.z))); }

// intersectGP (SCNM:102-183): fixed-step march, secant refinement with the reference's quirks.
// Returns true on a zero crossing; t is the crossing (or far_t).
SCN_FN bool intersect_gp(const Params &P, f3 pos, f3 dir, float near_t, float far_t, float u, RayState &st, const Frame &fr, uint32_t seed, double &t_out, uint32_t &evals) {
    dbl3 o = to_dbl3(pos), rd = to_dbl3(dir);
    float step = xdiv(xsub(far_t, near_t), (float)P.min_step);
    if (P.step_size < step) step = P.step_size;
    int gp_id = 0;
    st.c.vs = 0.f; st.c.gs = F3(0.f, 0.f, 0.f); st.c.org = F3(0.f, 0.f, 0.f); // fresh realization per query (SCNM:104)
    if (!st.first_scatter && P.activate_conditioning) {
        conditioning(P, st.c, fr, pos, dir, st.last_val, st.last_aniso, st.info);
        evals += (P.sampling_1d && P.ctx == SCN_CTX_RENEWAL_PLUS) ? 2u : 1u;
    }
    double t = (double)near_t;
    double f0 = (double)eval_value(P, st.c, fr, ray_at(o, rd, t), seed, gp_id); evals++;
    int sign0 = f0 < 0 ? -1 : 1;
    double pf = f0;
    t = (double)xadd(near_t, xmul(step, u));
    int nstep = 0;
    const double dstep = (double)step, dfar = (double)far_t;
    while (t < dfar) {
        nstep++;
        double fc = (double)eval_value(P, st.c, fr, ray_at(o, rd, t), seed, gp_id); evals++;
        int signc = fc < 0 ? -1 : 1;
        if (!st.first_scatter && nstep == 1) sign0 = signc;
        else if (signc != sign0) {
            double intp = ydiv(pf, ysub(pf, fc));
            double ta = ysub(t, dstep);
            double t_test, t_prev = yadd(ymul(ta, ysub(1.0, intp)), ymul(t, intp));
            for (;;) {
                t_test = yadd(ymul(ta, ysub(1.0, intp)), ymul(t, intp));
                double ft = (double)eval_value(P, st.c, fr, ray_at(o, rd, t_test), seed, gp_id); evals++;
                int sign_t = ft < 0 ? -1 : 1;
                if (sign_t == sign0) break;
                intp = ymul(intp, 0.9);
                if (intp <= 0.01) { t_prev = t_test = 0; break; }
                t_prev = t_test;
            }
            t = t_prev;
            st.last_gp_id = gp_id;
            st.scheme = cov_amplitude(P, ray_at(o, rd, t)) < 1e-6f ? SCN_SCHEME_UNI : P.scheme;
            st.last_val = 0.0f;
            t_out = t;
            return true;
        }
        pf = fc;
        t = yadd(t, dstep);
    }
    t = dfar;
    st.scheme = SCN_SCHEME_UNI;
    st.last_val = eval_value(P, st.c, fr, ray_at(o, rd, t), seed, gp_id); evals++;
    t_out = t;
    return false;
}

// GaussianProcessMedium::sampleDistance / transmittance share this walk (GPM:221-393).  The sparse-convolution
// intersectGP always consumes the whole ray, so the reference's do/while runs once.
SCN_FN void medium_query(const Params &P, f3 pos, f3 dir, float near_t, float far_t, float u, RayState &st, Sample &sm, bool transmittance_only) {
    sm.ok = false; sm.exited = false; sm.t = 0.f; sm.cont_t = 0.f; sm.aniso = F3(0.f, 0.f, 0.f); sm.gp_id = 0; sm.scheme = SCN_SCHEME_UNI; sm.weight = 0.f; sm.n_evals = 0;
    double start_t = (double)near_t;
    if (!isfinite(far_t)) far_t = (float)(start_t + 2000.0);
    const float max_t = far_t;
    // GPM:250-258: an absorption-only medium answers sampleDistance with weight = transmittance(); that walk touches
    // the state (lastAniso, firstScatter) but advance() is not called.
    const bool absorb = !transmittance_only && P.absorption_only;
    if (!transmittance_only) {
        if (st.bounce >= P.max_bounces) return;
        if (max_t == 0.f) { sm.ok = true; sm.t = max_t; sm.weight = 1.f; sm.exited = true; return; }
    }
    Frame fr = ray_frame(P, dir);
    uint32_t seed = compute_seed(P, st.info);
    double t = (double)max_t;
    dbl3 ro = to_dbl3(pos), rdn = to_dbl3(dir);
    { double inv = ydiv(1.0, ysqrt(yadd(yadd(ymul(rdn.x, rdn.x), ymul(rdn.y, rdn.y)), ymul(rdn.z, rdn.z)))); rdn.x = ymul(rdn.x, inv); rdn.y = ymul(rdn.y, inv); rdn.z = ymul(rdn.z, inv); }
    uint32_t evals = 0;
    bool hit = intersect_gp(P, pos, dir, (float)start_t, far_t, u, st, fr, seed, t, evals);
    bool exited = !hit;
    f3 grad = F3(0.f, 0.f, 0.f);
    if (t < (double)max_t) {
        f3 ip = F3((float)yadd(ro.x, ymul(rdn.x, t)), (float)yadd(ro.y, ymul(rdn.y, t)), (float)yadd(ro.z, ymul(rdn.z, t)));
        grad = eval_gradient(P, st.c, fr, ip, (float)t, dir, st.info, seed); evals++;
        st.last_aniso = grad; st.first_scatter = false;
        sm.aniso = grad;
        double avg = ((double)grad.x + (double)grad.y + (double)grad.z) * (1.0 / 3.0);
        if (!isfinite(avg)) {
            sm.aniso = F3(1.f, 0.f, 0.f); sm.n_evals = evals; sm.weight = 0.f;
            if (absorb) { sm.ok = true; sm.t = max_t; sm.exited = true; sm.gp_id = st.last_gp_id; st.info.t += sm.t; sm.aniso = F3(0.f, 0.f, 0.f); }
            return;
        }
    }
    sm.n_evals = evals;
    if (transmittance_only) { sm.ok = true; sm.exited = exited; sm.weight = exited ? 1.f : 0.f; return; }
    if (absorb) { sm.ok = true; sm.t = max_t; sm.weight = exited ? 1.f : 0.f; sm.exited = true; sm.gp_id = st.last_gp_id; st.info.t += sm.t; sm.aniso = F3(0.f, 0.f, 0.f); return; }
    sm.exited = exited;
    if (!exited) {
        double dt = (double)grad.x * (double)dir.x + (double)grad.y * (double)dir.y + (double)grad.z * (double)dir.z;
        if (dt > 0) return;
        double l2 = (double)grad.x * (double)grad.x + (double)grad.y * (double)grad.y + (double)grad.z * (double)grad.z;
        if (l2 < (double)0.0000001f) { sm.aniso = F3(1.f, 0.f, 0.f); return; }
        sm.weight = 1.f;
    } else {
        f3 ip = F3((float)yadd(ro.x, ymul(rdn.x, t)), (float)yadd(ro.y, ymul(rdn.y, t)), (float)yadd(ro.z, ymul(rdn.z, t)));
        sm.aniso = eval_gradient(P, st.c, fr, ip, (float)t, dir, st.info, seed); sm.n_evals++;
        sm.weight = 1.f;
    }
    sm.t = (float)t < max_t ? (float)t : max_t;
    sm.cont_t = (float)t;
    sm.weight *= P.sigma_s_over_t;
    st.last_aniso = sm.aniso;
    sm.scheme = st.scheme;
    st.first_scatter = false; st.scheme = SCN_SCHEME_UNI; st.bounce++; // state.advance()
    sm.gp_id = st.last_gp_id;
    st.info.t += sm.t;
    sm.ok = true;
}

} // namespace scn
#endif
