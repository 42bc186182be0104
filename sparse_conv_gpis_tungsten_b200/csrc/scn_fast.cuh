// scn_fast.cuh — the roofline kernel: 3D sparse-convolution noise value + gradient for the stationary
// squared-exponential kernel (evaluateNoise3D, math/SparseConvolutionNoise.cpp:101-124,251-320,362-395),
// isotropic-ray space and world space.  One thread per query point; 27 cells x N impulses regenerated on the fly
// from the reference's cell hash + PCG32 stream (bit-exact), rejection test on exactly rounded operands, then
// FMA/SFU kernel evaluation on the accepted ~15.5 % (4*pi/81).
//
// Differences from the generic path in scn_device.cuh (all on the tolerance side of the rejection test):
//   * exp(-d^T A d) is one MUFU.EX2 on  -(log2 e) * R^2 * sum_j A_j t_j^2  with t = p - p_i in cell units
//     (one multiply of the already computed |t|^2 when A is isotropic);
//   * the gradient is accumulated as sum w f t_j and scaled by -2 A_j R once per evaluation;
//   * cells are visited dz-outermost so that two of the three hash rounds are hoisted out of the inner loops
//     (the reference visits dx-outermost; only the floating-point summation order changes);
//   * the impulse loop is branch-free (the accepted branch is taken by some lane of a warp 99.5 % of the time, so
//     a branch only adds BSSY/BSYNC and divergence) and the 64-bit LCG step is three IMADs written in PTX.
// ncu (profiles/): the kernel is issue-bound with the ALU pipe (shifts, logic ops of the PCG output function) the
// fullest pipe; see DESIGN.md "k_fast_noise3d" for the per-impulse instruction budget.
#ifndef SCN_FAST_CUH_
#define SCN_FAST_CUH_

#include "scn_device.cuh"

namespace scn {

inline bool fast_noise3d_supported(const Params &P) { return P.stationary && !P.use_mtx && P.kernel == SCN_COV_SQUARED_EXPONENTIAL && !P.sampling_1d && P.n_imp > 0; }

constexpr int FAST_BLOCK = 128;

#if defined(__CUDACC__)
#ifndef SCN_LCG_JUMP
#define SCN_LCG_JUMP 1
#endif
// One PCG32 state transition s' = s * 6364136223846793005 + 1 on a (lo, hi) register pair: IMAD.WIDE + 2 IMAD.
// `inc` is the 64-bit increment (1) held in a register pair; it comes in as a kernel argument so that ptxas keeps
// it resident instead of re-materialising two MOVs in front of every IMAD.WIDE.
__device__ __forceinline__ void lcg_step(uint32_t &lo, uint32_t &hi, uint64_t inc) {
    uint32_t nlo, nhi;
    asm("{\n\t.reg .u64 t, c;\n\t.reg .u32 th, x, one, dummy;\n\t"
        "mov.b64 {one, dummy}, %4;\n\t"
        "mul.lo.u32 x, %3, 0x4c957f2d;\n\t"
        "mov.b64 c, {one, x};\n\t"
        "mad.wide.u32 t, %2, 0x4c957f2d, c;\n\t"
        "mov.b64 {%0, th}, t;\n\t"
        "mad.lo.u32 %1, %2, 0x5851f42d, th;\n\t}"
        : "=r"(nlo), "=r"(nhi) : "r"(lo), "r"(hi), "l"(inc));
    lo = nlo; hi = nhi;
}
// PCG XSH-RR output of the state (lo, hi): xs = bits [27,59) of s ^ (s >> 18), rotated right by s >> 59
__device__ __forceinline__ uint32_t pcg_out(uint32_t lo, uint32_t hi) {
    // SCN_PCG_VARIANT picks which of the two plain shifts of `hi` go through IMAD.HI (FMA-heavy pipe) instead of SHF (ALU pipe);
    // both pipes are close to full in the impulse loop (profiles/), the variants were timed on B200 and the fastest kept
#ifndef SCN_PCG_VARIANT
#define SCN_PCG_VARIANT 0
#endif
#if SCN_PCG_VARIANT == 0
    uint32_t xs = __funnelshift_r(lo, hi, 27) ^ (hi >> 13);
    return __funnelshift_r(xs, xs, hi >> 27);
#elif SCN_PCG_VARIANT == 1
    uint32_t xs = __funnelshift_r(lo, hi, 27) ^ (hi >> 13);
    return __funnelshift_r(xs, xs, __umulhi(hi, 1u << 5));
#else
    uint32_t xs = __funnelshift_r(lo, hi, 27) ^ __umulhi(hi, 1u << 19);
    return __funnelshift_r(xs, xs, __umulhi(hi, 1u << 5));
#endif
}
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// PCG32 state after j transitions from the 32-bit seed x (hi word 0): s_j = A_j x + C_j (mod 2^64) with A_j = M^j and
// C_j = (M^j - 1) / (M - 1), compile-time constants.  Because x has no high word this is IMAD.WIDE + one IMAD per draw instead of
// IMAD.WIDE + 2 IMAD for a chained step, and the draws of a cell no longer depend on each other.
template <int J> struct LcgJump { static constexpr uint64_t A = LcgJump<J - 1>::A * PCG_MULT, C = LcgJump<J - 1>::C * PCG_MULT + 1ULL; };
template <> struct LcgJump<0> { static constexpr uint64_t A = 1ULL, C = 0ULL; };
template <int J>
__device__ __forceinline__ uint32_t pcg_draw_at(uint32_t x) {
    const uint64_t w = (uint64_t)x * (uint32_t)LcgJump<J>::A + LcgJump<J>::C;
    return pcg_out((uint32_t)w, (uint32_t)(w >> 32) + x * (uint32_t)(LcgJump<J>::A >> 32));
}

// impulse K of a cell seeded with x = h + 1 (draws 2 + 4K .. 5 + 4K of the stream: set_state discards two), accumulated
template <bool ISO_A, int K, int NIMP>
struct ImpulseUnrolled {
    static __device__ __forceinline__ void run(uint32_t x, float cx, float cy, float cz, float a0, float a1, float a2, float &sv, float &s0, float &s1, float &s2) {
        const float uz = u32_to_unit(pcg_draw_at<2 + 4 * K>(x)), uy = u32_to_unit(pcg_draw_at<3 + 4 * K>(x)), ux = u32_to_unit(pcg_draw_at<4 + 4 * K>(x));
        const uint32_t rs = pcg_draw_at<5 + 4 * K>(x);
        const float tz = xsub(cz, uz), ty = xsub(cy, uy), tx = xsub(cx, ux);
        const float l2 = xadd(xadd(xmul(tx, tx), xmul(ty, ty)), xmul(tz, tz));
        const float arg = ISO_A ? -a0 * l2 : -(a0 * tx * tx + (a1 * ty * ty + a2 * tz * tz));
        float f = ex2_approx(arg);
        f = __uint_as_float(__float_as_uint(f) ^ (~rs & 0x80000000u));
        f = l2 < 1.0f ? f : 0.0f;
        sv += f; s0 = fmaf(f, tx, s0); s1 = fmaf(f, ty, s1); s2 = fmaf(f, tz, s2);
        ImpulseUnrolled<ISO_A, K + 1, NIMP>::run(x, cx, cy, cz, a0, a1, a2, sv, s0, s1, s2);
    }
};
template <bool ISO_A, int NIMP>
struct ImpulseUnrolled<ISO_A, NIMP, NIMP> {
    static __device__ __forceinline__ void run(uint32_t, float, float, float, float, float, float, float &, float &, float &, float &) {}
};

// All N impulses of one cell (cell3D, SCN:375-395) against the point whose cell-local coordinates are (cx, cy, cz).
// ISO_A: the kernel's inverse covariance is a multiple of the identity in the sampling space (a0 only).
// NIMP > 0: the impulse count is a compile-time constant and the loop is unrolled (no loop-control instructions).
template <bool ISO_A, int NIMP>
__device__ __forceinline__ void cell_accumulate(uint32_t h, uint32_t nimp_dyn, uint64_t inc, float cx, float cy, float cz, float a0, float a1, float a2,
                                                float &sv, float &s0, float &s1, float &s2) {
#if SCN_LCG_JUMP
    if (NIMP > 0) { ImpulseUnrolled<ISO_A, 0, (NIMP > 0 ? NIMP : 0)>::run(h + 1u, cx, cy, cz, a0, a1, a2, sv, s0, s1, s2); return; }
#endif
    // set_state(h + 1u) (the +1 wraps in 32 bits), then two discarded draws (sampling/UniformSampler.hpp:71-75)
    uint32_t lo = h + 1u, hi = 0u;
    lcg_step(lo, hi, inc); lcg_step(lo, hi, inc);
    const uint32_t nimp = NIMP > 0 ? (uint32_t)NIMP : nimp_dyn;
#pragma unroll(NIMP > 0 ? NIMP : 1)
    for (uint32_t k = 0; k < nimp; ++k) {
        // next3D draws z, y, x (g++ argument order), then the Bernoulli sign (SCN:384-385)
        const float uz = u32_to_unit(pcg_out(lo, hi)); lcg_step(lo, hi, inc);
        const float uy = u32_to_unit(pcg_out(lo, hi)); lcg_step(lo, hi, inc);
        const float ux = u32_to_unit(pcg_out(lo, hi)); lcg_step(lo, hi, inc);
        const uint32_t rs = pcg_out(lo, hi); lcg_step(lo, hi, inc);
        const float tz = xsub(cz, uz), ty = xsub(cy, uy), tx = xsub(cx, ux);
        const float l2 = xadd(xadd(xmul(tx, tx), xmul(ty, ty)), xmul(tz, tz));
        const float arg = ISO_A ? -a0 * l2 : -(a0 * tx * tx + (a1 * ty * ty + a2 * tz * tz));
        float f = ex2_approx(arg);
        // Bernoulli(u < 0.5 ? -1 : +1): u < 0.5 <=> top bit of the 32-bit draw clear -> flip the sign bit of f
        f = __uint_as_float(__float_as_uint(f) ^ (~rs & 0x80000000u));
        f = l2 < 1.0f ? f : 0.0f;
        sv += f; s0 = fmaf(f, tx, s0); s1 = fmaf(f, ty, s1); s2 = fmaf(f, tz, s2);
    }
}

template <bool ISO, int NIMP>
__global__ void __launch_bounds__(FAST_BLOCK) k_fast_noise3d(const __grid_constant__ Params P, const float *__restrict__ px, const float *__restrict__ py,
                                                             const float *__restrict__ pz, const float *__restrict__ dx, const float *__restrict__ dy,
                                                             const float *__restrict__ dz, uint32_t seed, uint64_t inc, float *__restrict__ oval,
                                                             float *__restrict__ ogx, float *__restrict__ ogy, float *__restrict__ ogz, size_t n) {
    size_t i = (size_t)blockIdx.x * FAST_BLOCK + threadIdx.x;
    if (i >= n) return;
    f3 p = F3(__ldg(px + i), __ldg(py + i), __ldg(pz + i));
    Frame fr;
    f3 q;        // point in the sampling space (iso-ray or world)
    float R;     // cell size in that space
    float A0, A1, A2; // log2(e) * R^2 * diag(inverse covariance)
    if (ISO) {
        fr = ray_frame(P, F3(__ldg(dx + i), __ldg(dy + i), __ldg(dz + i)));
        q = to_local(fr, cov_pos_w2l(P, p, 1.0f));
        R = P.R_iso;
        A0 = A1 = A2 = 1.44269504f * 0.5f * R * R;
    } else {
        q = p; R = P.R_world;
        float k = 1.44269504f * 0.5f * R * R;
        A0 = k * P.l_aniso_inv[0] * P.l_aniso_inv[0]; A1 = k * P.l_aniso_inv[1] * P.l_aniso_inv[1]; A2 = k * P.l_aniso_inv[2] * P.l_aniso_inv[2];
    }
    const bool iso_a = ISO || (A0 == A1 && A1 == A2); // uniform across the grid
    // noise3D (SCN:362-373): grid coordinates, exactly rounded
    f3 pg = xdiv3(q, R);
    f3 fl = F3(floorf(pg.x), floorf(pg.y), floorf(pg.z));
    f3 frac = xsub3(pg, fl);
    const int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z;
    const uint32_t nimp = P.n_imp;
    // keep the increment in a per-thread register pair: as a uniform kernel argument ptxas feeds it through the uniform
    // datapath and splits every IMAD.WIDE into IMAD.WIDE + IADD3 + IADD3.X (two more ALU-pipe instructions per draw)
    inc += (uint64_t)(__float_as_uint(p.x) == 0x7fc54321u); // never true for a finite coordinate; opaque to ptxas
    const float cxm = xsub(frac.x, -1.0f), cx0 = frac.x, cxp = xsub(frac.x, 1.0f);
    float sv = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll 1
    for (int ddz = -1; ddz <= 1; ++ddz) {
        const float cz = xsub(frac.z, (float)ddz);
        // xxhash32(Vec4u(k, j, i, seed)) (math/MathUtil.hpp:211-224): k is mixed first
        const uint32_t hz = XP4 * rotl17(seed + XP5 + (uint32_t)(iz + ddz) * XP3);
#pragma unroll 1
        for (int ddy = -1; ddy <= 1; ++ddy) {
            const float cy = xsub(frac.y, (float)ddy);
            const uint32_t hy = XP4 * rotl17(hz + (uint32_t)(iy + ddy) * XP3);
#pragma unroll(NIMP > 4 ? 1 : 3)
            for (int ddx = -1; ddx <= 1; ++ddx) {
                const float cx = ddx < 0 ? cxm : (ddx == 0 ? cx0 : cxp);
                const uint32_t h = xx_final(XP4 * rotl17(hy + (uint32_t)(ix + ddx) * XP3));
                float cv = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f;
                if (iso_a) cell_accumulate<true, NIMP>(h, nimp, inc, cx, cy, cz, A0, A1, A2, cv, c0, c1, c2);
                else cell_accumulate<false, NIMP>(h, nimp, inc, cx, cy, cz, A0, A1, A2, cv, c0, c1, c2);
                sv += cv; s0 += c0; s1 += c1; s2 += c2;
            }
        }
    }
    // d/dq_j exp(-sum A'_j (R t_j)^2) = -2 A'_j R t_j f with A' = A / (log2e R^2)
    const float gk = -2.0f / (1.44269504f * R);
    f3 gq = F3(gk * A0 * s0, gk * A1 * s1, gk * A2 * s2);
    float inv_nf;
    f3 gw;
    if (ISO) {
        f3 gi = to_global(fr, gq);
        gw = F3(gi.x * P.l_aniso_inv[0], gi.y * P.l_aniso_inv[1], gi.z * P.l_aniso_inv[2]);
        inv_nf = 1.0f / P.nf_iso;
    } else { gw = gq; inv_nf = 1.0f / P.nf_world; }
    oval[i] = sv * inv_nf; ogx[i] = gw.x * inv_nf; ogy[i] = gw.y * inv_nf; ogz[i] = gw.z * inv_nf;
}

inline void launch_fast_noise3d(const Params &P, const float *px, const float *py, const float *pz, const float *dx, const float *dy, const float *dz,
                                uint32_t seed, float *oval, float *ogx, float *ogy, float *ogz, size_t n, cudaStream_t s) {
    unsigned grid = (unsigned)((n + FAST_BLOCK - 1) / FAST_BLOCK);
    const uint64_t inc = 1ULL; // the PCG32 stream increment (_sequence | 1, sampling/UniformSampler.hpp:17-27)
#define SCN_FAST_LAUNCH(ISO_, N_) k_fast_noise3d<ISO_, N_><<<grid, FAST_BLOCK, 0, s>>>(P, px, py, pz, dx, dy, dz, seed, inc, oval, ogx, ogy, ogz, n)
#define SCN_FAST_BY_N(ISO_) switch (P.n_imp) { case 1: SCN_FAST_LAUNCH(ISO_, 1); break; case 2: SCN_FAST_LAUNCH(ISO_, 2); break; case 3: SCN_FAST_LAUNCH(ISO_, 3); break; \
                                               case 4: SCN_FAST_LAUNCH(ISO_, 4); break; case 8: SCN_FAST_LAUNCH(ISO_, 8); break; \
                                               case 16: SCN_FAST_LAUNCH(ISO_, 16); break; case 32: SCN_FAST_LAUNCH(ISO_, 32); break; default: SCN_FAST_LAUNCH(ISO_, 0); }
    if (P.iso3d) { SCN_FAST_BY_N(true) } else { SCN_FAST_BY_N(false) }
#undef SCN_FAST_BY_N
#undef SCN_FAST_LAUNCH
}
#endif // __CUDACC__

} // namespace scn
#endif
