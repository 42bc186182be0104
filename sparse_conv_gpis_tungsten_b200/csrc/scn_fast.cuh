// scn_fast.cuh — the roofline kernel: 3D sparse-convolution noise value + gradient for the stationary
// squared-exponential kernel (evaluateNoise3D, math/SparseConvolutionNoise.cpp:101-124,251-320,362-395),
// isotropic-ray space and world space.  One thread per query point; 27 cells x N impulses regenerated on the fly
// from the reference's cell hash + PCG32 stream (bit-exact), rejection test on exactly rounded operands, then
// FMA/SFU kernel evaluation on the accepted ~15.5 % (4*pi/81).
//
// Differences from the generic path in scn_device.cuh (all on the tolerance side of the rejection test):
//   * exp(-d^T A d) is one MUFU.EX2 on  -(log2 e) * R^2 * sum_j A_j t_j^2  with t = p - p_i in cell units;
//   * the gradient is accumulated as sum w f t_j and scaled by -2 A_j R once per evaluation;
//   * cells are visited dz-outermost so that two of the three hash rounds are hoisted out of the inner loops
//     (the reference visits dx-outermost; only the floating-point summation order changes).
#ifndef SCN_FAST_CUH_
#define SCN_FAST_CUH_

#include "scn_device.cuh"

namespace scn {

inline bool fast_noise3d_supported(const Params &P) { return P.stationary && !P.sampling_1d && P.n_imp > 0; }

constexpr int FAST_BLOCK = 128;

template <bool ISO>
__global__ void __launch_bounds__(FAST_BLOCK) k_fast_noise3d(const __grid_constant__ Params P, const float *__restrict__ px, const float *__restrict__ py,
                                                             const float *__restrict__ pz, const float *__restrict__ dx, const float *__restrict__ dy,
                                                             const float *__restrict__ dz, uint32_t seed, float *__restrict__ oval, float *__restrict__ ogx,
                                                             float *__restrict__ ogy, float *__restrict__ ogz, size_t n) {
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
    // noise3D (SCN:362-373): grid coordinates, exactly rounded
    f3 pg = xdiv3(q, R);
    f3 fl = F3(floorf(pg.x), floorf(pg.y), floorf(pg.z));
    f3 frac = xsub3(pg, fl);
    const int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z;
    const uint32_t nimp = P.n_imp;
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
#pragma unroll 1
            for (int ddx = -1; ddx <= 1; ++ddx) {
                const float cx = xsub(frac.x, (float)ddx);
                const uint32_t h = xx_final(XP4 * rotl17(hy + (uint32_t)(ix + ddx) * XP3));
                Pcg g; g.seed32(h);
                float cv = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll 1
                for (uint32_t k = 0; k < nimp; ++k) {
                    // next3D draws z, y, x (g++ argument order), then the Bernoulli sign (SCN:384-385)
                    const uint64_t st0 = g.s, st1 = st0 * PCG_MULT + 1ULL, st2 = st1 * PCG_MULT + 1ULL, st3 = st2 * PCG_MULT + 1ULL;
                    g.s = st3 * PCG_MULT + 1ULL;
                    const float tz = xsub(cz, u32_to_unit(pcg_output(st0)));
                    const float ty = xsub(cy, u32_to_unit(pcg_output(st1)));
                    const float tx = xsub(cx, u32_to_unit(pcg_output(st2)));
                    const float l2 = xadd(xadd(xmul(tx, tx), xmul(ty, ty)), xmul(tz, tz));
                    if (l2 < 1.0f) {
                        // Bernoulli(u < 0.5 ? -1 : +1): u < 0.5 <=> top bit of the 32-bit draw clear
                        const uint32_t r = pcg_output(st3);
                        float f = exp2f(-(A0 * tx * tx + (A1 * ty * ty + A2 * tz * tz)));
                        f = (r & 0x80000000u) ? f : -f;
                        cv += f; c0 = fmaf(f, tx, c0); c1 = fmaf(f, ty, c1); c2 = fmaf(f, tz, c2);
                    }
                }
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
    if (P.iso3d) k_fast_noise3d<true><<<grid, FAST_BLOCK, 0, s>>>(P, px, py, pz, dx, dy, dz, seed, oval, ogx, ogy, ogz, n);
    else k_fast_noise3d<false><<<grid, FAST_BLOCK, 0, s>>>(P, px, py, pz, dx, dy, dz, seed, oval, ogx, ogy, ogz, n);
}

} // namespace scn
#endif
