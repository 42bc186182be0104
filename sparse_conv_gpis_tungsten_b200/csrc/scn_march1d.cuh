// scn_march1d.cuh — wavefront march for 1D-sampling media (BASELINE configs C3 / C4):
//   SparseConvolutionNoiseMedium::intersectGP + GaussianProcessMedium::sampleDistance / transmittance (SCNM:93-183, GPM:221-393)
//   over evaluateValueNoise1D / evaluateNoise1DNormalized / noise1D / cell1D (SCN:126-135, 322-360, 397-428).
//
// Same skeleton as k_march (scn_march.cuh): one ray per lane, a per-lane state machine with ONE value-evaluation site, persistent
// warps that pull rays from a global counter.  What differs is the cache: a 1D evaluation reads the 3 N impulses (position, sign)
// of three consecutive cells of the ray-space z axis, and z grows monotonically along the ray, so every lane keeps those 3 N floats
// in REGISTERS and, when it steps into the next cell, shifts two cells down and regenerates one (xxhash32(Vec2u(i, seed)) + 2 + 2 N
// PCG draws) instead of regenerating all three at every step (~21 steps per cell at the default step size).  The rejection test uses
// the reference's operands in the reference's association order (tc = frac - dx, tp = tc - t_i, tp * tp < 1), so accepted sets are
// identical to the CPU's.
//
// The once-per-ray work — Renewal / Renewal+ conditioning at the ray origin (SCN:519-595) and the gradient at the crossing or exit
// (double-precision Box-Muller draws, SCN:137-239) — costs a few thousand instructions per ray and would run with one or two active
// lanes inside the march loop (rays of a warp finish at different iterations; measured: 63 % of all issued instructions at 1.1 active
// lanes).  It is therefore split off into two converged one-thread-per-ray kernels around the march:
//     k_cond1d  ->  k_march1d (values only)  ->  k_finish1d (gradient + the tail of sampleDistance)
// which hand the per-ray intermediate state (conditioning coefficients, crossing distance as a hi/lo float pair, flags) through a
// stream-ordered scratch buffer (cudaMallocAsync on the caller's stream: no synchronisation, safe for concurrent streams).
//
// Covered: 1D sampling with the stationary squared-exponential kernel and with its brute-force non-stationary form (BRUTE: kernel scale,
// variance normalisation and conditioning scale follow ls(p) at the query point) and its multi-resolution form (two levels, each with its
// own register window), 1 <= N <= 4 (compile-time), every mean (procedural means through the PROC instantiations).
// Everything else stays on the generic kernels of scn_api.cu.
#ifndef SCN_MARCH1D_CUH_
#define SCN_MARCH1D_CUH_

#include "scn_march.cuh"

namespace scn {

inline bool march1d_supported(const Params &P) {
    return P.sampling_1d && !P.has_aniso && P.kernel == SCN_COV_SQUARED_EXPONENTIAL && P.n_imp >= 1 && P.n_imp <= 4;
}

#if defined(__CUDACC__)

constexpr int MARCH1D_BLOCK = 128;
constexpr size_t SCRATCH1D_BYTES_PER_RAY = 13 * sizeof(float); // Scratch1D below: 13 arrays of 4 bytes per ray

#if defined(SCN_MARCH1D_IMPL) // the kernels and their launchers: scn_march1d.cu only (scn_api.cu takes the declarations at the end)

// the 2 + 2 N draws of cell1D (SCN:408-428) for cell index `cell`: positions and a sign bit per impulse (1 = +1)
template <int NIMP>
__device__ __forceinline__ void cell1d_generate(uint32_t cell, uint32_t seed, float (&ti)[NIMP], uint32_t &plus) {
    Pcg g; g.seed32(xxhash32_2(cell, seed));
    plus = 0u;
#pragma unroll
    for (int k = 0; k < NIMP; ++k) {
        ti[k] = g.next();
        const float ws = g.next();
        plus |= (ws < 0.5f ? 0u : 1u) << k; // Bernoulli: rv < 0.5 ? -1 : +1 (math/MathUtil.hpp:226-228)
    }
}

// Register cache of one lattice level: a window of FOUR consecutive cells cb .. cb + 3 (an evaluation in cell i reads i - 1, i, i + 1, so
// the window serves i = cb + 1 and i = cb + 2).  The spare cell is what lets the regeneration run converged: a lane MUST move its window
// when it enters cb + 3, and whenever any lane of the warp must, every lane already in cb + 2 moves too (one cell early).  After such a
// round every lane is at least one full cell away from its next forced move, so rounds happen about once per cell length (~21 steps) with
// most lanes taking part, instead of each lane regenerating on its own iteration with one or two lanes active.
template <int NIMP>
struct Win1D { int cb, ck; bool valid; float w0[NIMP], w1[NIMP], w2[NIMP], w3[NIMP]; uint32_t sig; /* sign bits: 4 per slot */ };

// noise1D (SCN:397-406) for this lane's point (cell i, fraction frac) of the level with table index kidx and seed seedl, through the
// window; `ka` = -log2(e) / (2 ls^2).  Called by all 32 lanes together (window moves are warp-cooperative); returns the unnormalised sum.
template <int NIMP>
__device__ __forceinline__ float win1d_noise(Win1D<NIMP> &W, const bool active, const int i, const float frac, const uint32_t seedl, const int kidx,
                                             const float R, const float ka, const int lane) {
    const unsigned FULL = 0xffffffffu;
    int nshift = 0; bool must = false;
    if (active) {
        if (!W.valid || W.ck != kidx || i < W.cb + 1 || i > W.cb + 3) nshift = 4; // new ray / level change / jump: build cells i - 1 .. i + 2
        else if (i == W.cb + 3) { nshift = 1; must = true; }
        else if (i == W.cb + 2) nshift = 1; // one cell early, taken only when some lane of the warp must move
    }
    // a lane that starts a ray (or jumped): lanes 0..3 generate its four cells side by side and hand them over by shuffles
    unsigned ms = __ballot_sync(FULL, nshift == 4);
    while (ms) {
        const int T = __ffs(ms) - 1; ms &= ms - 1;
        const int ti = __shfl_sync(FULL, i, T); const uint32_t ts = __shfl_sync(FULL, seedl, T);
        float g[NIMP]; uint32_t plus = 0u;
#pragma unroll
        for (int k = 0; k < NIMP; ++k) g[k] = 0.f;
        if (lane < 4) cell1d_generate<NIMP>((uint32_t)(ti - 1 + lane), ts, g, plus);
#pragma unroll
        for (int k = 0; k < NIMP; ++k) {
            const float a0 = __shfl_sync(FULL, g[k], 0), a1 = __shfl_sync(FULL, g[k], 1), a2 = __shfl_sync(FULL, g[k], 2), a3 = __shfl_sync(FULL, g[k], 3);
            if (lane == T) { W.w0[k] = a0; W.w1[k] = a1; W.w2[k] = a2; W.w3[k] = a3; }
        }
        const uint32_t pw = __shfl_sync(FULL, plus, 0) | (__shfl_sync(FULL, plus, 1) << 4) | (__shfl_sync(FULL, plus, 2) << 8) | (__shfl_sync(FULL, plus, 3) << 12);
        if (lane == T) { W.sig = pw; W.cb = ti - 1; W.ck = kidx; W.valid = true; }
    }
    if (__any_sync(FULL, must)) { // a round: every lane in the upper half of its window moves one cell forward
        if (nshift == 1) {
#pragma unroll
            for (int k = 0; k < NIMP; ++k) { W.w0[k] = W.w1[k]; W.w1[k] = W.w2[k]; W.w2[k] = W.w3[k]; }
            W.cb += 1;
            uint32_t plus; cell1d_generate<NIMP>((uint32_t)(W.cb + 3), seedl, W.w3, plus);
            W.sig = (W.sig >> 4) | (plus << 12);
        }
    }
    float sv = 0.f;
    if (active) {
        const bool hi = i == W.cb + 2; // which three of the four cached cells are i - 1, i, i + 1
        float im[NIMP], i0[NIMP], ip[NIMP];
#pragma unroll
        for (int k = 0; k < NIMP; ++k) { im[k] = hi ? W.w1[k] : W.w0[k]; i0[k] = hi ? W.w2[k] : W.w1[k]; ip[k] = hi ? W.w3[k] : W.w2[k]; }
        const uint32_t sg = hi ? W.sig >> 4 : W.sig, sm = sg & 15u, s0 = (sg >> 4) & 15u, sp = (sg >> 8) & 15u;
        // cells dx = -1, 0, +1 in this order, per-cell partial sums; the rejection test on the reference's operands
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) {
            const float tc = xsub(frac, (float)dx);
            const float Rtc = xmul(R, tc);
            float cv = 0.f;
#pragma unroll
            for (int k = 0; k < NIMP; ++k) {
                const float tik = dx < 0 ? im[k] : (dx == 0 ? i0[k] : ip[k]);
                const uint32_t sb = ((dx < 0 ? sm : (dx == 0 ? s0 : sp)) >> k) & 1u;
                const float tp = xsub(tc, tik);
                const float ab = xsub(Rtc, xmul(R, tik));
                float kv = ex2_approx(ka * ab * ab); // exp(-ab^2 / (2 ls^2)): splattingKernel1DVal (GPF:835-841)
                kv = xmul(tp, tp) < 1.0f ? kv : 0.f;
                cv += sb ? kv : -kv;
            }
            sv += cv;
        }
    }
    return sv;
}

// per-ray intermediate state between the three kernels (device scratch, n elements per array)
struct Scratch1D {
    float *c[7];          // conditioning coefficients (value_scale, gradient_scale, ray_origin); null when the medium never conditions
    float *t_hi, *t_lo;   // crossing / exit distance (double split into two floats)
    float *last_val;
    uint32_t *flags;      // bit 0: hit, bit 1: nothing left to do (refused / trivial exit), bits 8-15: scheme at the hit
    uint32_t *evals;
    int32_t *gp_id;       // gp id at the hit: the full int32 of the JSON "id", as the generic and 3D kernels carry it
};
inline Scratch1D scratch1d_carve(void *base, size_t n) {
    Scratch1D X; float *f = (float *)base;
    for (int j = 0; j < 7; ++j) X.c[j] = f + (size_t)j * n;
    X.t_hi = f + 7 * n; X.t_lo = f + 8 * n; X.last_val = f + 9 * n; X.flags = (uint32_t *)(f + 10 * n); X.evals = (uint32_t *)(f + 11 * n);
    X.gp_id = (int32_t *)(f + 12 * n);
    return X;
}

__device__ __forceinline__ Info load_info1d(const Rays &r, const State &s, size_t i) {
    Info in = {r.px ? r.px[i] : 0u, r.py ? r.py[i] : 0u, r.spp ? r.spp[i] : 0u, r.seg ? r.seg[i] : 0u, r.scene_seed, s.info_t[i]};
    return in;
}

// conditioning1D (SCN:519-595) for every ray that carries Renewal state: one thread per ray, all lanes busy
__global__ void __launch_bounds__(MARCH1D_BLOCK) k_cond1d(const __grid_constant__ Params P, const Rays r, const State s, const Scratch1D X, unsigned n) {
    const unsigned i = blockIdx.x * MARCH1D_BLOCK + threadIdx.x;
    if (i >= n) return;
    Coeff c; c.vs = 0.f; c.gs = F3(0, 0, 0); c.org = F3(0, 0, 0);
    if (!s.first_scatter[i]) { // (also for rays sampleDistance will refuse: transmittance has no bounce limit, GPM:343-393)
        const f3 pos = F3(r.ox[i], r.oy[i], r.oz[i]), dir = F3(r.dx[i], r.dy[i], r.dz[i]);
        const Frame fr = ray_frame(P, dir);
        conditioning(P, c, fr, pos, dir, s.last_val[i], F3(s.lgx[i], s.lgy[i], s.lgz[i]), load_info1d(r, s, i));
    }
    X.c[0][i] = c.vs; X.c[1][i] = c.gs.x; X.c[2][i] = c.gs.y; X.c[3][i] = c.gs.z; X.c[4][i] = c.org.x; X.c[5][i] = c.org.y; X.c[6][i] = c.org.z;
}

// VAR: 0 = stationary kernel, 1 = brute-force non-stationary (point-dependent kernel scale), 2 = multi-resolution (two levels, SCN:126-135)
// PROC: the instantiation carries the procedural means (proc_mean: SDF functions, a field behind the mean) and the full anisotropy matrix
// ("useAnisoMtx"); the media without either keep that code compiled out
template <bool TRANSMIT, int NIMP, int VAR, bool PROC = false>
// resident blocks per SM the register allocation must allow: 5 (<= 102 registers) for stationary media, 4 / 3 for the brute-force and
// multi-resolution variants; the calls into the procedural-field code (cold for most media) then spill around the call instead of
// raising the whole kernel's register count
__global__ void __launch_bounds__(MARCH1D_BLOCK, VAR == 0 ? 5 : (VAR == 1 ? 4 : 3)) k_march1d(const __grid_constant__ Params P, const MarchArgs A, const Scratch1D X) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    // world -> the kernel's local frame: the anisotropy scales, or the full matrix of "useAnisoMtx" (PROC instantiations only)
    auto w2l = [&](const f3 v) -> f3 { return (PROC && P.use_mtx) ? mtx_mulv(P.mW, v) : xmul3(l_inv(P), v); };
    const unsigned lt = (1u << lane) - 1u;
    const bool absorb = !TRANSMIT && P.absorption_only;
    const bool cond_on = P.activate_conditioning != 0;
    const float R = P.R_iso, nf = P.nf_1d;
    constexpr bool BRUTE = VAR == 1, MULTI = VAR == 2;
    constexpr int NLEV = MULTI ? 2 : 1;

    int phase = PH_FETCH; unsigned ri = 0;
    dbl3 o = {0, 0, 0}, rd = {0, 0, 0};
    f3 frn = F3(0, 0, 1);
    uint32_t seed = 0, evals = 0; float near_t = 0.f, oz = 0.f, cvs = 0.f, cgz = 0.f;
    f3 corg = F3(0, 0, 0); // conditioning origin (BRUTE: the covariance scale between the point and the origin is point-dependent)
    double t = 0, pf = 0, dstep = 0, dfar = 0, intp = 0, ta = 0, t_prev = 0, t_test = 0;
    int sign0 = 1, gp_id = 0; bool first_step = false, first_scatter = true, conditioned = false;
    float osf_lo = 1.f, osf_hi = 1.f; // MULTI: blend weights of the conditioning origin's two levels (SCN:340-343)
    Win1D<NIMP> W[NLEV];
#pragma unroll
    for (int l = 0; l < NLEV; ++l) {
        W[l].cb = 0; W[l].ck = 0; W[l].valid = false; W[l].sig = 0u;
#pragma unroll
        for (int k = 0; k < NIMP; ++k) { W[l].w0[k] = 0.f; W[l].w1[k] = 0.f; W[l].w2[k] = 0.f; W[l].w3[k] = 0.f; }
    }

    for (;;) {
        // ------------------------------------------------------------ fetch rays for idle lanes (warp-aggregated atomic)
        const unsigned mf = __ballot_sync(FULL, phase == PH_FETCH);
        if (mf) {
            const int leader = __ffs(mf) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(A.counter, (unsigned)__popc(mf));
            base = __shfl_sync(FULL, base, leader);
            if (phase == PH_FETCH) {
                const unsigned idx = base + __popc(mf & lt);
                if (idx >= A.n) phase = PH_EXIT;
                else {
                    ri = idx; SCN_CHECK(ri < A.n, 6);
                    const f3 pos = F3(A.r.ox[ri], A.r.oy[ri], A.r.oz[ri]), dir = F3(A.r.dx[ri], A.r.dy[ri], A.r.dz[ri]);
                    near_t = A.r.tmin[ri]; float far_t = A.r.tmax[ri]; const float u = A.r.u[ri];
                    if (!isfinite(far_t)) far_t = (float)((double)near_t + 2000.0); // GPM:229-231
                    first_scatter = A.s.first_scatter[ri] != 0;
                    evals = 0;
                    bool trivial = false;
                    if (!TRANSMIT) {
                        if (A.s.bounce[ri] >= P.max_bounces) { // GPM:235-237: refused
                            trivial = true;
                            A.o.ok[ri] = 0; A.o.exited[ri] = 0; A.o.t[ri] = 0.f; A.o.cont_t[ri] = 0.f; A.o.gx[ri] = 0.f; A.o.gy[ri] = 0.f; A.o.gz[ri] = 0.f;
                            A.o.gp_id[ri] = A.s.last_gp_id[ri]; A.o.scheme[ri] = SCN_SCHEME_UNI; A.o.weight[ri] = 0.f;
                        } else if (far_t == 0.f) { // GPM:239-248: trivial exit
                            trivial = true;
                            A.o.ok[ri] = 1; A.o.exited[ri] = 1; A.o.t[ri] = far_t; A.o.cont_t[ri] = 0.f; A.o.gx[ri] = 0.f; A.o.gy[ri] = 0.f; A.o.gz[ri] = 0.f;
                            A.o.gp_id[ri] = A.s.last_gp_id[ri]; A.o.scheme[ri] = SCN_SCHEME_UNI; A.o.weight[ri] = 1.f;
                        }
                        if (trivial) {
                            if (A.o.n_evals) A.o.n_evals[ri] = 0;
                            if (A.s.scheme) A.s.scheme[ri] = SCN_SCHEME_UNI;
                            if (A.s.ctx[0]) for (int j = 0; j < 7; ++j) A.s.ctx[j][ri] = 0.f;
                            X.flags[ri] = 2u;
                        }
                    }
                    if (!trivial) {
                        const Info info = load_info1d(A.r, A.s, ri);
                        seed = compute_seed(P, info);
                        o = to_dbl3(pos); rd = to_dbl3(dir);
                        const Frame fr = ray_frame(P, dir);
                        frn = fr.n;
                        float step = xdiv(xsub(far_t, near_t), (float)P.min_step); // SCNM:107-109
                        if (P.step_size < step) step = P.step_size;
                        dstep = (double)step; dfar = (double)far_t;
                        t = (double)xadd(near_t, xmul(step, u)); // SCNM:129
#pragma unroll
                        for (int l = 0; l < NLEV; ++l) W[l].valid = false;
                        conditioned = !first_scatter && cond_on;
                        if (conditioned) { // coefficients from k_cond1d; o_ray of evaluateNoise1DNormalized is constant along the ray
                            cvs = X.c[0][ri]; cgz = X.c[3][ri];
                            corg = F3(X.c[4][ri], X.c[5][ri], X.c[6][ri]);
                            oz = xdot(fr.n, BRUTE ? cov_pos_w2l(P, corg, 1.0f) : w2l(corg));
                            if (MULTI) { const Levels oi = level_ratio(P, corg); osf_lo = oi.r_lo; osf_hi = oi.r_hi; }
                            evals = P.ctx == SCN_CTX_RENEWAL_PLUS ? 2u : 1u;
                        }
                        phase = PH_F0;
                    }
                }
            }
        }
        if (__all_sync(FULL, phase == PH_EXIT)) break;
        const bool active = phase != PH_FETCH && phase != PH_EXIT;

        // ------------------------------------------------------------ evaluateValue (SCN:73-89) through the register cache
        float fval = 0.f; int id = 0;
        f3 p = F3(0.f, 0.f, 0.f);
        if (active) {
            const double te = phase == PH_F0 ? (double)near_t : (phase == PH_MARCH ? t : (phase == PH_REFINE ? t_test : dfar));
            p = ray_at(o, rd, te);
        }
        Levels L; L.s_lo = L.s_hi = 1.f; L.r_lo = 1.f; L.r_hi = 0.f; L.k_lo = 0;
        if (MULTI && active) L = level_ratio(P, p);
        const float lsq = (BRUTE && active) ? cov_ns_scale(P, p) : 1.0f; // the kernel's local scale at the QUERY point (SCN:425-426)
        const float ka = BRUTE ? -0.72134752f / (lsq * lsq) : -0.72134752f;
        float nv = 0.f;
#pragma unroll
        for (int l = 0; l < NLEV; ++l) {
            const float kss = MULTI ? (l ? L.s_hi : L.s_lo) : 1.0f;
            const int kidx = MULTI ? L.k_lo + l : 0;
            const uint32_t seedl = seed + (uint32_t)(MULTI ? P.lvl_add_seed[kidx] : 0);
            // transformPosDirWorldtoLocal (GPF:631-633; stationary: the division by 1 is exact; BRUTE: max ls; MULTI: the level's scale), z of the ray frame
            float z = 0.f, frac = 0.f; int i = 0;
            if (active) {
                z = xdot(frn, MULTI ? xdiv3(w2l(p), kss) : (BRUTE ? cov_pos_w2l(P, p, 1.0f) : w2l(p)));
                const float tg = xdiv(z, R), fl = floorf(tg);
                frac = xsub(tg, fl); i = (int)fl;
            }
            const float sv = win1d_noise<NIMP>(W[l], active, i, frac, seedl, kidx, R, ka, lane);
            float nl = 0.f;
            if (active) {
                nl = sv / (BRUTE ? sqrtf(se_variance1d(P, R, lsq)) : nf); // SCN:334-336
                if (conditioned) { // SCN:338-357
                    const float ozl = MULTI ? xdot(frn, xdiv3(w2l(corg), kss)) : oz;
                    const float ab = z - ozl;
                    const float cls = BRUTE ? cov_splat_cov1d(P, p, corg) : 1.0f;
                    const float osf = MULTI ? (l ? osf_hi : osf_lo) : 1.0f;
                    nl += osf * (cvs * se_c1d_val(ab, cls) + (kss * cgz) * se_c1d_grad(ab, cls));
                }
            }
            nv = MULTI ? nv + (l ? L.r_hi : L.r_lo) * nl : nl; // SCN:131-134
        }
        if (active) {
            const float amp = VAR == 0 ? P.sigma : cov_amplitude(P, p); // stationary instantiations carry no procedural-field code
            double mean; gp_mean_and_id<PROC>(P, p, mean, id);
            if (P.surf_vol_sep) id = 1.f < P.surf_vol_thresh ? 0 : 1;
            fval = (float)yadd((double)xmul(amp, nv), mean);
        }
        const int sgn = (double)fval < 0 ? -1 : 1;

        // ------------------------------------------------------------ state machine (SCNM:102-183)
        bool finished = false, hit = false; double t_fin = 0; float last_val = 0.f;
        if (phase == PH_F0) {
            evals++; gp_id = id;
            sign0 = sgn; pf = (double)fval; first_step = true;
            phase = t < dfar ? PH_MARCH : PH_END;
        } else if (phase == PH_MARCH) {
            evals++; gp_id = id;
            const double fc = (double)fval;
            bool advance = true;
            if (!first_scatter && first_step) sign0 = sgn; // SCNM:138-140
            else if (sgn != sign0) {
                intp = ydiv(pf, ysub(pf, fc)); ta = ysub(t, dstep);
                t_test = yadd(ymul(ta, ysub(1.0, intp)), ymul(t, intp)); t_prev = t_test;
                phase = PH_REFINE; advance = false;
            }
            first_step = false;
            if (advance) { pf = fc; t = yadd(t, dstep); if (!(t < dfar)) phase = PH_END; }
        } else if (phase == PH_REFINE) { // SCNM:144-162
            evals++; gp_id = id;
            bool done = false;
            if (sgn == sign0) done = true;
            else {
                intp = ymul(intp, 0.9);
                if (intp <= 0.01) { t_prev = 0; done = true; }
                else { t_prev = t_test; t_test = yadd(ymul(ta, ysub(1.0, intp)), ymul(t, intp)); }
            }
            if (done) { finished = true; hit = true; t_fin = t_prev; last_val = 0.0f; }
        } else if (phase == PH_END) {
            evals++;
            finished = true; hit = false; t_fin = dfar; last_val = fval;
        }
        if (finished) {
            SCN_CHECK(ri < A.n && active, 7);
            if (TRANSMIT) { // GPM:343-393: the gradient at a crossing is evaluated (one more evaluation) but cannot change T
                A.T[ri] = hit ? 0.f : 1.f;
                if (A.n_evals) A.n_evals[ri] = evals + (hit ? 1u : 0u);
            } else {
                uint32_t scheme = SCN_SCHEME_UNI;
                if (hit) scheme = (VAR == 0 ? P.sigma : cov_amplitude(P, ray_at(o, rd, t_fin))) < 1e-6f ? SCN_SCHEME_UNI : (uint32_t)P.scheme; // samplingScheme (SCN:51-58)
                const float th = (float)t_fin;
                X.t_hi[ri] = th; X.t_lo[ri] = (float)(t_fin - (double)th); X.last_val[ri] = last_val;
                X.flags[ri] = (hit ? 1u : 0u) | (scheme << 8); X.gp_id[ri] = gp_id;
                X.evals[ri] = evals;
            }
            phase = PH_FETCH;
        }
        (void)absorb;
    }
}

// evaluateGradient at the crossing / exit (GPM:276-296, 319; SCN:92-99, 137-239) and the tail of sampleDistance (GPM:303-339):
// one thread per ray, all lanes busy
__global__ void __launch_bounds__(MARCH1D_BLOCK) k_finish1d(const __grid_constant__ Params P, const MarchArgs A, const Scratch1D X) {
    const unsigned ri = blockIdx.x * MARCH1D_BLOCK + threadIdx.x;
    if (ri >= A.n) return;
    const uint32_t fl = X.flags[ri];
    if (fl & 2u) return; // refused / trivial exit: written by k_march1d
    const f3 pos = F3(A.r.ox[ri], A.r.oy[ri], A.r.oz[ri]), dir = F3(A.r.dx[ri], A.r.dy[ri], A.r.dz[ri]);
    float far_t = A.r.tmax[ri];
    if (!isfinite(far_t)) far_t = (float)((double)A.r.tmin[ri] + 2000.0);
    Coeff c; c.vs = 0.f; c.gs = F3(0, 0, 0); c.org = F3(0, 0, 0);
    if (P.activate_conditioning && !A.s.first_scatter[ri]) { c.vs = X.c[0][ri]; c.gs = F3(X.c[1][ri], X.c[2][ri], X.c[3][ri]); c.org = F3(X.c[4][ri], X.c[5][ri], X.c[6][ri]); }
    Finish F; F.hit = (fl & 1u) != 0u; F.grad_valid = false; F.t = (double)X.t_hi[ri] + (double)X.t_lo[ri]; F.grad = F3(0, 0, 0); F.last_val = X.last_val[ri];
    F.gp_id_hit = X.gp_id[ri]; F.scheme_hit = (int)((fl >> 8) & 0xffu); F.evals = X.evals[ri];
    if (F.hit || !P.absorption_only) {
        const Info info = load_info1d(A.r, A.s, ri);
        const uint32_t seed = compute_seed(P, info);
        const Frame fr = ray_frame(P, dir);
        const dbl3 o = to_dbl3(pos), rd = to_dbl3(dir);
        const double inv = ydiv(1.0, ysqrt(yadd(yadd(ymul(rd.x, rd.x), ymul(rd.y, rd.y)), ymul(rd.z, rd.z))));
        const f3 ipt = F3((float)yadd(o.x, ymul(ymul(rd.x, inv), F.t)), (float)yadd(o.y, ymul(ymul(rd.y, inv), F.t)), (float)yadd(o.z, ymul(ymul(rd.z, inv), F.t)));
        F.grad = eval_gradient(P, c, fr, ipt, (float)F.t, dir, info, seed);
        F.evals++;
    }
    finish_ray<false>(P, A, ri, dir, far_t, c, F);
}

// scratch: stream-ordered allocation on the caller's stream (no synchronisation; concurrent streams get distinct buffers) out of
// the MEDIUM's own memory pool — the device's default pool, which belongs to the host application, is never touched
template <bool TRANSMIT>
inline cudaError_t launch_march1d_t(const Params &P, const MarchArgs &A, int sm_count, cudaMemPool_t pool, cudaStream_t s) {
    void *scratch = nullptr;
    cudaError_t e = cudaMallocFromPoolAsync(&scratch, SCRATCH1D_BYTES_PER_RAY * (size_t)A.n, pool, s);
    if (e != cudaSuccess) return e;
    const Scratch1D X = scratch1d_carve(scratch, A.n);
    const unsigned per_ray_grid = (A.n + MARCH1D_BLOCK - 1) / MARCH1D_BLOCK;
    if (P.activate_conditioning) k_cond1d<<<per_ray_grid, MARCH1D_BLOCK, 0, s>>>(P, A.r, A.s, X, A.n);
    const int var = P.multires ? 2 : (P.stationary ? 0 : 1);
    const bool proc = has_procedural_mean(P) || P.use_mtx;
    auto launch = [&](auto kern) -> cudaError_t {
        int per_sm = 0; // resident blocks per SM (registers): one wave of persistent blocks
        cudaError_t er = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MARCH1D_BLOCK, 0);
        if (er != cudaSuccess) return er;
        unsigned grid = (unsigned)(sm_count * (per_sm < 1 ? 1 : per_sm)); if (grid > per_ray_grid) grid = per_ray_grid;
        kern<<<grid, MARCH1D_BLOCK, 0, s>>>(P, A, X);
        return cudaSuccess;
    };
#define SCN_M1D_P(N, PR) (var == 2 ? launch(k_march1d<TRANSMIT, N, 2, PR>) : (var == 1 ? launch(k_march1d<TRANSMIT, N, 1, PR>) : launch(k_march1d<TRANSMIT, N, 0, PR>)))
#define SCN_M1D(N) (proc ? SCN_M1D_P(N, true) : SCN_M1D_P(N, false))
    switch (P.n_imp) {
    case 1: e = SCN_M1D(1); break;
    case 2: e = SCN_M1D(2); break;
    case 3: e = SCN_M1D(3); break;
    default: e = SCN_M1D(4); break;
    }
#undef SCN_M1D
#undef SCN_M1D_P
    if (e != cudaSuccess) { cudaFreeAsync(scratch, s); return e; }
    if (!TRANSMIT) k_finish1d<<<per_ray_grid, MARCH1D_BLOCK, 0, s>>>(P, A, X);
    e = cudaGetLastError();
    cudaError_t e2 = cudaFreeAsync(scratch, s);
    return e != cudaSuccess ? e : e2;
}

// The same three launches over scratch the CALLER of this function owns (scn_sample_distance_batch_static: a file-scope device array,
// scn_api.cu): nothing is allocated or freed here.  `scratch` is a DEVICE address of at least SCRATCH1D_BYTES_PER_RAY * A.n bytes; the part
// in use is cleared in-stream first (the array outlives the call and still holds the previous wavefront's flags and coefficients).
inline cudaError_t launch_march1d_sample_on(const Params &P, const MarchArgs &A, int sm_count, void *scratch, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(scratch, 0, SCRATCH1D_BYTES_PER_RAY * (size_t)A.n, s);
    if (e != cudaSuccess) return e;
    const Scratch1D X = scratch1d_carve(scratch, A.n);
    const unsigned per_ray_grid = (A.n + MARCH1D_BLOCK - 1) / MARCH1D_BLOCK;
    if (P.activate_conditioning) k_cond1d<<<per_ray_grid, MARCH1D_BLOCK, 0, s>>>(P, A.r, A.s, X, A.n);
    const int var = P.multires ? 2 : (P.stationary ? 0 : 1);
    const bool proc = has_procedural_mean(P) || P.use_mtx;
    auto launch = [&](auto kern) -> cudaError_t {
        int per_sm = 0;
        cudaError_t er = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MARCH1D_BLOCK, 0);
        if (er != cudaSuccess) return er;
        unsigned grid = (unsigned)(sm_count * (per_sm < 1 ? 1 : per_sm)); if (grid > per_ray_grid) grid = per_ray_grid;
        kern<<<grid, MARCH1D_BLOCK, 0, s>>>(P, A, X);
        return cudaSuccess;
    };
#define SCN_M1D_P(N, PR) (var == 2 ? launch(k_march1d<false, N, 2, PR>) : (var == 1 ? launch(k_march1d<false, N, 1, PR>) : launch(k_march1d<false, N, 0, PR>)))
#define SCN_M1D(N) (proc ? SCN_M1D_P(N, true) : SCN_M1D_P(N, false))
    switch (P.n_imp) {
    case 1: e = SCN_M1D(1); break;
    case 2: e = SCN_M1D(2); break;
    case 3: e = SCN_M1D(3); break;
    default: e = SCN_M1D(4); break;
    }
#undef SCN_M1D
#undef SCN_M1D_P
    if (e != cudaSuccess) return e;
    k_finish1d<<<per_ray_grid, MARCH1D_BLOCK, 0, s>>>(P, A, X);
    return cudaGetLastError();
}

#endif // SCN_MARCH1D_IMPL

// instantiated in scn_march1d.cu (a translation unit of its own); scn_api.cu calls these
cudaError_t march1d_launch_sample(const Params &P, const MarchArgs &A, int sm_count, cudaMemPool_t pool, cudaStream_t s);
cudaError_t march1d_launch_transmit(const Params &P, const MarchArgs &A, int sm_count, cudaMemPool_t pool, cudaStream_t s);
cudaError_t march1d_launch_sample_static(const Params &P, const MarchArgs &A, int sm_count, void *scratch, cudaStream_t s);

#endif // __CUDACC__

} // namespace scn
#endif
