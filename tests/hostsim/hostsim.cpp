// tests/hostsim/hostsim.cpp — TEST INFRASTRUCTURE ONLY.
//
// Compiles the product's device header (sparse_conv_gpis_tungsten_b200/csrc/scn_device.cuh) for the HOST with g++ so
// that the control flow of the kernels' device functions can be checked against the oracle on the CPU-only CI leg
// (`pytest -m "not gpu"`).  It is never shipped, never loaded by the package, and is not a CPU fallback: the
// product library has no entry point that reaches this file.  Same array-of-structs signatures as
// oracle/ref_harness.cpp (ref_*) and oracle/scn_oracle.c (orc_*), prefix sim_.
#include <cstring>
#include <new>

#include "../../sparse_conv_gpis_tungsten_b200/csrc/scn_device.cuh"

using namespace scn;

struct SimMedium { Params P; };

static Info mkinfo(const uint32_t *pix, size_t i, uint32_t ss, const float *info_t) {
    Info in; in.px = pix ? pix[4 * i] : 0; in.py = pix ? pix[4 * i + 1] : 0; in.spp = pix ? pix[4 * i + 2] : 0; in.seg = pix ? pix[4 * i + 3] : 0;
    in.scene_seed = ss; in.t = info_t ? info_t[i] : 0.f; return in;
}
static Coeff mkcoeff(const float *c, size_t i) {
    Coeff k; k.vs = 0; k.gs = F3(0, 0, 0); k.org = F3(0, 0, 0);
    if (c) { k.vs = c[7 * i]; k.gs = F3(c[7 * i + 1], c[7 * i + 2], c[7 * i + 3]); k.org = F3(c[7 * i + 4], c[7 * i + 5], c[7 * i + 6]); }
    return k;
}
#define P3(a, i) F3((a)[3 * (i)], (a)[3 * (i) + 1], (a)[3 * (i) + 2])

extern "C" {

void *sim_medium_create(const scn_medium_desc *d) {
    SimMedium *m = new SimMedium();
    if (params_from_desc(*d, m->P) != 0) { delete m; return nullptr; }
    return m;
}
void sim_medium_destroy(void *h) { delete (SimMedium *)h; }
const char *sim_last_error() { return last_error(); }

int sim_eval_noise3d(void *h, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int) {
    const Params &P = ((SimMedium *)h)->P;
    Coeff c = mkcoeff(nullptr, 0);
    for (size_t i = 0; i < n; ++i) {
        Frame fr = ray_frame(P, P3(d, i));
        f4 v = P.sampling_1d ? noise1d_norm(P, c, fr, P3(p, i), seed, 1.0f, 0, false, false) : eval_noise3d(P, c, fr, P3(p, i), seed, false);
        out4[4 * i] = v.x; out4[4 * i + 1] = v.y; out4[4 * i + 2] = v.z; out4[4 * i + 3] = v.w;
    }
    return 0;
}
int sim_eval(void *h, size_t n, const float *p, const float *d, const float *t, const uint32_t *pix, uint32_t ss, const float *info_t,
             const float *coeffs, float *out_val, float *out_grad, int *out_gpid, int) {
    const Params &P = ((SimMedium *)h)->P;
    for (size_t i = 0; i < n; ++i) {
        Info in = mkinfo(pix, i, ss, info_t); Coeff c = mkcoeff(coeffs, i);
        Frame fr = ray_frame(P, P3(d, i)); uint32_t seed = compute_seed(P, in);
        if (out_val) { int id; out_val[i] = eval_value(P, c, fr, P3(p, i), seed, id); if (out_gpid) out_gpid[i] = id; }
        if (out_grad) { f3 g = eval_gradient(P, c, fr, P3(p, i), t ? t[i] : 0.f, P3(d, i), in, seed); out_grad[3 * i] = g.x; out_grad[3 * i + 1] = g.y; out_grad[3 * i + 2] = g.z; }
    }
    return 0;
}
int sim_conditioning(void *h, size_t n, const float *p, const float *d, const float *tv, const float *tg, const uint32_t *pix, uint32_t ss,
                     const float *info_t, float *out, int) {
    const Params &P = ((SimMedium *)h)->P;
    for (size_t i = 0; i < n; ++i) {
        Info in = mkinfo(pix, i, ss, info_t); Coeff c = mkcoeff(nullptr, 0);
        Frame fr = ray_frame(P, P3(d, i));
        conditioning(P, c, fr, P3(p, i), P3(d, i), tv[i], P3(tg, i), in);
        float *o = out + 7 * i; o[0] = c.vs; o[1] = c.gs.x; o[2] = c.gs.y; o[3] = c.gs.z; o[4] = c.org.x; o[5] = c.org.y; o[6] = c.org.z;
    }
    return 0;
}
int sim_sample_distance(void *h, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                        const uint32_t *pix, uint32_t ss, uint8_t *fs, float *lv, double *lg, int *bounce, int *lgp, float *info_t,
                        uint8_t *ok, uint8_t *exited, float *t_out, float *cont_t, double *grad, int *gp_id, uint8_t *scheme,
                        float *coeffs, float *weight, int) {
    const Params &P = ((SimMedium *)h)->P;
    for (size_t i = 0; i < n; ++i) {
        RayState st; st.first_scatter = fs[i] != 0; st.last_val = lv[i]; st.last_aniso = F3((float)lg[3 * i], (float)lg[3 * i + 1], (float)lg[3 * i + 2]);
        st.bounce = bounce[i]; st.last_gp_id = lgp[i]; st.scheme = SCN_SCHEME_UNI; st.info = mkinfo(pix, i, ss, info_t); st.c = mkcoeff(nullptr, 0);
        Sample sm;
        medium_query(P, P3(o, i), P3(d, i), tmin[i], tmax[i], u[i], st, sm, false);
        ok[i] = sm.ok; exited[i] = sm.exited; t_out[i] = sm.t; cont_t[i] = sm.cont_t;
        grad[3 * i] = sm.aniso.x; grad[3 * i + 1] = sm.aniso.y; grad[3 * i + 2] = sm.aniso.z;
        gp_id[i] = sm.ok ? sm.gp_id : st.last_gp_id; scheme[i] = (uint8_t)sm.scheme;
        if (weight) weight[3 * i] = weight[3 * i + 1] = weight[3 * i + 2] = sm.weight;
        float *c = coeffs + 7 * i; c[0] = st.c.vs; c[1] = st.c.gs.x; c[2] = st.c.gs.y; c[3] = st.c.gs.z; c[4] = st.c.org.x; c[5] = st.c.org.y; c[6] = st.c.org.z;
        fs[i] = st.first_scatter; lv[i] = st.last_val; lg[3 * i] = st.last_aniso.x; lg[3 * i + 1] = st.last_aniso.y; lg[3 * i + 2] = st.last_aniso.z;
        bounce[i] = st.bounce; lgp[i] = st.last_gp_id; info_t[i] = st.info.t;
    }
    return 0;
}
int sim_transmittance(void *h, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                      const uint32_t *pix, uint32_t ss, const uint8_t *fs, const float *lv, const double *lg, const int *bounce,
                      const int *lgp, const float *info_t, float *T, int) {
    const Params &P = ((SimMedium *)h)->P;
    for (size_t i = 0; i < n; ++i) {
        RayState st; st.first_scatter = fs[i] != 0; st.last_val = lv[i]; st.last_aniso = F3((float)lg[3 * i], (float)lg[3 * i + 1], (float)lg[3 * i + 2]);
        st.bounce = bounce[i]; st.last_gp_id = lgp[i]; st.scheme = SCN_SCHEME_UNI; st.info = mkinfo(pix, i, ss, info_t); st.c = mkcoeff(nullptr, 0);
        Sample sm;
        medium_query(P, P3(o, i), P3(d, i), tmin[i], tmax[i], u[i], st, sm, true);
        T[i] = sm.weight;
    }
    return 0;
}
int sim_nee(void *h, size_t n, const float *dir, const float *nrm, const float *p, const float *ts, const uint32_t *pix, uint32_t ss,
            const float *info_t, const float *coeffs, float *pdf, float *grad, int) {
    const Params &P = ((SimMedium *)h)->P;
    for (size_t i = 0; i < n; ++i) {
        Info in = mkinfo(pix, i, ss, info_t); Coeff c = mkcoeff(coeffs, i);
        if (pdf) pdf[i] = nee_pdf(P, c, P3(dir, i), P3(nrm, i), P3(p, i), ts[i], in);
        if (grad) { Nee o; nee_shared(P, c, P3(dir, i), P3(nrm, i), P3(p, i), in, o); grad[3 * i] = o.dist * nrm[3 * i]; grad[3 * i + 1] = o.dist * nrm[3 * i + 1]; grad[3 * i + 2] = o.dist * nrm[3 * i + 2]; }
    }
    return 0;
}
}
