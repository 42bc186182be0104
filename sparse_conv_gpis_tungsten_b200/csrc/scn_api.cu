// scn_api.cu — C ABI (include/scn.h) + kernels of the B200-native sparse-convolution-noise GPIS medium.
//
// One thread marches one ray / evaluates one point; per-ray Renewal state (Coeff: 7 floats) stays in registers for
// the whole query.  There is no CPU fallback: without a CUDA device every compute entry point returns SCN_ECUDA.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <new>

#include "scn_device.cuh"
#include "scn_fast.cuh"
#include "scn_march.cuh"
#include "scn_march1d.cuh"
#include <atomic>

namespace scn {

static int cuda_fail(cudaError_t e, const char *what) { return set_error(SCN_ECUDA, "%s: %s", what, cudaGetErrorString(e)); }
#define SCN_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail(e_, #x); } while (0)

// ---------------------------------------------------------------- kernels
constexpr int BLOCK = 128;

struct EvalIn { const float *px, *py, *pz, *dx, *dy, *dz, *t, *info_t; const uint32_t *ix, *iy, *is, *ig; uint32_t scene_seed, noise_seed; const float *ctx[7]; };
struct EvalOut { float *val, *gx, *gy, *gz; int32_t *gp_id; };

__device__ __forceinline__ Info load_info(const uint32_t *ix, const uint32_t *iy, const uint32_t *is, const uint32_t *ig, uint32_t scene_seed, const float *info_t, size_t i) {
    Info in; in.px = ix ? ix[i] : 0u; in.py = iy ? iy[i] : 0u; in.spp = is ? is[i] : 0u; in.seg = ig ? ig[i] : 0u; in.scene_seed = scene_seed; in.t = info_t ? info_t[i] : 0.f;
    return in;
}
__device__ __forceinline__ Coeff load_coeff(const float *const ctx[7], size_t i) {
    Coeff c; c.vs = 0.f; c.gs = F3(0.f, 0.f, 0.f); c.org = F3(0.f, 0.f, 0.f);
    if (ctx[0]) { c.vs = ctx[0][i]; c.gs = F3(ctx[1][i], ctx[2][i], ctx[3][i]); c.org = F3(ctx[4][i], ctx[5][i], ctx[6][i]); }
    return c;
}

// generic noise-level evaluation (any configuration)
__global__ void __launch_bounds__(BLOCK) k_eval_noise(const __grid_constant__ Params P, const EvalIn in, const EvalOut out, size_t n) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    f3 p = F3(in.px[i], in.py[i], in.pz[i]), d = F3(in.dx[i], in.dy[i], in.dz[i]);
    Coeff c; c.vs = 0.f; c.gs = F3(0.f, 0.f, 0.f); c.org = F3(0.f, 0.f, 0.f);
    Frame fr = ray_frame(P, d);
    f4 v = P.sampling_1d ? noise1d_norm(P, c, fr, p, in.noise_seed, 1.0f, 0, false, false) : eval_noise3d(P, c, fr, p, in.noise_seed, false);
    if (out.val) out.val[i] = v.x;
    if (out.gx) { out.gx[i] = v.y; out.gy[i] = v.z; out.gz[i] = v.w; }
}

__global__ void __launch_bounds__(BLOCK) k_eval(const __grid_constant__ Params P, const EvalIn in, const EvalOut out, size_t n, int what) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    f3 p = F3(in.px[i], in.py[i], in.pz[i]), d = F3(in.dx[i], in.dy[i], in.dz[i]);
    Info info = load_info(in.ix, in.iy, in.is, in.ig, in.scene_seed, in.info_t, i);
    Coeff c = load_coeff(in.ctx, i);
    Frame fr = ray_frame(P, d);
    uint32_t seed = compute_seed(P, info);
    if (what & SCN_EVAL_VALUE) { int id; float v = eval_value(P, c, fr, p, seed, id); out.val[i] = v; if (out.gp_id) out.gp_id[i] = id; }
    if (what & SCN_EVAL_GRADIENT) { f3 g = eval_gradient(P, c, fr, p, in.t ? in.t[i] : 0.f, d, info, seed); out.gx[i] = g.x; out.gy[i] = g.y; out.gz[i] = g.z; }
}

struct CondArgs { const float *tv, *tgx, *tgy, *tgz; float *ctx[7]; };
__global__ void __launch_bounds__(BLOCK) k_conditioning(const __grid_constant__ Params P, const EvalIn in, const CondArgs a, size_t n) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    f3 p = F3(in.px[i], in.py[i], in.pz[i]), d = F3(in.dx[i], in.dy[i], in.dz[i]);
    Info info = load_info(in.ix, in.iy, in.is, in.ig, in.scene_seed, in.info_t, i);
    Coeff c; c.vs = 0.f; c.gs = F3(0.f, 0.f, 0.f); c.org = F3(0.f, 0.f, 0.f);
    Frame fr = ray_frame(P, d);
    conditioning(P, c, fr, p, d, a.tv[i], F3(a.tgx[i], a.tgy[i], a.tgz[i]), info);
    a.ctx[0][i] = c.vs; a.ctx[1][i] = c.gs.x; a.ctx[2][i] = c.gs.y; a.ctx[3][i] = c.gs.z; a.ctx[4][i] = c.org.x; a.ctx[5][i] = c.org.y; a.ctx[6][i] = c.org.z;
}


__device__ __forceinline__ RayState load_state(const State &s, const Rays &r, size_t i) {
    RayState st;
    st.first_scatter = s.first_scatter[i] != 0; st.last_val = s.last_val[i]; st.last_aniso = F3(s.lgx[i], s.lgy[i], s.lgz[i]);
    st.bounce = s.bounce[i]; st.last_gp_id = s.last_gp_id[i]; st.scheme = SCN_SCHEME_UNI;
    st.info = load_info(r.px, r.py, r.spp, r.seg, r.scene_seed, s.info_t, i);
    st.c.vs = 0.f; st.c.gs = F3(0.f, 0.f, 0.f); st.c.org = F3(0.f, 0.f, 0.f);
    return st;
}

__global__ void __launch_bounds__(BLOCK) k_sample_distance(const __grid_constant__ Params P, const Rays r, const State s, const Samples o, size_t n) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    RayState st = load_state(s, r, i);
    Sample sm;
    medium_query(P, F3(r.ox[i], r.oy[i], r.oz[i]), F3(r.dx[i], r.dy[i], r.dz[i]), r.tmin[i], r.tmax[i], r.u[i], st, sm, false);
    o.ok[i] = sm.ok; o.exited[i] = sm.exited; o.t[i] = sm.t; o.cont_t[i] = sm.cont_t;
    o.gx[i] = sm.aniso.x; o.gy[i] = sm.aniso.y; o.gz[i] = sm.aniso.z;
    o.gp_id[i] = sm.ok ? sm.gp_id : st.last_gp_id; o.scheme[i] = (uint8_t)sm.scheme; o.weight[i] = sm.weight;
    if (o.n_evals) o.n_evals[i] = sm.n_evals;
    s.first_scatter[i] = st.first_scatter; s.last_val[i] = st.last_val; s.lgx[i] = st.last_aniso.x; s.lgy[i] = st.last_aniso.y; s.lgz[i] = st.last_aniso.z;
    s.bounce[i] = st.bounce; s.last_gp_id[i] = st.last_gp_id; s.info_t[i] = st.info.t; if (s.scheme) s.scheme[i] = (uint8_t)st.scheme;
    if (s.ctx[0]) { s.ctx[0][i] = st.c.vs; s.ctx[1][i] = st.c.gs.x; s.ctx[2][i] = st.c.gs.y; s.ctx[3][i] = st.c.gs.z; s.ctx[4][i] = st.c.org.x; s.ctx[5][i] = st.c.org.y; s.ctx[6][i] = st.c.org.z; }
}

__global__ void __launch_bounds__(BLOCK) k_transmittance(const __grid_constant__ Params P, const Rays r, const State s, float *T, uint32_t *n_evals, size_t n) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    RayState st = load_state(s, r, i);
    Sample sm;
    medium_query(P, F3(r.ox[i], r.oy[i], r.oz[i]), F3(r.dx[i], r.dy[i], r.dz[i]), r.tmin[i], r.tmax[i], r.u[i], st, sm, true);
    T[i] = sm.weight;
    if (n_evals) n_evals[i] = sm.n_evals;
}

struct NeeIn { const float *dx, *dy, *dz, *nx, *ny, *nz, *px, *py, *pz, *ts, *info_t; const uint32_t *ix, *iy, *is, *ig; uint32_t scene_seed; const float *ctx[7]; };
__global__ void __launch_bounds__(BLOCK) k_nee(const __grid_constant__ Params P, const NeeIn in, float *pdf, float *gx, float *gy, float *gz, size_t n) {
    size_t i = (size_t)blockIdx.x * BLOCK + threadIdx.x;
    if (i >= n) return;
    f3 d = F3(in.dx[i], in.dy[i], in.dz[i]), nn = F3(in.nx[i], in.ny[i], in.nz[i]), p = F3(in.px[i], in.py[i], in.pz[i]);
    Info info = load_info(in.ix, in.iy, in.is, in.ig, in.scene_seed, in.info_t, i);
    Coeff c = load_coeff(in.ctx, i);
    if (pdf) pdf[i] = nee_pdf(P, c, d, nn, p, in.ts ? in.ts[i] : 0.f, info);
    if (gx) { Nee o; nee_shared(P, c, d, nn, p, info, o); gx[i] = o.dist * nn.x; gy[i] = o.dist * nn.y; gz[i] = o.dist * nn.z; }
}

__global__ void k_debug_hash4(const uint32_t *a, const uint32_t *b, const uint32_t *c, const uint32_t *d, uint32_t *out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = xxhash32_4(a[i], b[i], c[i], d[i]);
}
__global__ void k_debug_pcg(const uint32_t *seed32, float *out, size_t n, int draws) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Pcg g; g.seed32(seed32[i] - 1u); // seed32() adds the reference's +1u
    for (int k = 0; k < draws; ++k) out[i * (size_t)draws + k] = g.next();
}
__global__ void k_debug_cell(const uint32_t *ci, const uint32_t *cj, const uint32_t *ck, uint32_t seed, uint32_t count, float *out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Pcg g; g.seed32(xxhash32_4(ck[i], cj[i], ci[i], seed));
    for (uint32_t k = 0; k < count; ++k) {
        float z = g.next(), y = g.next(), x = g.next(), w = g.next() < 0.5f ? -1.f : 1.f;
        float *o = out + (i * (size_t)count + k) * 4; o[0] = x; o[1] = y; o[2] = z; o[3] = w;
    }
}

// rejection test only, same operands as k_fast_noise3d / noise3d
__global__ void k_debug_accept(const __grid_constant__ Params P, const float *px, const float *py, const float *pz, const float *dx, const float *dy, const float *dz,
                               uint32_t seed, uint32_t *out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    f3 q = F3(px[i], py[i], pz[i]); float R = P.R_world;
    if (P.iso3d) { Frame fr = ray_frame(P, F3(dx[i], dy[i], dz[i])); q = to_local(fr, cov_pos_w2l(P, q, 1.0f)); R = P.R_iso; }
    f3 pg = xdiv3(q, R); f3 fl = F3(floorf(pg.x), floorf(pg.y), floorf(pg.z)); f3 frac = xsub3(pg, fl);
    int ix = (int)fl.x, iy = (int)fl.y, iz = (int)fl.z; uint32_t cnt = 0;
    for (int a = -1; a <= 1; ++a) for (int b = -1; b <= 1; ++b) for (int c = -1; c <= 1; ++c) {
        Pcg g; g.seed32(xxhash32_4((uint32_t)(iz + c), (uint32_t)(iy + b), (uint32_t)(ix + a), seed));
        float cx = xsub(frac.x, (float)a), cy = xsub(frac.y, (float)b), cz = xsub(frac.z, (float)c);
        for (uint32_t k = 0; k < P.n_imp; ++k) {
            float z = g.next(), y = g.next(), x = g.next(); (void)g.next();
            float tx = xsub(cx, x), ty = xsub(cy, y), tz = xsub(cz, z);
            if (xadd(xadd(xmul(tx, tx), xmul(ty, ty)), xmul(tz, tz)) < 1.0f) cnt++;
        }
    }
    out[i] = cnt;
}

} // namespace scn

// ---------------------------------------------------------------- C ABI
using namespace scn;

struct scn_medium {
    scn_medium_desc desc;
    Params P;
    int device;
    // staging arena of the *_host entry points
    std::mutex mu;
    cudaStream_t streams[3] = {nullptr, nullptr, nullptr};
    char *arena[3] = {nullptr, nullptr, nullptr};
    size_t arena_cap = 0;
    cudaEvent_t ev_in[3] = {nullptr, nullptr, nullptr}, ev_k[3] = {nullptr, nullptr, nullptr}, ev_out[3] = {nullptr, nullptr, nullptr};
    // k_march: ring of ray counters (one per launch in flight, zeroed in-stream before the launch) and the SM count
    static constexpr unsigned COUNTERS = 1024;
    unsigned int *counters = nullptr;
    mutable std::atomic<unsigned> next_counter{0};
    int sm_count = 0;
    int march_block = 0; // 128 / 64, 0 = generic kernels only
    bool march1d = false; // 1D-sampling media covered by k_march1d (scn_march1d.cuh)
    cudaMemPool_t pool = nullptr; // k_march1d's per-call scratch (stream-ordered allocations of this medium only)
    bool force_generic = false;   // scn_debug_force_generic: tests compare the tuned kernels with the one-thread-per-ray ones
    double *grids[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; // device copies of the regular_grid fields ("ls", "var", "aniso", the two procedural means' "f")
};

static inline unsigned grid_for(size_t n) { return (unsigned)((n + BLOCK - 1) / BLOCK); }
static int check_launch(const char *what) { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) return cuda_fail(e, what); return SCN_OK; }

struct DeviceGuard {
    int prev = -1; bool ok = true;
    explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; } if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// cached wavefront march (scn_march.cuh, scn_march1d.cuh) for the configurations it covers
static bool use_march(const scn_medium *m) { return (m->march_block != 0 || m->march1d) && !m->force_generic; }
constexpr size_t MAX_WAVEFRONT = 0x40000000u; // the persistent kernels' ray counter is 32 bits wide
template <bool TRANSMIT>
static int launch_march(const scn_medium *m, const Rays &kr, const State &ks, const Samples &ko, float *T, uint32_t *n_evals, size_t n, cudaStream_t s) {
    MarchArgs A; A.r = kr; A.s = ks; A.o = ko; A.T = T; A.n_evals = n_evals; A.n = (unsigned)n; A.K = march_cache_entries(m->P.n_imp, march_rayz(m->P));
    A.counter = m->counters + (m->next_counter.fetch_add(1) % scn_medium::COUNTERS);
    SCN_CUDA(cudaMemsetAsync(A.counter, 0, sizeof(unsigned int), s));
    cudaError_t e = m->march1d ? (TRANSMIT ? march1d_launch_transmit(m->P, A, m->sm_count, m->pool, s) : march1d_launch_sample(m->P, A, m->sm_count, m->pool, s))
                               : (TRANSMIT ? march_launch_transmit(m->P, A, m->march_block, m->sm_count, s) : march_launch_sample(m->P, A, m->march_block, m->sm_count, s));
    if (e != cudaSuccess) return cuda_fail(e, m->march1d ? "k_march1d" : "k_march");
    return SCN_OK;
}

// scn_sample_distance_batch_static: k_march1d's per-ray scratch for callers that may not allocate.  One array per GPU that loads the
// library, shared by every call on that GPU.  In host code `g_static_scratch1d` is only the symbol's host-side stand-in, never a device
// address: the address the kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_static_scratch1d[SCRATCH1D_BYTES_PER_RAY * (size_t)SCN_STATIC_SCRATCH_MAX_RAYS];
constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
static std::mutex g_static_scratch_mu;
static void *g_static_scratch_ptr[STATIC_SCRATCH_MAX_DEVICES] = {};
// the calling thread's current device must be `device`
static int static_scratch_for(int device, void **out) {
    if (device < 0 || device >= STATIC_SCRATCH_MAX_DEVICES) return set_error(SCN_EINVAL, "static scratch: device %d beyond the %d this library keeps an address for", device, STATIC_SCRATCH_MAX_DEVICES);
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (!g_static_scratch_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch1d);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
        if (!p) return set_error(SCN_ECUDA, "cudaGetSymbolAddress returned a null address for the static scratch on device %d", device);
        g_static_scratch_ptr[device] = p;
    }
    *out = g_static_scratch_ptr[device];
    return SCN_OK;
}

extern "C" {

const char *scn_last_error(void) { return last_error(); }
int scn_abi_version(void) { return SCN_ABI_VERSION; }
// internal (not in include/scn.h): the wavefront stage's NEE kernels (scn_render.cu) call the medium's device functions directly
extern "C" const void *scn_internal_params(const scn_medium *m) { return m ? &m->P : nullptr; }
int scn_desc_default(scn_medium_desc *d) { return desc_default(d); }
int scn_desc_from_json(const char *json, scn_medium_desc *d) { return desc_from_json(json, nullptr, d); }
int scn_desc_from_json_at(const char *json, const char *base_dir, scn_medium_desc *d) { return desc_from_json(json, base_dir, d); }
void scn_desc_release(scn_medium_desc *d) { desc_release(d); }

int scn_medium_create(const scn_medium_desc *desc, int device, scn_medium **out) {
    if (!desc || !out) return set_error(SCN_EINVAL, "null argument");
    *out = nullptr;
    Params P;
    if (int rc = params_from_desc(*desc, P)) return rc;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return set_error(SCN_ECUDA, "no CUDA device: this library has no CPU path (%s)", e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
    if (device < 0 || device >= count) return set_error(SCN_EINVAL, "device %d out of range (0..%d)", device, count - 1);
    scn_medium *m = new (std::nothrow) scn_medium();
    if (!m) return set_error(SCN_ENOMEM, "out of memory");
    m->desc = *desc; m->P = P; m->device = device;
    m->march_block = march_block_size(P);
    m->march1d = march1d_supported(P);
    {
        DeviceGuard g(device);
        cudaError_t e1 = cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, device);
        cudaError_t e2 = cudaMalloc((void **)&m->counters, scn_medium::COUNTERS * sizeof(unsigned int));
        if (e1 != cudaSuccess || e2 != cudaSuccess) { delete m; return cuda_fail(e1 != cudaSuccess ? e1 : e2, "scn_medium_create"); }
        // regular_grid fields: the descriptor's host arrays -> device memory owned by the medium (the kernels read them through Params)
        FieldK *fk[5] = {&m->P.ls, &m->P.var, &m->P.aniso_f, &m->P.mean_f, &m->P.mean_add_f};
        scn_field_desc *fd[5] = {&m->desc.ls, &m->desc.var, &m->desc.aniso_field, &m->desc.mean.f, &m->desc.mean_additional.f};
        for (int i = 0; i < 5; ++i) {
            fd[i]->grid.values = nullptr; fd[i]->grid.owned = 0; // the medium keeps no reference to the caller's arrays
            if (fk[i]->type != SCN_FIELD_REGULAR_GRID) { fk[i]->grid = nullptr; continue; }
            const size_t bytes = (size_t)fk[i]->gres * fk[i]->gres * fk[i]->gres * fk[i]->gcomp * sizeof(double);
            cudaError_t eg = cudaMalloc((void **)&m->grids[i], bytes);
            if (eg == cudaSuccess) eg = cudaMemcpy(m->grids[i], fk[i]->grid, bytes, cudaMemcpyHostToDevice);
            if (eg != cudaSuccess) { for (int j = 0; j < 5; ++j) if (m->grids[j]) cudaFree(m->grids[j]); cudaFree(m->counters); delete m; return cuda_fail(eg, "scn_medium_create: regular_grid upload"); }
            fk[i]->grid = m->grids[i];
        }
        if (m->march1d) { // k_march1d's per-call scratch: a pool of this medium's own that keeps freed blocks cached instead of returning them at every sync
            cudaMemPoolProps props{}; props.allocType = cudaMemAllocationTypePinned; props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice; props.location.id = device;
            unsigned long long keep = ~0ULL;
            cudaError_t e3 = cudaMemPoolCreate(&m->pool, &props);
            if (e3 == cudaSuccess) e3 = cudaMemPoolSetAttribute(m->pool, cudaMemPoolAttrReleaseThreshold, &keep);
            if (e3 != cudaSuccess) { if (m->pool) cudaMemPoolDestroy(m->pool); cudaFree(m->counters); delete m; return cuda_fail(e3, "scn_medium_create: memory pool"); }
        }
    }
    *out = m;
    return SCN_OK;
}

void scn_medium_destroy(scn_medium *m) {
    if (!m) return;
    {
        DeviceGuard g(m->device);
        for (int i = 0; i < 3; ++i) { if (m->arena[i]) cudaFree(m->arena[i]); if (m->streams[i]) cudaStreamDestroy(m->streams[i]);
                                      if (m->ev_in[i]) cudaEventDestroy(m->ev_in[i]); if (m->ev_k[i]) cudaEventDestroy(m->ev_k[i]); if (m->ev_out[i]) cudaEventDestroy(m->ev_out[i]); }
        if (m->counters) cudaFree(m->counters);
        for (int i = 0; i < 5; ++i) if (m->grids[i]) cudaFree(m->grids[i]);
        if (m->pool) { cudaDeviceSynchronize(); cudaMemPoolDestroy(m->pool); }
    }
    delete m;
}

static void to_eval_in(const scn_eval_in *in, EvalIn &k) {
    k.px = in->px; k.py = in->py; k.pz = in->pz; k.dx = in->dx; k.dy = in->dy; k.dz = in->dz; k.t = in->t; k.info_t = in->info_t;
    k.ix = in->pix_x; k.iy = in->pix_y; k.is = in->pix_spp; k.ig = in->pix_seg; k.scene_seed = in->scene_seed; k.noise_seed = in->noise_seed;
    bool all = true; for (int i = 0; i < 7; ++i) all = all && in->ctx[i];
    for (int i = 0; i < 7; ++i) k.ctx[i] = all ? in->ctx[i] : nullptr;
}

int scn_eval_batch(const scn_medium *m, const scn_eval_in *in, const scn_eval_out *out, size_t n, int what, void *stream) {
    if (!m || !in || !out) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    if (!in->px || !in->py || !in->pz || !in->dx || !in->dy || !in->dz) return set_error(SCN_EINVAL, "eval: points and directions are required");
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    cudaStream_t s = (cudaStream_t)stream;
    EvalIn ki; to_eval_in(in, ki);
    EvalOut ko; ko.val = out->val; ko.gx = out->gx; ko.gy = out->gy; ko.gz = out->gz; ko.gp_id = out->gp_id;
    if (what == SCN_EVAL_NOISE) {
        if (!ko.val && !ko.gx) return set_error(SCN_EINVAL, "eval: no output requested");
        if (ko.gx && (!ko.gy || !ko.gz)) return set_error(SCN_EINVAL, "eval: gx, gy, gz go together");
        if (fast_noise3d_supported(m->P) && ko.val && ko.gx) launch_fast_noise3d(m->P, ki.px, ki.py, ki.pz, ki.dx, ki.dy, ki.dz, ki.noise_seed, ko.val, ko.gx, ko.gy, ko.gz, n, s);
        else k_eval_noise<<<grid_for(n), BLOCK, 0, s>>>(m->P, ki, ko, n);
        return check_launch("k_eval_noise");
    }
    if (!(what & (SCN_EVAL_VALUE | SCN_EVAL_GRADIENT)) || (what & ~(SCN_EVAL_VALUE | SCN_EVAL_GRADIENT))) return set_error(SCN_EINVAL, "eval: bad `what` %d", what);
    if ((what & SCN_EVAL_VALUE) && !ko.val) return set_error(SCN_EINVAL, "eval: val output required");
    if ((what & SCN_EVAL_GRADIENT) && (!ko.gx || !ko.gy || !ko.gz)) return set_error(SCN_EINVAL, "eval: gradient outputs required");
    k_eval<<<grid_for(n), BLOCK, 0, s>>>(m->P, ki, ko, n, what);
    return check_launch("k_eval");
}

int scn_conditioning_batch(const scn_medium *m, const scn_eval_in *in, const float *tv, const float *tgx, const float *tgy, const float *tgz,
                           float *const ctx_out[7], size_t n, void *stream) {
    if (!m || !in || !tv || !tgx || !tgy || !tgz || !ctx_out) return set_error(SCN_EINVAL, "null argument");
    for (int i = 0; i < 7; ++i) if (!ctx_out[i]) return set_error(SCN_EINVAL, "conditioning: all 7 coefficient outputs are required");
    if (n == 0) return SCN_OK;
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    EvalIn ki; to_eval_in(in, ki);
    CondArgs a; a.tv = tv; a.tgx = tgx; a.tgy = tgy; a.tgz = tgz; for (int i = 0; i < 7; ++i) a.ctx[i] = ctx_out[i];
    k_conditioning<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(m->P, ki, a, n);
    return check_launch("k_conditioning");
}

static int to_rays(const scn_rays *r, Rays &k) {
    if (!r->ox || !r->oy || !r->oz || !r->dx || !r->dy || !r->dz || !r->tmin || !r->tmax || !r->u) return set_error(SCN_EINVAL, "rays: origin, direction, tmin, tmax and u are required");
    k.ox = r->ox; k.oy = r->oy; k.oz = r->oz; k.dx = r->dx; k.dy = r->dy; k.dz = r->dz; k.tmin = r->tmin; k.tmax = r->tmax; k.u = r->u;
    k.px = r->px; k.py = r->py; k.spp = r->spp; k.seg = r->segment; k.scene_seed = r->scene_seed;
    return SCN_OK;
}
static int to_state(const scn_state *s, State &k) {
    if (!s->first_scatter || !s->last_val || !s->last_gx || !s->last_gy || !s->last_gz || !s->bounce || !s->last_gp_id || !s->info_t)
        return set_error(SCN_EINVAL, "state: first_scatter, last_val, last_g{x,y,z}, bounce, last_gp_id and info_t are required");
    k.first_scatter = s->first_scatter; k.last_val = s->last_val; k.lgx = s->last_gx; k.lgy = s->last_gy; k.lgz = s->last_gz;
    k.bounce = s->bounce; k.last_gp_id = s->last_gp_id; k.info_t = s->info_t; k.scheme = s->scheme;
    bool all = true; for (int i = 0; i < 7; ++i) all = all && s->ctx[i];
    for (int i = 0; i < 7; ++i) k.ctx[i] = all ? s->ctx[i] : nullptr;
    return SCN_OK;
}

int scn_sample_distance_batch(const scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, size_t n, void *stream) {
    if (!m || !rays || !state || !out) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK; // an empty wavefront is a no-op (empty buffers may be NULL)
    if (!out->ok || !out->exited || !out->t || !out->continued_t || !out->gx || !out->gy || !out->gz || !out->gp_id || !out->scheme || !out->weight)
        return set_error(SCN_EINVAL, "samples: every output but n_evals is required");
    if (n > MAX_WAVEFRONT) return set_error(SCN_EINVAL, "wavefronts above 2^30 rays are not supported (split the batch)"); // before anything is enqueued: the state is updated in place
    Rays kr; State ks;
    if (int rc = to_rays(rays, kr)) return rc;
    if (int rc = to_state(state, ks)) return rc;
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    Samples ko; ko.ok = out->ok; ko.exited = out->exited; ko.t = out->t; ko.cont_t = out->continued_t; ko.gx = out->gx; ko.gy = out->gy; ko.gz = out->gz;
    ko.gp_id = out->gp_id; ko.scheme = out->scheme; ko.weight = out->weight; ko.n_evals = out->n_evals;
    if (use_march(m)) return launch_march<false>(m, kr, ks, ko, nullptr, nullptr, n, (cudaStream_t)stream);
    k_sample_distance<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(m->P, kr, ks, ko, n);
    return check_launch("k_sample_distance");
}

// No counterpart in the reference (include/scn.h): k_cond1d -> k_march1d -> k_finish1d over the file-scope scratch array; no allocation.
int scn_sample_distance_batch_static(const scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, size_t n, void *stream) {
    if (!m || !rays || !state || !out) return set_error(SCN_EINVAL, "null argument");
    if (!m->march1d) return set_error(SCN_EUNSUPPORTED, "scn_sample_distance_batch_static serves the media whose kernel needs scratch memory (1D sampling, SCN_KERNEL_MARCH1D); "
                                                        "this medium's kernel uses none: call scn_sample_distance_batch, which allocates nothing for it");
    if (n > (size_t)SCN_STATIC_SCRATCH_MAX_RAYS) return set_error(SCN_EINVAL, "scn_sample_distance_batch_static: %zu rays exceed the static scratch of SCN_STATIC_SCRATCH_MAX_RAYS = %d rays "
                                                                  "(split the wavefront or call scn_sample_distance_batch)", n, (int)SCN_STATIC_SCRATCH_MAX_RAYS);
    if (n == 0) return SCN_OK;
    if (!out->ok || !out->exited || !out->t || !out->continued_t || !out->gx || !out->gy || !out->gz || !out->gp_id || !out->scheme || !out->weight)
        return set_error(SCN_EINVAL, "samples: every output but n_evals is required");
    Rays kr; State ks;
    if (int rc = to_rays(rays, kr)) return rc;
    if (int rc = to_state(state, ks)) return rc;
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    void *scratch = nullptr;
    if (int rc = static_scratch_for(m->device, &scratch)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    MarchArgs A; A.r = kr; A.s = ks; A.T = nullptr; A.n_evals = nullptr; A.n = (unsigned)n; A.K = march_cache_entries(m->P.n_imp, march_rayz(m->P));
    A.o.ok = out->ok; A.o.exited = out->exited; A.o.t = out->t; A.o.cont_t = out->continued_t; A.o.gx = out->gx; A.o.gy = out->gy; A.o.gz = out->gz;
    A.o.gp_id = out->gp_id; A.o.scheme = out->scheme; A.o.weight = out->weight; A.o.n_evals = out->n_evals;
    A.counter = m->counters + (m->next_counter.fetch_add(1) % scn_medium::COUNTERS); // the medium's ring, allocated at scn_medium_create
    SCN_CUDA(cudaMemsetAsync(A.counter, 0, sizeof(unsigned int), s));
    cudaError_t e = march1d_launch_sample_static(m->P, A, m->sm_count, scratch, s);
    if (e != cudaSuccess) return cuda_fail(e, "k_march1d (static scratch)");
    return SCN_OK;
}

static int transmittance_impl(const scn_medium *m, const scn_rays *rays, const scn_state *state, float *T, uint32_t *n_evals, size_t n, cudaStream_t s) {
    if (n > MAX_WAVEFRONT) return set_error(SCN_EINVAL, "wavefronts above 2^30 rays are not supported (split the batch)");
    Rays kr; State ks;
    if (int rc = to_rays(rays, kr)) return rc;
    if (int rc = to_state(state, ks)) return rc;
    if (n == 0) return SCN_OK;
    if (use_march(m)) { Samples none{}; return launch_march<true>(m, kr, ks, none, T, n_evals, n, s); }
    k_transmittance<<<grid_for(n), BLOCK, 0, s>>>(m->P, kr, ks, T, n_evals, n);
    return check_launch("k_transmittance");
}
int scn_transmittance_batch(const scn_medium *m, const scn_rays *rays, const scn_state *state, float *T_out, size_t n, void *stream) {
    if (!m || !rays || !state) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    if (!T_out) return set_error(SCN_EINVAL, "null argument");
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    return transmittance_impl(m, rays, state, T_out, nullptr, n, (cudaStream_t)stream);
}

static int nee_impl(const scn_medium *m, const scn_nee_in *in, float *pdf, float *gx, float *gy, float *gz, size_t n, void *stream) {
    if (!m || !in) return set_error(SCN_EINVAL, "null argument");
    if (!in->dx || !in->dy || !in->dz || !in->nx || !in->ny || !in->nz || !in->px || !in->py || !in->pz) return set_error(SCN_EINVAL, "nee: direction, normal and position are required");
    if (n == 0) return SCN_OK;
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    NeeIn k; k.dx = in->dx; k.dy = in->dy; k.dz = in->dz; k.nx = in->nx; k.ny = in->ny; k.nz = in->nz; k.px = in->px; k.py = in->py; k.pz = in->pz;
    k.ts = in->t_segment; k.info_t = in->info_t; k.ix = in->pix_x; k.iy = in->pix_y; k.is = in->pix_spp; k.ig = in->pix_seg; k.scene_seed = in->scene_seed;
    bool all = true; for (int i = 0; i < 7; ++i) all = all && in->ctx[i];
    for (int i = 0; i < 7; ++i) k.ctx[i] = all ? in->ctx[i] : nullptr;
    k_nee<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(m->P, k, pdf, gx, gy, gz, n);
    return check_launch("k_nee");
}
int scn_nee_pdf_batch(const scn_medium *m, const scn_nee_in *in, float *pdf_out, size_t n, void *stream) {
    if (!pdf_out) return set_error(SCN_EINVAL, "null output");
    return nee_impl(m, in, pdf_out, nullptr, nullptr, nullptr, n, stream);
}
int scn_nee_grad_batch(const scn_medium *m, const scn_nee_in *in, float *gx, float *gy, float *gz, size_t n, void *stream) {
    if (!gx || !gy || !gz) return set_error(SCN_EINVAL, "null output");
    return nee_impl(m, in, nullptr, gx, gy, gz, n, stream);
}

// debug build only (-DSCN_DEBUG_BOUNDS, tools/debug_bounds.py; not declared in include/scn.h): the eight assertion counters of the tuned
// kernels (scn_device.cuh: SCN_CHECK), read and cleared.  The product build answers SCN_EUNSUPPORTED.
int scn_debug_bounds_failures(unsigned long long out[8]) {
#if defined(SCN_DEBUG_BOUNDS)
    // the counters are per translation unit (static __device__): this unit's, then the three units that hold the tuned kernels
    cudaError_t e = cudaMemcpyFromSymbol(out, g_scn_bounds_fail, 8 * sizeof(unsigned long long));
    if (e != cudaSuccess) return cuda_fail(e, "scn_debug_bounds_failures");
    unsigned long long z[8] = {0};
    e = cudaMemcpyToSymbol(g_scn_bounds_fail, z, sizeof z);
    if (e == cudaSuccess) e = bounds_collect_march_sample(out);
    if (e == cudaSuccess) e = bounds_collect_march_transmit(out);
    if (e == cudaSuccess) e = bounds_collect_march1d(out);
    return e == cudaSuccess ? SCN_OK : cuda_fail(e, "scn_debug_bounds_failures");
#else
    (void)out; return set_error(SCN_EUNSUPPORTED, "built without SCN_DEBUG_BOUNDS");
#endif
}

// test hook: serve this medium with the one-thread-per-ray generic kernels even where a tuned kernel covers it
int scn_debug_force_generic(scn_medium *m, int on) {
    if (!m) return set_error(SCN_EINVAL, "null argument");
    m->force_generic = on != 0;
    return SCN_OK;
}

int scn_medium_info(const scn_medium *m, scn_medium_info_t *out) {
    if (!m || !out) return set_error(SCN_EINVAL, "null argument");
    memset(out, 0, sizeof(*out));
    const bool tuned = use_march(m);
    out->query_kernel = !tuned ? SCN_KERNEL_GENERIC : (m->march1d ? SCN_KERNEL_MARCH1D : SCN_KERNEL_MARCH3D);
    out->eval_kernel = fast_noise3d_supported(m->P) ? SCN_KERNEL_FAST_NOISE3D : SCN_KERNEL_GENERIC;
    out->march_block = m->march1d ? MARCH1D_BLOCK : m->march_block;
    out->cache_entries = m->march_block ? march_cache_entries(m->P.n_imp, march_rayz(m->P)) : 0;
    out->smem_bytes = m->march_block ? (uint32_t)march_smem_total(m->P, out->cache_entries, m->march_block) : 0;
    out->device = m->device; out->sm_count = m->sm_count;
    out->kernel_radius = m->P.iso3d || m->P.sampling_1d ? m->P.R_iso : m->P.R_world;
    return SCN_OK;
}

int scn_debug_hash4(const uint32_t *a, const uint32_t *b, const uint32_t *c, const uint32_t *d, uint32_t *out, size_t n, void *stream) {
    if (!a || !b || !c || !d || !out) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    k_debug_hash4<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(a, b, c, d, out, n);
    return check_launch("k_debug_hash4");
}
int scn_debug_pcg_stream(const uint32_t *seed32, float *out, size_t n, int draws, void *stream) {
    if (!seed32 || !out || draws < 0) return set_error(SCN_EINVAL, "bad argument");
    if (n == 0) return SCN_OK;
    k_debug_pcg<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(seed32, out, n, draws);
    return check_launch("k_debug_pcg");
}
int scn_debug_cell_impulses(const uint32_t *ci, const uint32_t *cj, const uint32_t *ck, uint32_t seed, uint32_t count, float *out, size_t n, void *stream) {
    if (!ci || !cj || !ck || !out) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    k_debug_cell<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(ci, cj, ck, seed, count, out, n);
    return check_launch("k_debug_cell");
}

int scn_debug_accepted_counts(const scn_medium *m, const float *px, const float *py, const float *pz, const float *dx, const float *dy, const float *dz,
                              uint32_t seed, uint32_t *out, size_t n, void *stream) {
    if (!m || !px || !py || !pz || !dx || !dy || !dz || !out) return set_error(SCN_EINVAL, "null argument");
    if (!m->P.stationary || m->P.sampling_1d) return set_error(SCN_EUNSUPPORTED, "accepted counts: stationary 3D media only");
    if (n == 0) return SCN_OK;
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    k_debug_accept<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(m->P, px, py, pz, dx, dy, dz, seed, out, n);
    return check_launch("k_debug_accept");
}

// ---------------------------------------------------------------- host-buffer entry points (the thin Tungsten shim calls these)
// Chunks of the batch travel host -> device -> kernel -> host through a three-slot arena (HostPipe below): the H2D copy of chunk
// k+1, the kernel of chunk k and the D2H copy of chunk k-1 overlap.  The arena is grown once and reused.
// The caller's buffers should be page-locked (cudaHostAlloc / cudaHostRegister): with pageable memory every cudaMemcpyAsync is
// staged by the driver and blocks the calling thread, so the three stages no longer overlap (results are identical, throughput is
// what bench.py reports as e2e.pageable).
static int ensure_arena(scn_medium *m, size_t bytes_per_slot) {
    for (int i = 0; i < 3; ++i) if (!m->streams[i]) SCN_CUDA(cudaStreamCreateWithFlags(&m->streams[i], cudaStreamNonBlocking));
    for (int i = 0; i < 3; ++i) if (!m->ev_out[i]) {
        SCN_CUDA(cudaEventCreateWithFlags(&m->ev_in[i], cudaEventDisableTiming)); SCN_CUDA(cudaEventCreateWithFlags(&m->ev_k[i], cudaEventDisableTiming));
        SCN_CUDA(cudaEventCreateWithFlags(&m->ev_out[i], cudaEventDisableTiming));
    }
    if (bytes_per_slot <= m->arena_cap) return SCN_OK;
    for (int i = 0; i < 3; ++i) { if (m->arena[i]) { SCN_CUDA(cudaFree(m->arena[i])); m->arena[i] = nullptr; } }
    m->arena_cap = 0;
    for (int i = 0; i < 3; ++i) SCN_CUDA(cudaMalloc((void **)&m->arena[i], bytes_per_slot));
    m->arena_cap = bytes_per_slot;
    return SCN_OK;
}
// Queries per pipeline stage of the *_host entry points: at most 2^22, at least 2^16, and small enough that a batch makes eight
// stages, so that mid-sized batches overlap their copies too.
static size_t host_chunk(size_t n) {
    const size_t cap = (size_t)1 << 22;
    size_t c = ((n + 7) / 8 + 255) & ~(size_t)255;
    if (c < ((size_t)1 << 16)) c = (size_t)1 << 16;
    if (c > cap) c = cap;
    return c < n ? c : n;
}

} // extern "C"
struct Carver { // carves aligned sub-buffers out of one arena slot
    char *base; size_t off = 0;
    template <typename T> T *take(size_t count) { T *p = (T *)(base + off); off += (count * sizeof(T) + 255) & ~(size_t)255; return p; }
};
template <typename T> static cudaError_t h2d(T *dst, const T *src, size_t first, size_t count, cudaStream_t s) { return cudaMemcpyAsync(dst, src + first, count * sizeof(T), cudaMemcpyHostToDevice, s); }
template <typename T> static cudaError_t d2h(T *dst, size_t first, const T *src, size_t count, cudaStream_t s) { return cudaMemcpyAsync(dst + first, src, count * sizeof(T), cudaMemcpyDeviceToHost, s); }
extern "C" {

// One stream per stage: every H2D copy on `up`, every kernel on `run`, every D2H copy on `down`, chained by events per arena slot
// (three slots), so that each copy engine sees its transfers back to back in batch order and the two directions stay busy at once:
// H2D of chunk k+1 || kernel of chunk k || D2H of chunk k-1.  `body(b, c, slot, up, run, down)` enqueues one chunk in three parts
// through the three callbacks below.  On any failure every stream is drained before returning, so that no copy into or out of the
// caller's buffers (or the arena) is still in flight when the caller frees them or retries.
struct HostPipe {
    scn_medium *m; cudaStream_t up, run, down; size_t k = 0; int slot = 0;
    explicit HostPipe(scn_medium *m_) : m(m_), up(m_->streams[0]), run(m_->streams[1]), down(m_->streams[2]) {}
    int begin_upload() { if (k >= 3) SCN_CUDA(cudaStreamWaitEvent(up, m->ev_k[slot], 0)); return SCN_OK; } // the kernel of chunk k-3 has read this slot's inputs
    int begin_kernel() {
        SCN_CUDA(cudaEventRecord(m->ev_in[slot], up)); SCN_CUDA(cudaStreamWaitEvent(run, m->ev_in[slot], 0));
        if (k >= 3) SCN_CUDA(cudaStreamWaitEvent(run, m->ev_out[slot], 0)); // chunk k-3's results have left this slot
        return SCN_OK;
    }
    int begin_download() { SCN_CUDA(cudaEventRecord(m->ev_k[slot], run)); SCN_CUDA(cudaStreamWaitEvent(down, m->ev_k[slot], 0)); return SCN_OK; }
    int end_chunk() { SCN_CUDA(cudaEventRecord(m->ev_out[slot], down)); slot = (slot + 1) % 3; ++k; return SCN_OK; }
    int drain(int rc) { // rc of the loop; secondary errors while draining are ignored
        cudaError_t first = cudaSuccess;
        for (int i = 0; i < 3; ++i) { cudaError_t e = cudaStreamSynchronize(m->streams[i]); if (first == cudaSuccess) first = e; }
        if (rc != SCN_OK) { cudaGetLastError(); return rc; }
        return first == cudaSuccess ? SCN_OK : cuda_fail(first, "host pipeline");
    }
};
#define SCN_PIPE(x) do { if (int rc_ = (x)) return rc_; } while (0)

static int eval_host_chunks(scn_medium *m, const scn_eval_in *in, const scn_eval_out *out, size_t n, int what, size_t chunk, HostPipe &hp) {
    bool has_ctx = true; for (int i = 0; i < 7; ++i) has_ctx = has_ctx && in->ctx[i];
    for (size_t b = 0; b < n; b += chunk) {
        size_t c = n - b < chunk ? n - b : chunk;
        Carver cv{m->arena[hp.slot]};
        scn_eval_in di = *in; scn_eval_out dd = *out;
        float *f;
        SCN_PIPE(hp.begin_upload());
#define IN_F(field) if (in->field) { f = cv.take<float>(c); SCN_CUDA(h2d(f, in->field, b, c, hp.up)); di.field = f; }
#define IN_U(field) if (in->field) { uint32_t *q = cv.take<uint32_t>(c); SCN_CUDA(h2d(q, in->field, b, c, hp.up)); di.field = q; }
        IN_F(px) IN_F(py) IN_F(pz) IN_F(dx) IN_F(dy) IN_F(dz) IN_F(t) IN_F(info_t) IN_U(pix_x) IN_U(pix_y) IN_U(pix_spp) IN_U(pix_seg)
        for (int i = 0; i < 7; ++i) { di.ctx[i] = nullptr; if (has_ctx) { f = cv.take<float>(c); SCN_CUDA(h2d(f, in->ctx[i], b, c, hp.up)); di.ctx[i] = f; } }
#undef IN_F
#undef IN_U
        if (out->val) dd.val = cv.take<float>(c);
        if (out->gx) dd.gx = cv.take<float>(c);
        if (out->gy) dd.gy = cv.take<float>(c);
        if (out->gz) dd.gz = cv.take<float>(c);
        if (out->gp_id) dd.gp_id = cv.take<int32_t>(c);
        SCN_PIPE(hp.begin_kernel());
        SCN_PIPE(scn_eval_batch(m, &di, &dd, c, what, hp.run));
        SCN_PIPE(hp.begin_download());
        if (out->val) SCN_CUDA(d2h(out->val, b, dd.val, c, hp.down));
        if (out->gx) SCN_CUDA(d2h(out->gx, b, dd.gx, c, hp.down));
        if (out->gy) SCN_CUDA(d2h(out->gy, b, dd.gy, c, hp.down));
        if (out->gz) SCN_CUDA(d2h(out->gz, b, dd.gz, c, hp.down));
        if (out->gp_id) SCN_CUDA(d2h(out->gp_id, b, dd.gp_id, c, hp.down));
        SCN_PIPE(hp.end_chunk());
    }
    return SCN_OK;
}

int scn_eval_batch_host(scn_medium *m, const scn_eval_in *in, const scn_eval_out *out, size_t n, int what) {
    if (!m || !in || !out) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    if (!in->px || !in->py || !in->pz || !in->dx || !in->dy || !in->dz) return set_error(SCN_EINVAL, "eval: points and directions are required");
    std::lock_guard<std::mutex> lock(m->mu);
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    size_t chunk = host_chunk(n);
    size_t per = ((chunk * 4 + 255) & ~(size_t)255) * 32; // <= 32 arrays of 4-byte elements
    if (int rc = ensure_arena(m, per)) return rc;
    HostPipe hp(m);
    return hp.drain(eval_host_chunks(m, in, out, n, what, chunk, hp));
}

static int query_host_chunks(scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, float *T_out, size_t n, size_t chunk, bool has_ctx, HostPipe &hp) {
    for (size_t b = 0; b < n; b += chunk) {
        size_t c = n - b < chunk ? n - b : chunk;
        Carver cv{m->arena[hp.slot]};
        scn_rays dr = *rays; scn_state ds = *state;
        float *f; uint32_t *q; uint8_t *u8; int32_t *i32;
        cudaStream_t s = hp.up;
        SCN_PIPE(hp.begin_upload());
#define RAY_F(field) f = cv.take<float>(c); SCN_CUDA(h2d(f, rays->field, b, c, s)); dr.field = f;
#define RAY_U(field) if (rays->field) { q = cv.take<uint32_t>(c); SCN_CUDA(h2d(q, rays->field, b, c, s)); dr.field = q; }
        RAY_F(ox) RAY_F(oy) RAY_F(oz) RAY_F(dx) RAY_F(dy) RAY_F(dz) RAY_F(tmin) RAY_F(tmax) RAY_F(u) RAY_U(px) RAY_U(py) RAY_U(spp) RAY_U(segment)
#undef RAY_F
#undef RAY_U
        u8 = cv.take<uint8_t>(c); SCN_CUDA(h2d(u8, (const uint8_t *)state->first_scatter, b, c, s)); ds.first_scatter = u8;
#define ST_F(field) f = cv.take<float>(c); SCN_CUDA(h2d(f, (const float *)state->field, b, c, s)); ds.field = f;
        ST_F(last_val) ST_F(last_gx) ST_F(last_gy) ST_F(last_gz) ST_F(info_t)
#undef ST_F
        i32 = cv.take<int32_t>(c); SCN_CUDA(h2d(i32, (const int32_t *)state->bounce, b, c, s)); ds.bounce = i32;
        i32 = cv.take<int32_t>(c); SCN_CUDA(h2d(i32, (const int32_t *)state->last_gp_id, b, c, s)); ds.last_gp_id = i32;
        ds.scheme = state->scheme ? cv.take<uint8_t>(c) : nullptr;
        for (int i = 0; i < 7; ++i) ds.ctx[i] = has_ctx ? cv.take<float>(c) : nullptr;
        if (out) {
            scn_samples dd = *out;
            dd.ok = cv.take<uint8_t>(c); dd.exited = cv.take<uint8_t>(c); dd.t = cv.take<float>(c); dd.continued_t = cv.take<float>(c);
            dd.gx = cv.take<float>(c); dd.gy = cv.take<float>(c); dd.gz = cv.take<float>(c); dd.gp_id = cv.take<int32_t>(c); dd.scheme = cv.take<uint8_t>(c);
            dd.weight = cv.take<float>(c); dd.n_evals = out->n_evals ? cv.take<uint32_t>(c) : nullptr;
            SCN_PIPE(hp.begin_kernel());
            SCN_PIPE(scn_sample_distance_batch(m, &dr, &ds, &dd, c, hp.run));
            SCN_PIPE(hp.begin_download());
            s = hp.down;
            SCN_CUDA(d2h(out->ok, b, dd.ok, c, s)); SCN_CUDA(d2h(out->exited, b, dd.exited, c, s)); SCN_CUDA(d2h(out->t, b, dd.t, c, s));
            SCN_CUDA(d2h(out->continued_t, b, dd.continued_t, c, s)); SCN_CUDA(d2h(out->gx, b, dd.gx, c, s)); SCN_CUDA(d2h(out->gy, b, dd.gy, c, s));
            SCN_CUDA(d2h(out->gz, b, dd.gz, c, s)); SCN_CUDA(d2h(out->gp_id, b, dd.gp_id, c, s)); SCN_CUDA(d2h(out->scheme, b, dd.scheme, c, s));
            SCN_CUDA(d2h(out->weight, b, dd.weight, c, s)); if (out->n_evals) SCN_CUDA(d2h(out->n_evals, b, dd.n_evals, c, s));
            // the state is read-modify-write
            SCN_CUDA(d2h(state->first_scatter, b, ds.first_scatter, c, s)); SCN_CUDA(d2h(state->last_val, b, ds.last_val, c, s));
            SCN_CUDA(d2h(state->last_gx, b, ds.last_gx, c, s)); SCN_CUDA(d2h(state->last_gy, b, ds.last_gy, c, s)); SCN_CUDA(d2h(state->last_gz, b, ds.last_gz, c, s));
            SCN_CUDA(d2h(state->bounce, b, ds.bounce, c, s)); SCN_CUDA(d2h(state->last_gp_id, b, ds.last_gp_id, c, s)); SCN_CUDA(d2h(state->info_t, b, ds.info_t, c, s));
            if (state->scheme) SCN_CUDA(d2h(state->scheme, b, ds.scheme, c, s));
            if (has_ctx) for (int i = 0; i < 7; ++i) SCN_CUDA(d2h(state->ctx[i], b, ds.ctx[i], c, s));
        } else {
            float *T = cv.take<float>(c);
            SCN_PIPE(hp.begin_kernel());
            SCN_PIPE(transmittance_impl(m, &dr, &ds, T, nullptr, c, hp.run));
            SCN_PIPE(hp.begin_download());
            SCN_CUDA(d2h(T_out, b, T, c, hp.down));
        }
        SCN_PIPE(hp.end_chunk());
    }
    return SCN_OK;
}

static int query_host(scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, float *T_out, size_t n) {
    if (!m || !rays || !state) return set_error(SCN_EINVAL, "null argument");
    if (n == 0) return SCN_OK;
    Rays kr; State ks;
    if (int rc = to_rays(rays, kr)) return rc;
    if (int rc = to_state(state, ks)) return rc;
    if (out && (!out->ok || !out->exited || !out->t || !out->continued_t || !out->gx || !out->gy || !out->gz || !out->gp_id || !out->scheme || !out->weight))
        return set_error(SCN_EINVAL, "samples: every output but n_evals is required");
    std::lock_guard<std::mutex> lock(m->mu);
    DeviceGuard g(m->device); if (!g.ok) return set_error(SCN_ECUDA, "cannot select device %d", m->device);
    size_t chunk = host_chunk(n);
    size_t per = ((chunk * 4 + 255) & ~(size_t)255) * 48;
    if (int rc = ensure_arena(m, per)) return rc;
    HostPipe hp(m);
    return hp.drain(query_host_chunks(m, rays, state, out, T_out, n, chunk, ks.ctx[0] != nullptr, hp));
}
int scn_sample_distance_batch_host(scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, size_t n) {
    if (!out) return set_error(SCN_EINVAL, "null argument");
    return query_host(m, rays, state, out, nullptr, n);
}
int scn_transmittance_batch_host(scn_medium *m, const scn_rays *rays, const scn_state *state, float *T_out, size_t n) {
    if (!T_out) return set_error(SCN_EINVAL, "null argument");
    return query_host(m, rays, state, nullptr, T_out, n);
}

} // extern "C"
