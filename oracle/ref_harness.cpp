// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Thin extern "C" batch entry points over the UNMODIFIED reference classes, compiled together with
// the reference sources where they lie (see oracle/Makefile, target `ref`). It is what pins the CPU
// restatement (oracle/scn_oracle.c), generates tests/golden/*, and is the timed "reference" CPU
// baseline of bench.py. Nothing here restates reference arithmetic: every number comes out of
//   SparseConvolutionNoiseRealization (src/core/math/SparseConvolutionNoise.cpp),
//   SparseConvolutionNoiseMedium / GaussianProcessMedium (src/core/media/*.cpp),
//   MathUtil::xxhash32, UniformSampler, rand_normal_2 (src/core/math, src/core/sampling).
// The medium is always built through the reference's own JSON path (Scene::fetchMedium -> fromJson),
// which is mandatory: the programmatic constructors leave the covariance transforms uninitialised.
// Compiled with -fno-access-control so the harness can read the medium's private JSON parameters.

#include "io/Scene.hpp"
#include "io/JsonDocument.hpp"
#include "media/SparseConvolutionNoiseMedium.hpp"
#include "math/SparseConvolutionNoise.hpp"
#include "math/MathUtil.hpp"
#include "sampling/UniformSampler.hpp"
#include "sampling/PathSampleGenerator.hpp"
#include "sampling/Gaussian.hpp"
#include "math/Ray.hpp"
#include "renderer/TraceableScene.hpp"
#include "integrators/Integrator.hpp"
#include "cameras/Camera.hpp"
#include "thread/ThreadUtils.hpp"

#include <chrono>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <thread>
#include <vector>

using namespace Tungsten;

namespace {

thread_local std::string g_err;

struct RefMedium {
    std::unique_ptr<Scene> scene;
    std::unique_ptr<JsonDocument> doc;
    std::shared_ptr<Medium> medium;
    SparseConvolutionNoiseMedium *scn = nullptr;
};

// The one external random input of the path (SparseConvolutionNoiseMedium.cpp:129): the march jitter.
struct FixedSampler : public PathSampleGenerator {
    float u = 0.f;
    UniformSampler uni;
    void startPath(uint32, uint32) override {}
    void advancePath() override {}
    void saveState(OutputStreamHandle &) override {}
    void loadState(InputStreamHandle &) override {}
    bool nextBoolean(float pTrue) override { return u < pTrue; }
    int nextDiscrete(int) override { return 0; }
    float next1D() override { return u; }
    Vec2f next2D() override { return Vec2f(u, u); }
    UniformSampler &uniformGenerator() override { return uni; }
};

// H-uninit: with "useAnisoMtx": false the reference only ever writes the DIAGONALS of the
// default-constructed Eigen matrices _local_to_world / _world_to_local (GPFunctions.cpp:661-667); the off-diagonals
// are whatever the heap held.  On a fresh heap that is zero, which is what every reference render relies on;
// in a long-lived test process it is not (observed here: NaN gradients on the 3rd medium created).  The harness
// therefore pins the intended state: off-diagonals zero, transposes recomputed.  No arithmetic is restated.
void pinDiagonalTransforms(StationaryCovariance *cov) {
    auto *se = dynamic_cast<SquaredExponentialCovariance *>(cov);
    if (!se || se->_use_aniso_mtx) return;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            if (i != j) { se->_local_to_world(i, j) = 0.f; se->_world_to_local(i, j) = 0.f; }
    se->_local_to_world_transpose = se->_local_to_world.transpose();
    se->_world_to_local_transpose = se->_world_to_local.transpose();
}

SparseConvolutionNoiseRealization makeRealization(const RefMedium *m) {
    auto *s = m->scn;
    auto gp = s->_gp->get_gaussian_process();
    // Same argument list as SparseConvolutionNoiseMedium.cpp:104
    return SparseConvolutionNoiseRealization(gp, s->_ctxt, s->_globalSeed, s->_impulseDensity, s->_useSingleRealization,
                                             s->_isotropicSpace3DSampling, s->_1DSampling, s->_1DsamplingScheme, s->_correlationXY,
                                             s->_surfVolPhaseSeparate, s->_surfVolPhaseAmpThresh);
}

void setCoeffs(SparseConvolutionNoiseRealization &r, const float *c) {
    r.coeff_3D.value_scale = c[0];
    r.coeff_3D.gradient_scale = Vec3f(c[1], c[2], c[3]);
    r.coeff_3D.ray_origin = Vec3f(c[4], c[5], c[6]);
    r.coeff_1D.value_scale = c[0];
    r.coeff_1D.gradient_scale = Vec3f(c[1], c[2], c[3]);
    r.coeff_1D.ray_origin = Vec3f(c[4], c[5], c[6]);
}

void getCoeffs(const SparseConvolutionNoiseRealization &r, float *c) {
    if (r._1DSampling) {
        c[0] = r.coeff_1D.value_scale;
        c[1] = r.coeff_1D.gradient_scale.x(); c[2] = r.coeff_1D.gradient_scale.y(); c[3] = r.coeff_1D.gradient_scale.z();
        c[4] = r.coeff_1D.ray_origin.x(); c[5] = r.coeff_1D.ray_origin.y(); c[6] = r.coeff_1D.ray_origin.z();
    } else {
        c[0] = r.coeff_3D.value_scale;
        c[1] = r.coeff_3D.gradient_scale.x(); c[2] = r.coeff_3D.gradient_scale.y(); c[3] = r.coeff_3D.gradient_scale.z();
        c[4] = r.coeff_3D.ray_origin.x(); c[5] = r.coeff_3D.ray_origin.y(); c[6] = r.coeff_3D.ray_origin.z();
    }
}

RayInfo makeInfo(const uint32_t *pix, size_t i, uint32_t sceneSeed, const float *info_t) {
    RayInfo info;
    info.pixelSampleSegment = Vec4u(pix[4 * i + 0], pix[4 * i + 1], pix[4 * i + 2], pix[4 * i + 3]);
    info.sceneSeed = sceneSeed;
    info.t = info_t ? info_t[i] : 0.f;
    return info;
}

template <typename F>
void parallelFor(size_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f(size_t(0), n); return; }
    std::vector<std::thread> pool;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t b = std::min(n, chunk * t), e = std::min(n, chunk * (t + 1));
        if (b < e) pool.emplace_back([=] { f(b, e); });
    }
    for (auto &th : pool) th.join();
}

} // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }
int ref_hw_threads() { return (int)std::thread::hardware_concurrency(); }

void *ref_medium_create(const char *medium_json) {
    try {
        auto m = std::make_unique<RefMedium>();
        m->scene = std::make_unique<Scene>();
        m->doc = std::make_unique<JsonDocument>(Path("oracle-medium.json"), std::string(medium_json));
        m->medium = m->scene->fetchMedium(*m->doc);
        m->medium->loadResources();
        m->medium->prepareForRender();
        m->scn = dynamic_cast<SparseConvolutionNoiseMedium *>(m->medium.get());
        if (!m->scn) { g_err = "medium is not of type sparse_conv_noise"; return nullptr; }
        {
            auto gp = m->scn->_gp->get_gaussian_process();
            if (auto *st = dynamic_cast<StationaryCovariance *>(gp->_cov.get())) pinDiagonalTransforms(st);
            else if (auto *ns = dynamic_cast<NonstationaryCovariance *>(gp->_cov.get())) pinDiagonalTransforms(ns->_stationaryCov.get());
        }
        return m.release();
    } catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

void ref_medium_destroy(void *h) { delete (RefMedium *)h; }

// ---- H1-H3: hash / PRNG primitives (MathUtil.hpp:179-228, UniformSampler.hpp, Gaussian.cpp:21-33) ----
uint32_t ref_xxhash32_1(uint32_t a) { return MathUtil::xxhash32(a); }
uint32_t ref_xxhash32_2(uint32_t a, uint32_t b) { return MathUtil::xxhash32(Vec2u(a, b)); }
uint32_t ref_xxhash32_3(uint32_t a, uint32_t b, uint32_t c) { return MathUtil::xxhash32(Vec3u(a, b, c)); }
uint32_t ref_xxhash32_4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) { return MathUtil::xxhash32(Vec4u(a, b, c, d)); }

void ref_pcg_stream(uint64_t state, int n, float *out) {
    UniformSampler s;
    s.set_state(state);
    for (int i = 0; i < n; ++i) out[i] = s.next1D();
}

// set_state(state); next3D() -> out[0..2] = x,y,z ; out[3] = following next1D()
void ref_pcg_next3d(uint64_t state, float *out) {
    UniformSampler s;
    s.set_state(state);
    Vec3f v = s.next3D();
    out[0] = v.x(); out[1] = v.y(); out[2] = v.z();
    out[3] = s.next1D();
}

void ref_box_muller(uint64_t state, int npairs, double *out) {
    UniformSampler s;
    s.set_state(state);
    for (int i = 0; i < npairs; ++i) {
        Vec2d v = rand_normal_2(s);
        out[2 * i] = v.x(); out[2 * i + 1] = v.y();
    }
}

// ---- noise level: SparseConvolutionNoise.cpp:101-124 with an explicit seed, conditioning off ----
int ref_eval_noise3d(void *h, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        auto r = makeRealization(m);
        UniformSampler sampler;
        for (size_t i = b; i < e; ++i) {
            Vec4f v = r.evaluateNoise3D(Vec3f(p[3 * i], p[3 * i + 1], p[3 * i + 2]), Vec3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]), seed, sampler, false);
            out4[4 * i] = v.x(); out4[4 * i + 1] = v.y(); out4[4 * i + 2] = v.z(); out4[4 * i + 3] = v.w();
        }
    });
    return 0;
}

// ---- realization level: evaluateValue / evaluateGradient (SparseConvolutionNoise.cpp:73-99) ----
// coeffs: optional n x 7 conditioning coefficients (value_scale, gradient_scale xyz, ray_origin xyz)
int ref_eval(void *h, size_t n, const float *p, const float *d, const float *t, const uint32_t *pix, uint32_t sceneSeed,
             const float *info_t, const float *coeffs, float *out_val, float *out_grad, int *out_gpid, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        auto r = makeRealization(m);
        UniformSampler sampler;
        for (size_t i = b; i < e; ++i) {
            if (coeffs) setCoeffs(r, coeffs + 7 * i);
            RayInfo info = makeInfo(pix, i, sceneSeed, info_t);
            Vec3f P(p[3 * i], p[3 * i + 1], p[3 * i + 2]), D(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
            float ti = t ? t[i] : 0.f;
            int gpid = 0;
            if (out_val) { out_val[i] = r.evaluateValue(P, ti, D, info, sampler, gpid); if (out_gpid) out_gpid[i] = gpid; }
            if (out_grad) {
                Vec3f g = r.evaluateGradient(P, ti, D, info, sampler);
                out_grad[3 * i] = g.x(); out_grad[3 * i + 1] = g.y(); out_grad[3 * i + 2] = g.z();
            }
        }
    });
    return 0;
}

// ---- conditioning (SparseConvolutionNoise.cpp:431-595): returns the 7 coefficients per ray ----
int ref_conditioning(void *h, size_t n, const float *p, const float *d, const float *target_val, const float *target_grad,
                     const uint32_t *pix, uint32_t sceneSeed, const float *info_t, float *out_coeffs, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        UniformSampler sampler;
        for (size_t i = b; i < e; ++i) {
            auto r = makeRealization(m);
            RayInfo info = makeInfo(pix, i, sceneSeed, info_t);
            r.conditioning(Vec3f(p[3 * i], p[3 * i + 1], p[3 * i + 2]), Vec3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]), target_val[i],
                           Vec3f(target_grad[3 * i], target_grad[3 * i + 1], target_grad[3 * i + 2]), info, sampler);
            getCoeffs(r, out_coeffs + 7 * i);
        }
    });
    return 0;
}

// ---- medium level: Medium::sampleDistance (GaussianProcessMedium.cpp:221-341) ----
// rays: o[3n], d[3n], tmin[n], tmax[n] (inf allowed), u[n] (march jitter)
// state in/out: first_scatter[n] (u8), last_val[n], last_grad[3n] (f64), bounce[n], last_gp_id[n], info_t[n]
// out: ok[n], exited[n], t[n], cont_t[n], grad[3n] (f64, sample.aniso), gp_id[n], scheme[n], coeffs[7n], weight[3n], p_out[3n]
int ref_sample_distance(void *h, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                        const uint32_t *pix, uint32_t sceneSeed,
                        uint8_t *first_scatter, float *last_val, double *last_grad, int *bounce, int *last_gp_id, float *info_t,
                        uint8_t *ok, uint8_t *exited, float *t_out, float *cont_t, double *grad, int *gp_id, uint8_t *scheme,
                        float *coeffs, float *weight, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        FixedSampler sampler;
        for (size_t i = b; i < e; ++i) {
            Medium::MediumState st;
            st.reset();
            st.component = 0;
            st.firstScatter = first_scatter[i] != 0;
            st.lastVal = last_val[i];
            st.lastAniso = Vec3d(last_grad[3 * i], last_grad[3 * i + 1], last_grad[3 * i + 2]);
            st.bounce = bounce[i];
            st.lastGPId = last_gp_id[i];
            st.info = makeInfo(pix, i, sceneSeed, info_t);
            sampler.u = u[i];
            Ray ray(Vec3f(o[3 * i], o[3 * i + 1], o[3 * i + 2]), Vec3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin[i], tmax[i]);
            MediumSample s;
            std::memset(&s, 0, sizeof(s));
            s.aniso = Vec3d(0.);
            bool res = m->medium->sampleDistance(sampler, ray, st, s);
            ok[i] = res;
            exited[i] = s.exited;
            t_out[i] = s.t;
            cont_t[i] = s.continuedT;
            grad[3 * i] = s.aniso.x(); grad[3 * i + 1] = s.aniso.y(); grad[3 * i + 2] = s.aniso.z();
            gp_id[i] = res ? s.gpId : st.lastGPId;
            scheme[i] = (uint8_t)(int)s.sparseConv1DSamplingScheme;
            if (weight) { weight[3 * i] = s.weight.x(); weight[3 * i + 1] = s.weight.y(); weight[3 * i + 2] = s.weight.z(); }
            for (int k = 0; k < 7; ++k) coeffs[7 * i + k] = 0.f;
            if (st.gpContext) getCoeffs(((GPContextSparseConvNoise *)st.gpContext.get())->noise, coeffs + 7 * i);
            first_scatter[i] = st.firstScatter;
            last_val[i] = st.lastVal;
            last_grad[3 * i] = st.lastAniso.x(); last_grad[3 * i + 1] = st.lastAniso.y(); last_grad[3 * i + 2] = st.lastAniso.z();
            bounce[i] = st.bounce;
            last_gp_id[i] = st.lastGPId;
            info_t[i] = st.info.t;
        }
    });
    return 0;
}

// ---- Medium::transmittance (GaussianProcessMedium.cpp:343-393); state is read, a copy is advanced ----
int ref_transmittance(void *h, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                      const uint32_t *pix, uint32_t sceneSeed,
                      const uint8_t *first_scatter, const float *last_val, const double *last_grad, const int *bounce,
                      const int *last_gp_id, const float *info_t, float *T_out, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        FixedSampler sampler;
        for (size_t i = b; i < e; ++i) {
            Medium::MediumState st;
            st.reset();
            st.component = 0;
            st.firstScatter = first_scatter[i] != 0;
            st.lastVal = last_val[i];
            st.lastAniso = Vec3d(last_grad[3 * i], last_grad[3 * i + 1], last_grad[3 * i + 2]);
            st.bounce = bounce[i];
            st.lastGPId = last_gp_id[i];
            st.info = makeInfo(pix, i, sceneSeed, info_t);
            sampler.u = u[i];
            Ray ray(Vec3f(o[3 * i], o[3 * i + 1], o[3 * i + 2]), Vec3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin[i], tmax[i]);
            Vec3f T = m->medium->transmittance(sampler, ray, false, false, &st);
            T_out[i] = T.x();
        }
    });
    return 0;
}

// ---- 1D next-event estimation (SparseConvolutionNoise.cpp:601-743) on a realization carrying `coeffs` ----
int ref_nee(void *h, size_t n, const float *ray_dir, const float *normal, const float *p, const float *t_segment,
            const uint32_t *pix, uint32_t sceneSeed, const float *info_t, const float *coeffs,
            float *out_pdf, float *out_grad, int nthreads) {
    auto *m = (RefMedium *)h;
    parallelFor(n, nthreads, [=](size_t b, size_t e) {
        auto r = makeRealization(m);
        for (size_t i = b; i < e; ++i) {
            if (coeffs) setCoeffs(r, coeffs + 7 * i);
            RayInfo info = makeInfo(pix, i, sceneSeed, info_t);
            Vec3f D(ray_dir[3 * i], ray_dir[3 * i + 1], ray_dir[3 * i + 2]), N(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
            Vec3f P(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
            if (out_pdf) out_pdf[i] = r.neePDF(D, N, P, t_segment[i], info);
            if (out_grad) {
                Vec3f g = r.neeGrad(D, N, P, info);
                out_grad[3 * i] = g.x(); out_grad[3 * i + 1] = g.y(); out_grad[3 * i + 2] = g.z();
            }
        }
    });
    return 0;
}

// ---- the reference's own renderer (StandaloneRenderer::renderScene, src/tungsten/Shared.hpp:195-352, minus CLI and file output) ----
// Loads a scene file through Scene::load, renders it with the scene's integrator on `threads` pool threads and copies the linear
// framebuffer (Camera::getLinear, what Integrator::writeBuffers saves as the .pfm) into rgb (w*h*3 floats, row-major, top row first).
// The scene must set "renderer": {"scene_bvh": false}: TraceableScene then intersects its primitives in a plain loop
// (renderer/TraceableScene.hpp:186-193) and embree — which this recipe does not build — is never entered.
// Returns 0, or -1 with ref_last_error() set; *w and *h are written even when rgb is too small (cap < w*h*3).
static bool g_save_outputs = false;
// 1: ref_render also lets the reference write its result files (renderer.output_file / hdr_output_file, relative to the scene file)
void ref_render_save_outputs(int on) { g_save_outputs = on != 0; }
int ref_render(const char *scene_path, int threads, uint32_t seed, int *w, int *h, float *rgb, size_t cap, double *seconds) {
    try {
        static bool started = false;
        if (!started) { ThreadUtils::startThreads(threads > 0 ? threads : 1); started = true; }
        Path path(scene_path);
        Path dir = path.parent();
        std::unique_ptr<Scene> scene(Scene::load(path, nullptr, &dir));
        scene->loadResources();
        if (scene->rendererSettings().useSceneBvh()) { g_err = "ref_render needs \"scene_bvh\": false (embree is not built)"; return -1; }
        for (auto &m : scene->media())
            if (auto *scn = dynamic_cast<SparseConvolutionNoiseMedium *>(m.get())) {
                auto gp = scn->_gp->get_gaussian_process();
                if (auto *st = dynamic_cast<StationaryCovariance *>(gp->_cov.get())) pinDiagonalTransforms(st);
                else if (auto *ns = dynamic_cast<NonstationaryCovariance *>(gp->_cov.get())) pinDiagonalTransforms(ns->_stationaryCov.get());
            }
        std::unique_ptr<TraceableScene> flat(scene->makeTraceable(seed));
        Integrator &integrator = flat->integrator();
        auto t0 = std::chrono::steady_clock::now();
        while (!integrator.done()) {
            integrator.startRender([]() {});
            integrator.waitForCompletion();
        }
        if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (g_save_outputs) integrator.saveOutputs(); // Integrator::writeBuffers (integrators/Integrator.cpp:40-64): the scene's output_file / hdr_output_file
        Vec2u res = flat->cam().resolution();
        *w = (int)res.x(); *h = (int)res.y();
        if (!rgb || cap < (size_t)res.product() * 3) { g_err = "rgb buffer too small"; return -1; }
        for (uint32 y = 0; y < res.y(); ++y)
            for (uint32 x = 0; x < res.x(); ++x) {
                Vec3f c = flat->cam().getLinear(x, y);
                float *o = rgb + 3 * (x + (size_t)y * res.x());
                o[0] = c.x(); o[1] = c.y(); o[2] = c.z();
            }
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}

// Derived constants a port needs and that come out of libm on the host (multi-resolution tables etc.)
// scale = 2.5^k as computed by kernelScaleLevelRatio (float pow), and the seed offset the reference derives from it.
void ref_multires_level(int k, float *scale, int *additional_seed) {
    float base = 2.5f;
    float s = pow(base, (float)k);
    *scale = s;
    *additional_seed = (int)floor(log(s) / log(base));
}

// Exposes a few per-medium scalars so tests can assert the parser of the product sees the same medium.
int ref_medium_info(void *h, float *out, int nout) {
    auto *m = (RefMedium *)h;
    auto gp = m->scn->_gp->get_gaussian_process();
    float v[8] = {
        m->scn->_rayMarchStepSize, (float)m->scn->_rayMarchMinStep, (float)m->scn->_globalSeed, m->scn->_impulseDensity,
        gp->_cov->splattingKernelRadius(true, 1.0f), gp->_cov->splattingKernelRadius(false, 1.0f),
        gp->_cov->worldSamplingSpatialScale(), (float)(gp->_cov->isStationaryKernel() ? 1 : 0)};
    for (int i = 0; i < nout && i < 8; ++i) out[i] = v[i];
    return 0;
}

} // extern "C"
