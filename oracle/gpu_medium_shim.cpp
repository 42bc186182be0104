// oracle/gpu_medium_shim.cpp — the drop-in, compiled INTO the unmodified reference (TEST INFRASTRUCTURE; the product is
// libscn_b200.so, which this file only dlopen()s and calls through include/scn.h).
//
// GpuSparseConvolutionNoiseMedium is the class INTEGRATION.md §2 tells a Tungsten maintainer to add: a Medium registered under
// the factory string "sparse_conv_noise_gpu" (media/MediumFactory.cpp:13-22) whose sampleDistance / transmittance
// (media/Medium.hpp:104-111) forward to scn_sample_distance_batch_host / scn_transmittance_batch_host.  The reference's own
// PathTracer (integrators/path_tracer/PathTracer.cpp:68, integrators/TraceBase.cpp:118-122) then renders through the B200
// library without knowing it: a scene switches by changing "type": "sparse_conv_noise" to "type": "sparse_conv_noise_gpu".
// The scalar virtuals are wavefronts of n = 1; a maintainer who batches per bounce calls the same entry points with n = live paths.
//
// It derives from SparseConvolutionNoiseMedium only to inherit the JSON parsing of the host-side members the C ABI leaves
// on the host (phase functions, sigma_a / sigma_s, colour / emission lookups of the GP: include/scn.h "weight");
// intersectGP / sampleGradient of the base class are never reached — sampleDistance and transmittance are overridden whole.
//
// NEE side door: ConductorBsdf / MirrorBsdf call ((GPContextSparseConvNoise*)sample.ctxt)->noise.neePDF / neeGrad
// NON-virtually (bsdfs/ConductorBsdf.cpp:72-73,96-97,106-107), so an unmodified reference can only be handed a real
// SparseConvolutionNoiseRealization: the shim builds one with the reference's constructor arguments (SCNM.cpp:104) and
// loads the seven conditioning coefficients the GPU query wrote to scn_state.ctx[0..6] into it.  (A maintainer who edits
// the BSDFs calls scn_nee_pdf_batch / scn_nee_grad_batch with those seven floats instead.)
//
// Compiled with -fno-access-control like ref_harness.cpp (reads JsonPtr::_value and the medium's private members).

#include "media/SparseConvolutionNoiseMedium.hpp"
#include "media/MediumFactory.hpp"
#include "math/GaussianProcess.hpp"
#include "math/Ray.hpp"
#include "sampling/PathSampleGenerator.hpp"
#include "samplerecords/MediumSample.hpp"
#include "io/JsonPtr.hpp"

#include <rapidjson/stringbuffer.h>
#include <rapidjson/writer.h>

#include <dlfcn.h>

#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

#include "scn.h" // include/scn.h (-I../include)

using namespace Tungsten;

namespace {

struct ScnApi {
    void *handle = nullptr;
    int device = 0;
    decltype(&scn_last_error) last_error = nullptr;
    decltype(&scn_desc_from_json_at) desc_from_json_at = nullptr;
    decltype(&scn_desc_release) desc_release = nullptr;
    decltype(&scn_medium_create) medium_create = nullptr;
    decltype(&scn_medium_destroy) medium_destroy = nullptr;
    decltype(&scn_sample_distance_batch_host) sample_distance_host = nullptr;
    decltype(&scn_transmittance_batch_host) transmittance_host = nullptr;
};
ScnApi g_api;
std::mutex g_api_mu;
thread_local std::string g_shim_err;
std::atomic<unsigned long long> g_n_sample{0}, g_n_transmit{0}, g_n_media{0};
std::atomic<unsigned> g_thread_counter{0};

constexpr int HANDLES = 8; // the *_host entry points serialise per scn_medium: pool threads spread over a few replicas

} // namespace

namespace Tungsten {

class GpuSparseConvolutionNoiseMedium : public SparseConvolutionNoiseMedium
{
    std::string _json, _dir;
    scn_medium *_h[HANDLES] = {};

    scn_medium *handle() const
    {
        thread_local unsigned slot = g_thread_counter.fetch_add(1);
        return _h[slot % HANDLES];
    }

    SparseConvolutionNoiseRealization makeRealization(const float *c) const
    {
        // same argument list as media/SparseConvolutionNoiseMedium.cpp:104
        SparseConvolutionNoiseRealization r(_gp->get_gaussian_process(), _ctxt, _globalSeed, _impulseDensity, _useSingleRealization,
                                            _isotropicSpace3DSampling, _1DSampling, _1DsamplingScheme, _correlationXY, _surfVolPhaseSeparate,
                                            _surfVolPhaseAmpThresh);
        r.coeff_3D.value_scale = r.coeff_1D.value_scale = c[0];
        r.coeff_3D.gradient_scale = r.coeff_1D.gradient_scale = Vec3f(c[1], c[2], c[3]);
        r.coeff_3D.ray_origin = r.coeff_1D.ray_origin = Vec3f(c[4], c[5], c[6]);
        return r;
    }

    void release()
    {
        for (int i = 0; i < HANDLES; ++i)
            if (_h[i]) { g_api.medium_destroy(_h[i]); _h[i] = nullptr; }
    }

public:
    ~GpuSparseConvolutionNoiseMedium() { release(); }

    void fromJson(JsonPtr value, const Scene &scene) override
    {
        SparseConvolutionNoiseMedium::fromJson(value, scene);
        rapidjson::StringBuffer sb;
        rapidjson::Writer<rapidjson::StringBuffer> w(sb);
        value._value->Accept(w); // the medium's JSON object, handed to scn_desc_from_json_at unchanged
        _json = sb.GetString();
        _dir = scene.path().parent().absolute().asString(); // relative resources ("file" of a regular_grid field) resolve like Scene::fetchResource
    }

    void prepareForRender() override
    {
        SparseConvolutionNoiseMedium::prepareForRender();
        if (!g_api.handle) FAIL("sparse_conv_noise_gpu: libscn_b200.so is not loaded (ref_gpu_medium_register)");
        release();
        scn_medium_desc d;
        if (g_api.desc_from_json_at(_json.c_str(), _dir.empty() ? nullptr : _dir.c_str(), &d) != SCN_OK) FAIL("sparse_conv_noise_gpu: %s", g_api.last_error());
        for (int i = 0; i < HANDLES; ++i)
            if (g_api.medium_create(&d, g_api.device, &_h[i]) != SCN_OK) { g_api.desc_release(&d); FAIL("sparse_conv_noise_gpu: %s", g_api.last_error()); }
        g_api.desc_release(&d); // grids read from disk: every medium holds its own device copy
        g_n_media++;
    }

    // media/GaussianProcessMedium.cpp:221-341 with the march, the gradient and the state update done by the library
    bool sampleDistance(PathSampleGenerator &sampler, const Ray &ray, MediumState &state, MediumSample &sample) const override
    {
        sample.emission = Vec3f(0.0f);
        float nearT = ray.nearT(), farT = ray.farT();
        float maxT = std::isfinite(farT) ? farT : float(double(nearT) + 2000);
        if (state.bounce >= _maxBounce) return false;                 // GPM:235-237
        const bool trivial = maxT == 0.f;                             // GPM:239-248
        if (!trivial && _absorptionOnly && maxT == Ray::infinity()) return false; // GPM:251-252
        // the one draw of the march (SCNM:129); the trivial exit never reaches intersectGP
        float u = trivial ? 0.f : sampler.next1D();

        float o[3] = {ray.pos().x(), ray.pos().y(), ray.pos().z()}, d[3] = {ray.dir().x(), ray.dir().y(), ray.dir().z()};
        uint32_t pix[4] = {state.info.pixelSampleSegment.x(), state.info.pixelSampleSegment.y(), state.info.pixelSampleSegment.z(),
                           state.info.pixelSampleSegment.w()};
        scn_rays r;
        r.ox = &o[0]; r.oy = &o[1]; r.oz = &o[2]; r.dx = &d[0]; r.dy = &d[1]; r.dz = &d[2]; r.tmin = &nearT; r.tmax = &farT; r.u = &u;
        r.px = &pix[0]; r.py = &pix[1]; r.spp = &pix[2]; r.segment = &pix[3]; r.scene_seed = state.info.sceneSeed;
        uint8_t first = state.firstScatter, st_scheme = (uint8_t)(int)state.sparseConv1DSamplingScheme;
        float lastVal = state.lastVal, lg[3] = {float(state.lastAniso.x()), float(state.lastAniso.y()), float(state.lastAniso.z())};
        int32_t bounce = state.bounce, lastId = state.lastGPId;
        float infoT = state.info.t, ctx[7] = {0, 0, 0, 0, 0, 0, 0};
        scn_state s;
        s.first_scatter = &first; s.last_val = &lastVal; s.last_gx = &lg[0]; s.last_gy = &lg[1]; s.last_gz = &lg[2];
        s.bounce = &bounce; s.last_gp_id = &lastId; s.info_t = &infoT; s.scheme = &st_scheme;
        for (int i = 0; i < 7; ++i) s.ctx[i] = &ctx[i];
        uint8_t ok = 0, exited = 0, scheme = 0; float t = 0, contT = 0, g[3] = {0, 0, 0}, weight = 0; int32_t gpId = 0;
        scn_samples out;
        out.ok = &ok; out.exited = &exited; out.t = &t; out.continued_t = &contT; out.gx = &g[0]; out.gy = &g[1]; out.gz = &g[2];
        out.gp_id = &gpId; out.scheme = &scheme; out.weight = &weight; out.n_evals = nullptr;
        if (g_api.sample_distance_host(handle(), &r, &s, &out, 1) != SCN_OK) FAIL("sparse_conv_noise_gpu: %s", g_api.last_error());
        g_n_sample++;

        // MediumState as sampleDistance leaves it (GPM:292-293, 327-329, 338; SCNM:161-168, 176-181)
        state.firstScatter = first != 0; state.lastVal = lastVal; state.lastAniso = Vec3d(lg[0], lg[1], lg[2]);
        state.bounce = bounce; state.lastGPId = lastId; state.info.t = infoT;
        state.sparseConv1DSamplingScheme = (SparseConv1DSamplingScheme)(int)st_scheme;
        if (!trivial) {
            auto ctxt = std::make_shared<GPContextSparseConvNoise>();
            ctxt->noise = makeRealization(ctx);
            state.gpContext = ctxt;
        }
        sample.aniso = Vec3d(g[0], g[1], g[2]);
        if (!ok) return false;

        sample.exited = exited != 0;
        sample.t = t; sample.continuedT = contT; sample.pdf = 1.0f;
        sample.sparseConv1DSamplingScheme = (SparseConv1DSamplingScheme)(int)scheme;
        if (trivial) {
            sample.weight = Vec3f(1.f);
            sample.p = ray.pos() + sample.t * ray.dir();
            sample.phase = _phaseFunction.get();
            return true;
        }
        if (_absorptionOnly) sample.weight = Vec3f(weight); // GPM:253-257: the transmittance of the whole ray
        else {
            // colour / emission lookups stay on the host (include/scn.h: scn_samples.weight is sigmaS / sigmaT only)
            auto ro = vec_conv<Vec3d>(ray.pos()); auto rd = vec_conv<Vec3d>(ray.dir()).normalized();
            if (!sample.exited) {
                sample.weight = sample.continuedWeight = vec_conv<Vec3f>(_gp->color(ro + rd * double(contT)));
                sample.emission = vec_conv<Vec3f>(_gp->emission(ro + rd * double(contT)));
            } else sample.weight = sample.continuedWeight = Vec3f(1.f);
            sample.weight *= sigmaS(ray.pos() + sample.t * ray.dir()) / sigmaT(ray.pos() + sample.t * ray.dir());
            sample.continuedWeight *= sigmaS(ray.pos() + sample.continuedT * ray.dir()) / sigmaT(ray.pos() + sample.continuedT * ray.dir());
        }
        sample.p = ray.pos() + sample.t * ray.dir();
        sample.phase = _phaseFunctions[state.lastGPId].get();
        sample.gpId = state.lastGPId;
        sample.ctxt = state.gpContext.get();
        sample.rayInfo = state.info;
        return true;
    }

    // media/GaussianProcessMedium.cpp:343-393.  The reference advances *state while it walks; every caller on this path hands in a
    // scratch copy (integrators/TraceBase.cpp:77-85,118-122), so the shim reads it and leaves it alone, like scn_transmittance_batch.
    Vec3f transmittance(PathSampleGenerator &sampler, const Ray &ray, bool, bool, MediumState *state) const override
    {
        float nearT = ray.nearT(), farT = ray.farT();
        float u = sampler.next1D(); // intersectGP always runs (SCNM:129)
        float o[3] = {ray.pos().x(), ray.pos().y(), ray.pos().z()}, d[3] = {ray.dir().x(), ray.dir().y(), ray.dir().z()};
        uint32_t pix[4] = {state->info.pixelSampleSegment.x(), state->info.pixelSampleSegment.y(), state->info.pixelSampleSegment.z(),
                           state->info.pixelSampleSegment.w()};
        scn_rays r;
        r.ox = &o[0]; r.oy = &o[1]; r.oz = &o[2]; r.dx = &d[0]; r.dy = &d[1]; r.dz = &d[2]; r.tmin = &nearT; r.tmax = &farT; r.u = &u;
        r.px = &pix[0]; r.py = &pix[1]; r.spp = &pix[2]; r.segment = &pix[3]; r.scene_seed = state->info.sceneSeed;
        uint8_t first = state->firstScatter, st_scheme = 0;
        float lastVal = state->lastVal, lg[3] = {float(state->lastAniso.x()), float(state->lastAniso.y()), float(state->lastAniso.z())};
        int32_t bounce = state->bounce, lastId = state->lastGPId;
        float infoT = state->info.t;
        scn_state s;
        s.first_scatter = &first; s.last_val = &lastVal; s.last_gx = &lg[0]; s.last_gy = &lg[1]; s.last_gz = &lg[2];
        s.bounce = &bounce; s.last_gp_id = &lastId; s.info_t = &infoT; s.scheme = &st_scheme;
        for (int i = 0; i < 7; ++i) s.ctx[i] = nullptr;
        float T = 0.f;
        if (g_api.transmittance_host(handle(), &r, &s, &T, 1) != SCN_OK) FAIL("sparse_conv_noise_gpu: %s", g_api.last_error());
        g_n_transmit++;
        return Vec3f(T);
    }

    float pdf(PathSampleGenerator &, const Ray &, bool, bool) const override { return 1.0f; }
};

} // namespace Tungsten

extern "C" {

const char *ref_gpu_medium_last_error() { return g_shim_err.c_str(); }

// dlopen(libscn_b200.so) and add {"sparse_conv_noise_gpu", make_shared<GpuSparseConvolutionNoiseMedium>} to MediumFactory —
// the one line a maintainer adds to media/MediumFactory.cpp:13-22, done at run time because the reference tree is read-only.
int ref_gpu_medium_register(const char *libscn_path, int device)
{
    std::lock_guard<std::mutex> lock(g_api_mu);
    if (!g_api.handle) {
        void *h = dlopen(libscn_path, RTLD_NOW | RTLD_LOCAL);
        if (!h) { g_shim_err = std::string("dlopen failed: ") + dlerror(); return -1; }
        ScnApi a; a.handle = h; a.device = device;
#define SCN_SYM(field, name) a.field = (decltype(a.field))dlsym(h, #name); if (!a.field) { g_shim_err = "missing symbol " #name; dlclose(h); return -1; }
        SCN_SYM(last_error, scn_last_error) SCN_SYM(desc_from_json_at, scn_desc_from_json_at) SCN_SYM(desc_release, scn_desc_release) SCN_SYM(medium_create, scn_medium_create)
        SCN_SYM(medium_destroy, scn_medium_destroy) SCN_SYM(sample_distance_host, scn_sample_distance_batch_host)
        SCN_SYM(transmittance_host, scn_transmittance_batch_host)
#undef SCN_SYM
        g_api = a;
    }
    g_api.device = device;
    static const char *NAME = "sparse_conv_noise_gpu";
    auto &entries = MediumFactory::entries();
    for (const auto &e : entries) if (std::strcmp(e.first, NAME) == 0) return 0;
    entries.push_back({NAME, [] { return std::shared_ptr<Medium>(std::make_shared<GpuSparseConvolutionNoiseMedium>()); }});
    return 0;
}

// queries served by the library since the last reset: [0] sampleDistance, [1] transmittance, [2] media prepared
void ref_gpu_medium_stats(unsigned long long *out, int reset)
{
    out[0] = g_n_sample; out[1] = g_n_transmit; out[2] = g_n_media;
    if (reset) { g_n_sample = 0; g_n_transmit = 0; g_n_media = 0; }
}

} // extern "C"
