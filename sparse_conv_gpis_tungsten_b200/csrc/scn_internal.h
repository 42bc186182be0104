// scn_internal.h — shared between the host side of the library (scn_json.cpp, scn_api.cu) and the kernels.
#ifndef SCN_INTERNAL_H_
#define SCN_INTERNAL_H_

#include <stddef.h>
#include <stdint.h>

#include "../../include/scn.h"

namespace scn {

int set_error(int code, const char *fmt, ...);
const char *last_error();
int desc_default(scn_medium_desc *d);
int desc_from_json(const char *text, const char *base_dir, scn_medium_desc *d);
void desc_release(scn_medium_desc *d);

// Multi-resolution levels k in [LVL_MIN, LVL_MIN + LVL_COUNT): scale = powf(2.5f, k) and the seed offset
// floor(logf(scale)/logf(2.5f)) come out of the HOST libm (the reference derives an integer seed offset and
// the grid scaling from them, math/SparseConvolutionNoise.cpp:60-70,252,300,328), never out of device libm.
constexpr int LVL_MIN = -20;
constexpr int LVL_COUNT = 41;

// A procedural field with the per-call constants of the reference hoisted to the host (math/GPFunctions.cpp:53-138).
struct FieldK {
    int32_t type, _pad;
    double val[3];
    double la[2], lb[2], scale[2], off[2]; // [0] first ramp, [1] second ramp of bottom_top_left_right
    // regular_grid (math/GPFunctions.hpp:241-330): res^3 x ncomp doubles — device memory inside the kernels, the caller's host array in
    // params_from_desc and in tests/hostsim; bounds and the diagonal as the reference's getValue reads them
    const double *grid; double gmin[3], gdiag[3]; int32_t gres, ginterp, gcomp, _pad2;
    double fmin, fmax; // sandstone / rust scalar fields: "min" / "max" as given (lerp end points)
};

// Everything a kernel needs, derived once from scn_medium_desc in scn_medium_create (device copy lives in
// constant-bank kernel parameters: the struct is passed by value as a __grid_constant__).
struct Params {
    // SparseConvolutionNoiseRealization flags (math/SparseConvolutionNoise.cpp:8-37)
    uint32_t n_imp;            // uint(impulseDensity)
    float density;             // impulseDensity as given (feeds the variance normalisation untruncated)
    uint32_t global_seed;
    int32_t single_realization, iso3d, sampling_1d, ctx, corr_xy, activate_conditioning, scheme;
    int32_t stationary, multires;
    int32_t surf_vol_sep; float surf_vol_thresh;
    // medium (media/SparseConvolutionNoiseMedium.cpp, media/GaussianProcessMedium.cpp:152-158)
    float step_size; uint32_t min_step; int32_t max_bounces; int32_t absorption_only; float sigma_s_over_t;
    // squared-exponential kernel (math/GPFunctions.cpp:654-679)
    float sigma, kernel_scale;
    float l_aniso[3], l_aniso_inv[3];
    float l_aniso_max, ls_max;
    // radii / normalisation that do not depend on the query point
    float R_iso;               // splattingKernelRadius(true, .)
    float R_world;             // splattingKernelRadius(false, 1) (stationary and brute-force non-stationary)
    float nf_iso;              // sqrt(variance3D(isIdentity=true))            (stationary / multires)
    float nf_world;            // sqrt(variance3D(isIdentity=false, gs=world)) (stationary)
    float nf_1d;               // sqrt(variance1D(localScale=1))               (stationary / multires)
    float world_scale;         // worldSamplingSpatialScale()
    int32_t add_seed_world;    // floor(log(world_scale)/log(2.5)) on the host
    // multi-resolution tables
    float lvl_scale[LVL_COUNT];
    int32_t lvl_add_seed[LVL_COUNT];
    float lvl_R_world[LVL_COUNT];
    float lvl_nf_world[LVL_COUNT];
    // GP
    int32_t gp_id, has_mean_add, id_add, _pad;
    scn_mean_desc mean, mean_add;
    // the knob SDFs (math/SdfFunctions.cpp:100-179): Mat4f::rotAxis((1,0,0), -90 / +90 / -45 degrees) as {a11, cos, sin} each and the two
    // "cone" constants Vec2f(PI / 3.) and Vec2f(PI / 3.f), all out of the host's libm / float arithmetic
    float knob_rot[12]; // [0..8] the three rotations, [9] = float(PI / 3.), [10] = PI / 3.f
    FieldK mean_f, mean_add_f;  // procedural means: the scalar field "f" behind mean / mean_add (a regular_grid here points at device memory)
    FieldK ls, var;
    FieldK aniso_f;            // proc_nonstationary "aniso": rotation-anisotropy field (math/GPFunctions.cpp:1678-1690); has_aniso = field present
    int32_t has_aniso;
    // "useAnisoMtx" (math/GPFunctions.cpp:668-675), row-major: mL = _local_to_world = l_conv * anisoMtx, mW = its inverse, mC = mW^T mW
    int32_t use_mtx;
    float mL[9], mW[9], mC[9], _pad3;
    double mC_dsi;             // 1 / sqrt(det mC) (GPF:749), host libm
    float log_base;            // logf(2.5f) from the host libm
    // fp32 copies for k_march's fast multi-resolution level estimate (level_ratio_fast, scn_march.cuh): the single-ramp "ls" field
    // as 0.5 log2(e) (la + r (lb - la)), r = clamp(c ls_f_scale + ls_f_off); 1 / log2(2.5)
    float ls_f_a, ls_f_b, ls_f_scale, ls_f_off, inv_log2_base;
    // alternative stationary splat kernels (math/GPFunctions.cpp:1020-1082, 1127-1214): world-space sampling only
    int32_t kernel;            // SCN_COV_SQUARED_EXPONENTIAL | SCN_COV_MATERN | SCN_COV_GABOR_ANISO | SCN_COV_GABOR_ISO
    float k_v, k_l;            // Matern: _v, _l
    float k_a, k_f;            // Gabor: _a = 1 / a_inv, _f = 1 / f_inv (float members assigned from double quotients)
    float k_omega[3];          // Gabor anisotropic: normalised _omega
};

int params_from_desc(const scn_medium_desc &d, Params &p);

} // namespace scn

#endif
