// scn_params.cpp — host only: error slot and scn_medium_desc -> kernel parameters.
//
// Everything the reference derives once per medium with the HOST libm (kernel radii, variance normalisation,
// multi-resolution level tables whose integer seed offsets come out of logf/powf) is derived here with the same
// libm calls, so that no integer or rejection-test operand ever depends on device libm.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "scn_internal.h"
#include <initializer_list>

namespace scn {

static thread_local char g_err[512] = "";
int set_error(int code, const char *fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
    return code;
}
const char *last_error() { return g_err; }

// ---------------------------------------------------------------- host: descriptor -> kernel parameters
static float host_se_variance3d(const Params &p, float R, bool is_identity, float gs, float ls) { // GPF:741-754
    double per_unit = p.density / (R * R * R);
    double det_sqrt = 1.0;
    if (!is_identity) { det_sqrt = p.l_aniso[0] * p.l_aniso[1] * p.l_aniso[2]; if (p.use_mtx) det_sqrt = p.mC_dsi; det_sqrt *= std::pow((double)gs, 3); }
    det_sqrt *= std::pow((double)ls, 3);
    double integral = std::pow(M_PI, 1.5) * det_sqrt;
    return (float)(per_unit * integral);
}
// Eigen's fixed-size 3x3 inverse (cofactor form, LU/InverseImpl.h:141-179), row-major
static void host_mat3_inverse(const float a[9], float r[9]) {
    auto A = [&](int i, int j) { return a[(i % 3) * 3 + (j % 3)]; };
    auto cof = [&](int i, int j) { return A(i + 1, j + 1) * A(i + 2, j + 2) - A(i + 1, j + 2) * A(i + 2, j + 1); };
    float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    float inv = 1.0f / (c00 * a[0] + (c10 * a[3] + c20 * a[6]));
    r[0] = c00 * inv; r[1] = c10 * inv; r[2] = c20 * inv;
    r[3] = cof(0, 1) * inv; r[4] = cof(1, 1) * inv; r[5] = cof(2, 1) * inv;
    r[6] = cof(0, 2) * inv; r[7] = cof(1, 2) * inv; r[8] = cof(2, 2) * inv;
}
static void field_k(const scn_field_desc &f, FieldK &k) {
    memset(&k, 0, sizeof k);
    k.type = f.type;
    for (int i = 0; i < 3; ++i) k.val[i] = f.val[i];
    const double c = 1.; // _const (never read from JSON)
    double mn = f.min + c, mx = f.max + c, mn2 = f.min2 + c, mx2 = f.max2 + c;
    k.la[0] = std::log(mn * mn); k.lb[0] = std::log(mx * mx); k.scale[0] = 1.0 / (f.end - f.start); k.off[0] = -f.start * k.scale[0];
    k.la[1] = std::log(mn2 * mn2); k.lb[1] = std::log(mx2 * mx2); k.scale[1] = 1.0 / (f.end2 - f.start2); k.off[1] = -f.start2 * k.scale[1];
    k.fmin = f.min; k.fmax = f.max;
    if (f.type == SCN_FIELD_REGULAR_GRID) { // Box3d::diagonal() = max - min (math/Box.hpp:62-65)
        k.grid = f.grid.values; k.gres = (int32_t)f.grid.res; k.ginterp = f.grid.interp;
        for (int i = 0; i < 3; ++i) { k.gmin[i] = f.grid.bounds_min[i]; k.gdiag[i] = f.grid.bounds_max[i] - f.grid.bounds_min[i]; }
    }
}
static int check_grid(const scn_field_desc &f, const char *name) {
    if (f.type != SCN_FIELD_REGULAR_GRID) return SCN_OK;
    if (!f.grid.values) return set_error(SCN_EINVAL, "%s: regular_grid without values", name);
    if (f.grid.res == 0 || f.grid.res > 2048) return set_error(SCN_EINVAL, "%s: regular_grid res %u out of range", name, f.grid.res);
    if (f.grid.interp < SCN_INTERP_POINT || f.grid.interp > SCN_INTERP_QUADRATIC) return set_error(SCN_EINVAL, "%s: bad interpolate method", name);
    return SCN_OK;
}

int params_from_desc(const scn_medium_desc &d, Params &p) {
    memset(&p, 0, sizeof p);
    if (d.correlation_context < SCN_CTX_GLOBAL || d.correlation_context > SCN_CTX_NONE)
        return set_error(SCN_EINVAL, "correlation_context is mandatory (media/GaussianProcessMedium.cpp:104-106)");
    if (d.cov_type < SCN_COV_SQUARED_EXPONENTIAL || d.cov_type > SCN_COV_GABOR_ISO) return set_error(SCN_EINVAL, "bad cov_type");
    const bool alt = d.cov_type >= SCN_COV_MATERN;
    if (alt) {
        // these kernels have no second derivative and no 1D family (math/GPFunctions.hpp:1336-1345,1509-1516 print "not implemented"), and
        // their fromJson never writes _world_to_local (math/GPFunctions.cpp:870-880,1087-1097): only world-space 3D sampling without
        // Renewal conditioning is defined behaviour in the reference
        if (d.sampling_1d) return set_error(SCN_EUNSUPPORTED, "matern / gabor kernels have no 1D sampling in the reference");
        if (d.isotropic_3d_sampling) return set_error(SCN_EUNSUPPORTED, "matern / gabor kernels: isotropic_3D_sampling reads uninitialised transforms in the reference; use world-space sampling");
        if (!d.single_realization && (d.correlation_context == SCN_CTX_RENEWAL || d.correlation_context == SCN_CTX_RENEWAL_PLUS))
            return set_error(SCN_EUNSUPPORTED, "matern / gabor kernels have no second derivative: renewal / renewal+ conditioning is not defined for them");
        if (d.cov_type == SCN_COV_MATERN && d.matern_v != 0.5f && d.matern_v != 1.5f && d.matern_v != 2.5f)
            return set_error(SCN_EINVAL, "Matern kernel only implemented for v = 0.5, 1.5, 2.5");
    }
    if (d.mean.type < SCN_MEAN_HOMOGENEOUS || d.mean.type > SCN_MEAN_PROCEDURAL) return set_error(SCN_EINVAL, "bad mean type");
    for (const scn_mean_desc *md : {&d.mean, &d.mean_additional}) {
        if (md->type != SCN_MEAN_PROCEDURAL) continue;
        if (md->sdf_func < SCN_SDF_FIELD || md->sdf_func > SCN_SDF_KNOB_OUTER) return set_error(SCN_EINVAL, "procedural mean: bad sdf_func");
        if (md->sdf_func == SCN_SDF_FIELD && md->f.type == SCN_FIELD_NONE) return set_error(SCN_EINVAL, "procedural mean: neither a function nor a field");
    }
    if (!(d.impulse_density >= 0.f) || d.impulse_density > 4096.f) return set_error(SCN_EINVAL, "impulse_density out of range");
    if (d.min_step == 0) return set_error(SCN_EINVAL, "min_step must be positive");
    p.n_imp = (uint32_t)d.impulse_density; p.density = d.impulse_density; p.global_seed = d.seed;
    p.single_realization = d.single_realization; p.iso3d = d.isotropic_3d_sampling; p.sampling_1d = d.sampling_1d;
    p.ctx = d.correlation_context; p.corr_xy = d.correlation_xy;
    p.activate_conditioning = !d.single_realization && (d.correlation_context == SCN_CTX_RENEWAL || d.correlation_context == SCN_CTX_RENEWAL_PLUS);
    p.scheme = (!d.single_realization && d.sampling_1d) ? d.sampling_scheme_1d : SCN_SCHEME_UNI;
    p.stationary = d.cov_type != SCN_COV_PROC_NONSTATIONARY;
    p.kernel = alt ? d.cov_type : SCN_COV_SQUARED_EXPONENTIAL;
    p.k_v = d.matern_v; p.k_l = d.length_scale;
    p.k_a = (float)(1.0 / d.gabor_a_inv); p.k_f = (float)(1.0 / d.gabor_f_inv); // GPF:1090-1093
    for (int i = 0; i < 3; ++i) p.k_omega[i] = d.gabor_omega[i];
    p.multires = !p.stationary && d.multires;
    p.surf_vol_sep = d.surf_vol_phase_separate; p.surf_vol_thresh = d.surf_vol_phase_amp_thresh;
    p.step_size = d.step_size; p.min_step = d.min_step; p.max_bounces = d.max_bounces;
    { float sa = d.sigma_a * d.density, ss = d.sigma_s * d.density, st = sa + ss; p.absorption_only = ss == 0.0f; p.sigma_s_over_t = ss / st; }
    p.sigma = d.sigma; p.kernel_scale = d.kernel_scale;
    float l_conv = d.length_scale * std::sqrt(2.f) / 2; // GPF:658
    for (int i = 0; i < 3; ++i) {
        p.l_aniso[i] = l_conv * d.aniso[i];
        float inv = 1.0f / p.l_aniso[i];
        p.l_aniso_inv[i] = (std::isinf(inv) || std::isnan(inv)) ? 0.f : inv; // filterWithZero
    }
    p.l_aniso_max = std::fmax(std::fmax(p.l_aniso[0], p.l_aniso[1]), p.l_aniso[2]);
    p.use_mtx = d.use_aniso_mtx != 0;
    p.mC_dsi = 1.0;
    for (int i = 0; i < 9; ++i) p.mL[i] = p.mW[i] = p.mC[i] = (i % 4 == 0) ? 1.f : 0.f;
    if (p.use_mtx) { // GPF:668-675; Eigen's fixed-size inverse (cofactors) and three-term reductions x0 + (x1 + x2)
        if (alt) return set_error(SCN_EUNSUPPORTED, "useAnisoMtx: squared-exponential kernel only");
        for (int i = 0; i < 9; ++i) p.mL[i] = l_conv * d.aniso_mtx[i];
        host_mat3_inverse(p.mL, p.mW);
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) p.mC[i * 3 + j] = p.mW[i] * p.mW[j] + (p.mW[3 + i] * p.mW[3 + j] + p.mW[6 + i] * p.mW[6 + j]);
        const float *c = p.mC; // Eigen's bruteforce_det3_helper order
        const float det = c[0] * (c[4] * c[8] - c[5] * c[7]) - c[1] * (c[3] * c[8] - c[5] * c[6]) + c[2] * (c[3] * c[7] - c[4] * c[6]);
        p.mC_dsi = 1.0 / std::sqrt(det); // GPF:749: double division of a float square root (std::sqrt(float))
        // splattingKernelRadius: max component of col0 + col1 + col2 (GPF:704-706)
        p.l_aniso_max = std::fmax(std::fmax((p.mL[0] + p.mL[1]) + p.mL[2], (p.mL[3] + p.mL[4]) + p.mL[5]), (p.mL[6] + p.mL[7]) + p.mL[8]);
    }
    p.ls_max = 1.f;
    if (!p.stationary && d.ls.type != SCN_FIELD_NONE) {
        const scn_field_desc &f = d.ls;
        if (f.type == SCN_FIELD_CONSTANT) p.ls_max = (float)std::fmax(std::fmax(f.val[0], f.val[1]), f.val[2]);
        else if (f.type == SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT) p.ls_max = (float)(std::fmax(f.max, f.min) * std::fmax(f.max2, f.min2));
        else if (f.type == SCN_FIELD_SANDSTONE || f.type == SCN_FIELD_RUST) p.ls_max = 1.f; // ProceduralNoiseVec::maxVal (GPF:131-137)
        else if (f.type == SCN_FIELD_REGULAR_GRID) p.ls_max = 1.f; // accepted in the multi-resolution form only (below), which never reads it
        else p.ls_max = (float)std::fmax(f.max, f.min);
    }
    if (int rc = check_grid(d.ls, "ls")) return rc;
    if (int rc = check_grid(d.var, "var")) return rc;
    if (int rc = check_grid(d.aniso_field, "aniso")) return rc;
    // RegularGridVector does not override ProceduralVector::maxVal() (math/GPFunctions.hpp:474-477 prints "maxVal() not implemented!" and
    // returns 0), so sparseConvNoiseMaxLateralScale() is 0 for a grid "ls" and the brute-force form divides by it (GPF:1227, 1246, 1262-1282):
    // only the multi-resolution form, which never reads it, is defined behaviour in the reference
    if (!p.stationary && d.ls.type == SCN_FIELD_REGULAR_GRID && !d.multires)
        return set_error(SCN_EUNSUPPORTED, "\"ls\": regular_grid needs \"multiResolutionGrid\": true (the reference's maximum lateral scale of a grid field is 0: math/GPFunctions.hpp:474-477)");
    field_k(d.ls, p.ls); field_k(d.var, p.var); field_k(d.aniso_field, p.aniso_f);
    p.ls.gcomp = 3; p.var.gcomp = 1; p.aniso_f.gcomp = 1;
    if (p.stationary) { p.ls.type = SCN_FIELD_NONE; p.var.type = SCN_FIELD_NONE; p.aniso_f.type = SCN_FIELD_NONE; }
    p.has_aniso = p.aniso_f.type != SCN_FIELD_NONE;
    if (p.has_aniso && d.sampling_1d)
        return set_error(SCN_EUNSUPPORTED, "proc_nonstationary \"aniso\" with 1D sampling reads StationaryCovariance::_local_to_world, whose off-diagonal the reference never initialises; 3D sampling only");
    // sparseConvNoiseMaxAnisotropyScale = max(a, 1/a) with _anisotropyOnAxis = 1.5 (GPF:1741-1745, GPF.hpp:2372)
    const float aniso_max = p.has_aniso ? 1.5f : 1.f;
    p.log_base = std::log(2.5f);
    p.ls_f_a = (float)(0.5 * 1.4426950408889634 * p.ls.la[0]); p.ls_f_b = (float)(0.5 * 1.4426950408889634 * (p.ls.lb[0] - p.ls.la[0]));
    p.ls_f_scale = (float)p.ls.scale[0]; p.ls_f_off = (float)p.ls.off[0]; p.inv_log2_base = (float)(1.0 / std::log2(2.5));
    p.world_scale = p.stationary ? 1.f : p.ls_max;
    p.add_seed_world = (int)std::floor(std::log(p.world_scale) / std::log(2.5f));
    p.R_iso = d.kernel_scale;
    {   // splattingKernelRadius(false, 1.0) (GPF:696-709, 1245-1249)
        float ls = 1.0f;
        if (!p.stationary) { ls *= p.multires ? 1.0 : p.ls_max; ls *= aniso_max; }
        p.R_world = d.kernel_scale * ls * p.l_aniso_max;
    }
    p.nf_iso = std::sqrt(host_se_variance3d(p, p.R_iso, true, 1.0f, 1.0f));
    p.nf_world = std::sqrt(host_se_variance3d(p, p.R_world, false, p.world_scale, 1.0f));
    { double per_unit = p.density / p.R_iso; double integral = std::sqrt(M_PI) * 1.0f; p.nf_1d = std::sqrt((float)(per_unit * integral)); }
    for (int i = 0; i < LVL_COUNT; ++i) {
        int k = LVL_MIN + i;
        float s = std::pow(2.5f, (float)k); // powf, like pow(_base, level) in SCN:64-65
        p.lvl_scale[i] = s;
        p.lvl_add_seed[i] = (int)std::floor(std::log(s) / std::log(2.5f));
        p.lvl_R_world[i] = d.kernel_scale * (s * aniso_max) * p.l_aniso_max;
        p.lvl_nf_world[i] = std::sqrt(host_se_variance3d(p, p.lvl_R_world[i], false, s, 1.0f));
    }
    if (alt) {
        // splattingKernelRadius(false, 1.0) and sparseConvNoiseVariance3D of the alternative kernels, same float/double promotions
        const float sf = d.kernel_scale;
        double integral = 0.;
        if (d.cov_type == SCN_COV_MATERN) { // GPF:1020-1044
            float m = 0.f;
            for (int i = 0; i < 3; ++i) { float v = d.length_scale / std::sqrt(d.aniso[i]); if (std::isinf(v) || std::isnan(v)) v = 0.f; m = std::fmax(m, v); }
            p.R_world = (float)(sf * 1.0f * std::sqrt(2) / 2 * m);
            const float l = d.length_scale;
            if (d.matern_v == 0.5f) integral = 2.0 * M_PI * l;
            else if (d.matern_v == 1.5f) integral = std::pow(M_PI * l, 3) / (24 * std::sqrt(3));
            else integral = M_PI * std::pow(l, 3) / (5 * std::sqrt(5));
        } else if (d.cov_type == SCN_COV_GABOR_ANISO) { // GPF:1127-1140
            p.R_world = (float)(sf * std::sqrt(2) / 2 * 1.0 / p.k_a);
            float q = p.k_f / p.k_a; q = q * q;
            integral = std::pow(1.0 / p.k_a, 3) * (1 + std::exp(-2.0 * M_PI * q)) / (4 * std::sqrt(2));
        } else { // GPF:1185-1198
            p.R_world = (float)(sf * std::sqrt(2) / 4 * 1.0 / p.k_a);
            integral = 2 * std::sqrt(2) * M_PI * (p.k_f * p.k_f) / p.k_a * (1 - std::exp(-2 * M_PI * p.k_f / (p.k_a * p.k_a)));
        }
        double per_unit = p.density / (p.R_world * p.R_world * p.R_world);
        p.nf_world = std::sqrt((float)(per_unit * integral));
    }
    p.gp_id = d.gp_id; p.has_mean_add = d.has_mean_additional; p.id_add = d.id_additional;
    {   // Mat4f::rotAxis (math/Mat4f.cpp:132-148) about (1, 0, 0): a11 = c + 1 * 1 * c1, the rotation block is (c, -s; s, c); Angle::degToRad = a * (PI / 180.0f)
        const float PI_F = 3.1415926536f, angles[3] = {-90.f, 90.f, -45.f};
        for (int i = 0; i < 3; ++i) {
            const float a = angles[i] * (PI_F / 180.0f), sn = std::sin(a), cs = std::cos(a), c1 = 1.0f - cs;
            p.knob_rot[3 * i] = cs + 1.0f * 1.0f * c1; p.knob_rot[3 * i + 1] = cs; p.knob_rot[3 * i + 2] = sn;
        }
        p.knob_rot[9] = (float)(PI_F / 3.); p.knob_rot[10] = PI_F / 3.f;
    }
    p.mean = d.mean; p.mean_add = d.mean_additional;
    if (int rc = check_grid(d.mean.f, "mean.f")) return rc;
    if (int rc = check_grid(d.mean_additional.f, "mean_additional.f")) return rc;
    field_k(d.mean.f, p.mean_f); field_k(d.mean_additional.f, p.mean_add_f); p.mean_f.gcomp = p.mean_add_f.gcomp = 1;
    p.mean.f.grid.values = nullptr; p.mean_add.f.grid.values = nullptr; // the kernels read the grids through mean_f / mean_add_f only
    return SCN_OK;
}

} // namespace scn
