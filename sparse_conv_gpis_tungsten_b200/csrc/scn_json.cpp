// scn_json.cpp — medium JSON -> scn_medium_desc (host only, no CUDA).
//
// Accepts the reference's "sparse_conv_noise" medium JSON unchanged and flattens it.  Key names and
// defaults follow the reference's fromJson chain (paths relative to the reference's src/core/):
//   media/SparseConvolutionNoiseMedium.cpp:17-35,57-73   media/GaussianProcessMedium.cpp:97-126
//   media/Medium.cpp:28-38    math/GaussianProcess.cpp:25-39,172-190    math/GPFunctions.cpp:654-679
//   math/GPFunctions.hpp:604-800 (procedural fields), 867-1005 (means), 2199-2205 (non-stationary)
// Like the reference, unknown keys are ignored and invalid enum strings are an error
// (media/SparseConvolutionNoiseMedium.cpp:44, media/GaussianProcessMedium.cpp:40).
#include "scn_internal.h"

#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace scn {

namespace {

struct JVal {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;
    const JVal *get(const char *key) const {
        if (kind != Obj) return nullptr;
        for (auto &kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct Parser {
    const char *s; size_t i = 0, n; std::string err;
    explicit Parser(const char *text) : s(text), n(strlen(text)) {}
    void ws() { while (i < n && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) ++i; }
    bool fail(const std::string &m) { if (err.empty()) err = m + " at offset " + std::to_string(i); return false; }
    bool value(JVal &v, int depth = 0) {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (i >= n) return fail("unexpected end");
        char c = s[i];
        if (c == '{') {
            v.kind = JVal::Obj; ++i; ws();
            if (i < n && s[i] == '}') { ++i; return true; }
            for (;;) {
                ws(); JVal k;
                if (i >= n || s[i] != '"' || !string(k.str)) return fail("expected object key");
                ws(); if (i >= n || s[i] != ':') return fail("expected ':'"); ++i;
                JVal child; if (!value(child, depth + 1)) return false;
                v.obj.emplace_back(k.str, std::move(child));
                ws(); if (i < n && s[i] == ',') { ++i; continue; }
                if (i < n && s[i] == '}') { ++i; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            v.kind = JVal::Arr; ++i; ws();
            if (i < n && s[i] == ']') { ++i; return true; }
            for (;;) {
                JVal child; if (!value(child, depth + 1)) return false;
                v.arr.push_back(std::move(child));
                ws(); if (i < n && s[i] == ',') { ++i; continue; }
                if (i < n && s[i] == ']') { ++i; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') { v.kind = JVal::Str; return string(v.str); }
        if (!strncmp(s + i, "true", 4)) { v.kind = JVal::Bool; v.b = true; i += 4; return true; }
        if (!strncmp(s + i, "false", 5)) { v.kind = JVal::Bool; v.b = false; i += 5; return true; }
        if (!strncmp(s + i, "null", 4)) { v.kind = JVal::Null; i += 4; return true; }
        char *end = nullptr;
        double d = strtod(s + i, &end);
        if (end == s + i) return fail("unexpected character");
        v.kind = JVal::Num; v.num = d; i = (size_t)(end - s);
        return true;
    }
    bool string(std::string &out) {
        ++i; // opening quote
        while (i < n && s[i] != '"') {
            if (s[i] == '\\' && i + 1 < n) {
                char e = s[i + 1];
                switch (e) {
                case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                case 'u': out += '?'; i += 4; break;
                default: out += e;
                }
                i += 2;
            } else out += s[i++];
        }
        if (i >= n) return fail("unterminated string");
        ++i; return true;
    }
};

// JsonPtr::getField equivalents: leave the destination untouched when the key is absent.
bool getNum(const JVal &o, const char *k, double &d) {
    const JVal *v = o.get(k); if (!v) return false;
    if (v->kind == JVal::Num) { d = v->num; return true; }
    if (v->kind == JVal::Bool) { d = v->b ? 1 : 0; return true; }
    return false;
}
void getF(const JVal &o, const char *k, float &f) { double d; if (getNum(o, k, d)) f = (float)d; }
void getD(const JVal &o, const char *k, double &f) { double d; if (getNum(o, k, d)) f = d; }
void getU(const JVal &o, const char *k, uint32_t &u) { double d; if (getNum(o, k, d)) u = (uint32_t)(int64_t)d; }
void getI(const JVal &o, const char *k, int32_t &u) { double d; if (getNum(o, k, d)) u = (int32_t)d; }
void getB(const JVal &o, const char *k, int32_t &b) {
    const JVal *v = o.get(k); if (!v) return;
    if (v->kind == JVal::Bool) b = v->b; else if (v->kind == JVal::Num) b = v->num != 0;
}
bool getS(const JVal &o, const char *k, std::string &s) { const JVal *v = o.get(k); if (v && v->kind == JVal::Str) { s = v->str; return true; } return false; }
template <typename T> void getVec3(const JVal &o, const char *k, T *dst) {
    const JVal *v = o.get(k); if (!v) return;
    if (v->kind == JVal::Num) { dst[0] = dst[1] = dst[2] = (T)v->num; return; }
    if (v->kind == JVal::Arr && v->arr.size() >= 3) for (int i = 0; i < 3; ++i) if (v->arr[i].kind == JVal::Num) dst[i] = (T)v->arr[i].num;
}

thread_local std::string g_base_dir; // directory relative resource paths resolve against (scn_desc_from_json_at)

std::string joinPath(const std::string &dir, const std::string &rel) {
    if (rel.empty() || rel[0] == '/' || dir.empty()) return rel;
    return dir.back() == '/' ? dir + rel : dir + "/" + rel;
}
std::string parentDir(const std::string &path) { size_t k = path.find_last_of('/'); return k == std::string::npos ? std::string() : path.substr(0, k); }
bool readFile(const std::string &path, std::string &out) {
    FILE *fp = fopen(path.c_str(), "rb"); if (!fp) return false;
    char buf[1 << 16]; size_t k;
    while ((k = fread(buf, 1, sizeof buf, fp)) > 0) out.append(buf, k);
    fclose(fp); return true;
}
// RegularGridScalar / RegularGridVector::fromJson + loadResources -> load_grid (math/GPFunctions.hpp:386-401, 358-369, 335-344):
// "file" names a grid JSON {bounds_min, bounds_max, res, interpolate, path}; "path", relative to that file, holds the raw doubles.
int loadGrid(const JVal &j, scn_field_desc &f, bool vector) {
    std::string file;
    if (!getS(j, "file", file)) return set_error(SCN_EJSON, "regular_grid field: missing \"file\"");
    const std::string gpath = joinPath(g_base_dir, file);
    std::string text;
    if (!readFile(gpath, text)) return set_error(SCN_EJSON, "regular_grid field: cannot read '%s'", gpath.c_str());
    Parser p(text.c_str()); JVal g;
    if (!p.value(g) || g.kind != JVal::Obj) return set_error(SCN_EJSON, "regular_grid '%s': JSON parse error: %s", gpath.c_str(), p.err.c_str());
    scn_grid_desc &G = f.grid;
    getVec3(g, "bounds_min", G.bounds_min); getVec3(g, "bounds_max", G.bounds_max); getU(g, "res", G.res);
    std::string interp = "linear";
    getS(g, "interpolate", interp);
    if (interp == "point") G.interp = SCN_INTERP_POINT; else if (interp == "linear") G.interp = SCN_INTERP_LINEAR;
    else if (interp == "quadratic") G.interp = SCN_INTERP_QUADRATIC; else return set_error(SCN_EJSON, "Invalid interpolate method: '%s'", interp.c_str());
    std::string vpath;
    if (!getS(g, "path", vpath)) return set_error(SCN_EJSON, "regular_grid '%s': missing \"path\"", gpath.c_str());
    vpath = joinPath(parentDir(gpath), vpath);
    if (G.res == 0 || G.res > 2048) return set_error(SCN_EINVAL, "regular_grid '%s': res %u out of range", gpath.c_str(), G.res);
    const size_t count = (size_t)G.res * G.res * G.res * (vector ? 3 : 1);
    std::string raw;
    if (!readFile(vpath, raw)) return set_error(SCN_EJSON, "regular_grid: cannot read values file '%s'", vpath.c_str());
    if (raw.size() / sizeof(double) < count) return set_error(SCN_EINVAL, "regular_grid '%s': %zu values, res^3 needs %zu", vpath.c_str(), raw.size() / sizeof(double), count);
    double *vals = (double *)malloc(count * sizeof(double));
    if (!vals) return set_error(SCN_ENOMEM, "out of memory");
    memcpy(vals, raw.data(), count * sizeof(double));
    G.values = vals; G.owned = 1;
    f.type = SCN_FIELD_REGULAR_GRID;
    return SCN_OK;
}

int parseField(const JVal &j, scn_field_desc &f, bool vector) {
    std::string type;
    if (!getS(j, "type", type)) return set_error(SCN_EJSON, "procedural field: missing \"type\"");
    if (type == "regular_grid") return loadGrid(j, f, vector);
    if (type == "constant") {
        f.type = SCN_FIELD_CONSTANT;
        if (vector) getVec3(j, "val", f.val); else { getD(j, "val", f.val[0]); f.val[1] = f.val[2] = f.val[0]; }
        return SCN_OK;
    }
    if (type != "noise") return set_error(SCN_EUNSUPPORTED, "procedural field type '%s' is not covered", type.c_str());
    std::string noise = "bottom_top";
    getS(j, "noise", noise);
    if (noise == "bottom_top") f.type = SCN_FIELD_BOTTOM_TOP;
    else if (noise == "left_right") f.type = SCN_FIELD_LEFT_RIGHT;
    else if (noise == "front_back") f.type = SCN_FIELD_FRONT_BACK;
    else if (noise == "bottom_top_left_right") f.type = SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT;
    else if (noise == "sandstone") f.type = SCN_FIELD_SANDSTONE;
    else if (noise == "rust") f.type = SCN_FIELD_RUST;
    else return set_error(SCN_EJSON, "Invalid noise typ function: '%s'", noise.c_str());
    getD(j, "min", f.min); getD(j, "max", f.max); getD(j, "start", f.start); getD(j, "end", f.end);
    getD(j, "min2", f.min2); getD(j, "max2", f.max2); getD(j, "start2", f.start2); getD(j, "end2", f.end2);
    return SCN_OK;
}

// JsonPtr::get(Mat4f &) (io/JsonPtr.cpp:108-183): sixteen numbers row-major, or {position, scale, rotation} — the default axes pass
// gramSchmidt unchanged; look_at / up / explicit axes are not covered.  Float arithmetic in the reference's order.
int parseTransform(const JVal &t, float M[16]) {
    if (t.kind == JVal::Arr) {
        if (t.arr.size() != 16) return set_error(SCN_EJSON, "Trying to parse a matrix, but this array has the wrong size (need 16 elements, received %zu)", t.arr.size());
        for (int i = 0; i < 16; ++i) M[i] = (float)t.arr[i].num;
        return SCN_OK;
    }
    if (t.kind != JVal::Obj) return set_error(SCN_EJSON, "Parameter has wrong type: Expecting a matrix value here");
    if (t.get("look_at") || t.get("up") || t.get("x_axis") || t.get("y_axis") || t.get("z_axis"))
        return set_error(SCN_EUNSUPPORTED, "transform: look_at / up / explicit axes are not covered (position, scale, rotation, or 16 numbers)");
    float x[3] = {1, 0, 0}, y[3] = {0, 1, 0}, z[3] = {0, 0, 1}, pos[3] = {0, 0, 0}, sc[3], rot[3];
    getVec3(t, "position", pos);
    if (t.get("scale")) { sc[0] = sc[1] = sc[2] = 1.f; getVec3(t, "scale", sc); for (int i = 0; i < 3; ++i) { x[i] *= sc[0]; y[i] *= sc[1]; z[i] *= sc[2]; } }
    if (t.get("rotation")) { // Mat4f::rotYXZ (math/Mat4f.cpp:118-130), then tform * axis (transformVector)
        rot[0] = rot[1] = rot[2] = 0.f; getVec3(t, "rotation", rot);
        float r[3]; for (int i = 0; i < 3; ++i) r[i] = rot[i] * 3.1415926536f / 180.0f;
        const float c[3] = {std::cos(r[0]), std::cos(r[1]), std::cos(r[2])}, s[3] = {std::sin(r[0]), std::sin(r[1]), std::sin(r[2])};
        const float R[9] = {c[1] * c[2] - s[1] * s[0] * s[2], -c[1] * s[2] - s[1] * s[0] * c[2], -s[1] * c[0],
                            c[0] * s[2], c[0] * c[2], -s[0],
                            s[1] * c[2] + c[1] * s[0] * s[2], -s[1] * s[2] + c[1] * s[0] * c[2], c[1] * c[0]};
        auto apply = [&](float *v) { const float a = v[0], b = v[1], cc = v[2]; for (int i = 0; i < 3; ++i) v[i] = R[3 * i] * a + R[3 * i + 1] * b + R[3 * i + 2] * cc; };
        apply(x); apply(y); apply(z);
    }
    const float out[16] = {x[0], y[0], z[0], pos[0], x[1], y[1], z[1], pos[1], x[2], y[2], z[2], pos[2], 0.f, 0.f, 0.f, 1.f};
    memcpy(M, out, sizeof out);
    return SCN_OK;
}
// Mat4f::invert (math/Mat4f.hpp:75-101): cofactors in the reference's term order, identity when the determinant is 0
void mat4Invert(const float a[16], float out[16]) {
    float inv[16];
    inv[ 0] =  a[5]*a[10]*a[15] - a[5]*a[11]*a[14] - a[9]*a[6]*a[15] + a[9]*a[7]*a[14] + a[13]*a[6]*a[11] - a[13]*a[7]*a[10];
    inv[ 1] = -a[1]*a[10]*a[15] + a[1]*a[11]*a[14] + a[9]*a[2]*a[15] - a[9]*a[3]*a[14] - a[13]*a[2]*a[11] + a[13]*a[3]*a[10];
    inv[ 2] =  a[1]*a[ 6]*a[15] - a[1]*a[ 7]*a[14] - a[5]*a[2]*a[15] + a[5]*a[3]*a[14] + a[13]*a[2]*a[ 7] - a[13]*a[3]*a[ 6];
    inv[ 3] = -a[1]*a[ 6]*a[11] + a[1]*a[ 7]*a[10] + a[5]*a[2]*a[11] - a[5]*a[3]*a[10] - a[ 9]*a[2]*a[ 7] + a[ 9]*a[3]*a[ 6];
    inv[ 4] = -a[4]*a[10]*a[15] + a[4]*a[11]*a[14] + a[8]*a[6]*a[15] - a[8]*a[7]*a[14] - a[12]*a[6]*a[11] + a[12]*a[7]*a[10];
    inv[ 5] =  a[0]*a[10]*a[15] - a[0]*a[11]*a[14] - a[8]*a[2]*a[15] + a[8]*a[3]*a[14] + a[12]*a[2]*a[11] - a[12]*a[3]*a[10];
    inv[ 6] = -a[0]*a[ 6]*a[15] + a[0]*a[ 7]*a[14] + a[4]*a[2]*a[15] - a[4]*a[3]*a[14] - a[12]*a[2]*a[ 7] + a[12]*a[3]*a[ 6];
    inv[ 8] =  a[4]*a[ 9]*a[15] - a[4]*a[11]*a[13] - a[8]*a[5]*a[15] + a[8]*a[7]*a[13] + a[12]*a[5]*a[11] - a[12]*a[7]*a[ 9];
    inv[ 7] =  a[0]*a[ 6]*a[11] - a[0]*a[ 7]*a[10] - a[4]*a[2]*a[11] + a[4]*a[3]*a[10] + a[ 8]*a[2]*a[ 7] - a[ 8]*a[3]*a[ 6];
    inv[ 9] = -a[0]*a[ 9]*a[15] + a[0]*a[11]*a[13] + a[8]*a[1]*a[15] - a[8]*a[3]*a[13] - a[12]*a[1]*a[11] + a[12]*a[3]*a[ 9];
    inv[10] =  a[0]*a[ 5]*a[15] - a[0]*a[ 7]*a[13] - a[4]*a[1]*a[15] + a[4]*a[3]*a[13] + a[12]*a[1]*a[ 7] - a[12]*a[3]*a[ 5];
    inv[11] = -a[0]*a[ 5]*a[11] + a[0]*a[ 7]*a[ 9] + a[4]*a[1]*a[11] - a[4]*a[3]*a[ 9] - a[ 8]*a[1]*a[ 7] + a[ 8]*a[3]*a[ 5];
    inv[12] = -a[4]*a[ 9]*a[14] + a[4]*a[10]*a[13] + a[8]*a[5]*a[14] - a[8]*a[6]*a[13] - a[12]*a[5]*a[10] + a[12]*a[6]*a[ 9];
    inv[13] =  a[0]*a[ 9]*a[14] - a[0]*a[10]*a[13] - a[8]*a[1]*a[14] + a[8]*a[2]*a[13] + a[12]*a[1]*a[10] - a[12]*a[2]*a[ 9];
    inv[14] = -a[0]*a[ 5]*a[14] + a[0]*a[ 6]*a[13] + a[4]*a[1]*a[14] - a[4]*a[2]*a[13] - a[12]*a[1]*a[ 6] + a[12]*a[2]*a[ 5];
    inv[15] =  a[0]*a[ 5]*a[10] - a[0]*a[ 6]*a[ 9] - a[4]*a[1]*a[10] + a[4]*a[2]*a[ 9] + a[ 8]*a[1]*a[ 6] - a[ 8]*a[2]*a[ 5];
    const float det = a[0]*inv[0] + a[1]*inv[4] + a[2]*inv[8] + a[3]*inv[12];
    if (det == 0.0f) { for (int i = 0; i < 16; ++i) out[i] = (i % 5 == 0) ? 1.f : 0.f; return; }
    const float invDet = 1.0f / det;
    for (int i = 0; i < 16; ++i) out[i] = inv[i] * invDet;
}

int sdfFunction(const std::string &fn, int32_t &out) { // SdfFunctions::stringToFunction (math/SdfFunctions.cpp:22-36)
    if (fn == "knob") out = SCN_SDF_KNOB; else if (fn == "knob_inner") out = SCN_SDF_KNOB_INNER; else if (fn == "knob_outer") out = SCN_SDF_KNOB_OUTER;
    else if (fn == "two_spheres") out = SCN_SDF_TWO_SPHERES; else if (fn == "plane") out = SCN_SDF_PLANE;
    else return set_error(SCN_EJSON, "Invalid function: '%s'", fn.c_str());
    return SCN_OK;
}
int parseMean(const JVal &j, scn_mean_desc &m) {
    std::string type;
    if (!getS(j, "type", type)) return set_error(SCN_EJSON, "mean: missing \"type\"");
    if (j.get("color") || j.get("emission")) return set_error(SCN_EUNSUPPORTED, "mean: procedural \"color\"/\"emission\" fields are host-side and not covered");
    if (type == "homogeneous") { m.type = SCN_MEAN_HOMOGENEOUS; getF(j, "offset", m.offset); }
    else if (type == "spherical") { m.type = SCN_MEAN_SPHERICAL; getVec3(j, "center", m.center); getF(j, "radius", m.radius); }
    else if (type == "linear") {
        m.type = SCN_MEAN_LINEAR; getVec3(j, "reference_point", m.ref); getVec3(j, "direction", m.dir); getF(j, "scale", m.scale); getF(j, "min", m.min);
        // _dir.normalize() in double (math/GPFunctions.hpp:963, math/Vec.hpp:184-189)
        double l2 = 0.; l2 += m.dir[0] * m.dir[0]; l2 += m.dir[1] * m.dir[1]; l2 += m.dir[2] * m.dir[2];
        double inv = 1.0 / std::sqrt(l2);
        for (int i = 0; i < 3; ++i) m.dir[i] *= inv;
    } else if (type == "procedural") { // ProceduralMean::fromJson (math/GPFunctions.cpp:250-266)
        m.type = SCN_MEAN_PROCEDURAL;
        float T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, inv[16];
        if (const JVal *t = j.get("transform")) if (int rc = parseTransform(*t, T)) return rc;
        mat4Invert(T, inv);
        for (int i = 0; i < 12; ++i) m.inv_transform[i] = inv[i];
        std::string fn = "knob";
        if (getS(j, "func", fn)) {
            if (int rc = sdfFunction(fn, m.sdf_func)) return rc;
        } else if (const JVal *f = j.get("f")) {
            m.sdf_func = SCN_SDF_FIELD;
            std::string ft; getS(*f, "type", ft);
            if (ft == "sdf") { // ProceduralSdf (math/GPFunctions.hpp:579-603): the same function table
                std::string fs; getS(*f, "function", fs);
                if (int rc = sdfFunction(fs, m.sdf_func)) return rc;
            } else if (int rc = parseField(*f, m.f, false)) return rc;
        } else m.sdf_func = SCN_SDF_KNOB; // ProceduralMean's default member: ProceduralSdf(Function::Knob) (math/GPFunctions.hpp:1052-1055, 581)
        getF(j, "min", m.min); getF(j, "scale", m.scale); getF(j, "offset", m.offset);
    } else if (type == "tabulated" || type == "mesh" || type == "neural")
        return set_error(SCN_EUNSUPPORTED, "mean type '%s' needs an OpenVDB grid / a mesh BVH / an MLP and is not covered", type.c_str());
    else return set_error(SCN_EJSON, "Invalid mean: '%s'", type.c_str());
    return SCN_OK;
}

int parseSE(const JVal &j, scn_medium_desc &d, bool allowAlt) {
    std::string type;
    if (!getS(j, "type", type)) return set_error(SCN_EJSON, "covariance: missing \"type\"");
    getF(j, "localScale", d.kernel_scale); // StationaryCovariance::fromJson, math/GPFunctions.hpp:1481-1484
    if (allowAlt && (type == "matern" || type == "gabor_aniso" || type == "gabor_iso")) {
        getF(j, "sigma", d.sigma);
        if (type == "matern") { // math/GPFunctions.cpp:870-880
            d.cov_type = SCN_COV_MATERN;
            getF(j, "v", d.matern_v); getF(j, "lengthScale", d.length_scale); getVec3(j, "aniso", d.aniso);
            return SCN_OK;
        }
        d.cov_type = type == "gabor_aniso" ? SCN_COV_GABOR_ANISO : SCN_COV_GABOR_ISO; // math/GPFunctions.cpp:1087-1097, 1156-1163
        getF(j, "a_inv", d.gabor_a_inv); getF(j, "f_inv", d.gabor_f_inv);
        if (d.cov_type == SCN_COV_GABOR_ANISO) {
            if (!j.get("omega")) return set_error(SCN_EINVAL, "gabor_aniso: \"omega\" is mandatory (the reference reads it uninitialised when absent)");
            getVec3(j, "omega", d.gabor_omega);
            float *o = d.gabor_omega; // Vec3f::normalize(): multiply by the reciprocal length (math/Vec.hpp)
            float len = std::sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]), inv = 1.0f / len;
            o[0] *= inv; o[1] *= inv; o[2] *= inv;
        }
        return SCN_OK;
    }
    if (type != "squared_exponential") {
        static const char *known[] = {"rational_quadratic", "matern", "gabor_aniso", "gabor_iso", "dot_product", "periodic", "thin_plate",
                                      "grid_nonstationary", "mg_nonstationary", "neural_nonstationary", "proc_nonstationary"};
        for (const char *k : known) if (type == k) return set_error(SCN_EUNSUPPORTED, "covariance '%s' is not covered here (squared_exponential, matern, gabor_aniso, gabor_iso and proc_nonstationary over squared_exponential are)", type.c_str());
        return set_error(SCN_EJSON, "Invalid covariance: '%s'", type.c_str());
    }
    getF(j, "sigma", d.sigma); getF(j, "lengthScale", d.length_scale); getVec3(j, "aniso", d.aniso);
    getB(j, "useAnisoMtx", d.use_aniso_mtx);
    if (d.use_aniso_mtx) { // "anisoMtx": nine numbers, row-major (io/JsonPtr.cpp:188-198, math/GPFunctions.cpp:669)
        const JVal *v = j.get("anisoMtx");
        if (v) {
            if (v->kind != JVal::Arr || v->arr.size() != 9) return set_error(SCN_EJSON, "anisoMtx: expecting an array of 9 numbers");
            for (int i = 0; i < 9; ++i) { if (v->arr[i].kind != JVal::Num) return set_error(SCN_EJSON, "anisoMtx: expecting numbers"); d.aniso_mtx[i] = (float)v->arr[i].num; }
        }
    }
    return SCN_OK;
}

int parseCov(const JVal &j, scn_medium_desc &d) {
    std::string type;
    if (!getS(j, "type", type)) return set_error(SCN_EJSON, "covariance: missing \"type\"");
    if (type != "proc_nonstationary") { d.cov_type = SCN_COV_SQUARED_EXPONENTIAL; return parseSE(j, d, true); }
    d.cov_type = SCN_COV_PROC_NONSTATIONARY;
    const JVal *inner = j.get("cov");
    if (!inner) return set_error(SCN_EJSON, "proc_nonstationary: missing \"cov\"");
    if (int rc = parseSE(*inner, d, false)) return rc;
    getB(j, "multiResolutionGrid", d.multires);
    if (const JVal *v = j.get("aniso")) if (int rc = parseField(*v, d.aniso_field, false)) return rc; // math/GPFunctions.cpp:1602-1604
    if (const JVal *v = j.get("var")) if (int rc = parseField(*v, d.var, false)) return rc;
    if (const JVal *v = j.get("ls")) if (int rc = parseField(*v, d.ls, true)) return rc;
    return SCN_OK;
}

int parseStandardGP(const JVal &j, scn_medium_desc &d, scn_mean_desc &mean, int32_t &id, bool withCov) {
    if (const JVal *m = j.get("mean")) { if (int rc = parseMean(*m, mean)) return rc; }
    else { mean.type = SCN_MEAN_HOMOGENEOUS; mean.offset = 0.f; } // GaussianProcess() default, math/GaussianProcess.hpp:243
    if (withCov) if (const JVal *c = j.get("covariance")) if (int rc = parseCov(*c, d)) return rc;
    getI(j, "id", id);
    if (j.get("conditioning_data")) return set_error(SCN_EUNSUPPORTED, "\"conditioning_data\" is not used by the sparse-convolution path");
    return SCN_OK;
}

int parseGP(const JVal &j, scn_medium_desc &d) {
    std::string type;
    if (!getS(j, "type", type)) return set_error(SCN_EJSON, "gaussian_process: missing \"type\"");
    if (type == "standard") return parseStandardGP(j, d, d.mean, d.gp_id, true);
    if (type == "csg") { // math/GaussianProcess.cpp:25-39: left GP + the right GP's mean and id
        const JVal *l = j.get("left"), *r = j.get("right");
        if (!l || !r) return set_error(SCN_EJSON, "csg: needs \"left\" and \"right\"");
        std::string lt, rt; getS(*l, "type", lt); getS(*r, "type", rt);
        if (lt != "standard" || rt != "standard") return set_error(SCN_EUNSUPPORTED, "csg of non-standard gaussian processes is not covered");
        if (int rc = parseStandardGP(*l, d, d.mean, d.gp_id, true)) return rc;
        d.has_mean_additional = 1; d.id_additional = 0;
        int32_t rid = 0;
        if (int rc = parseStandardGP(*r, d, d.mean_additional, rid, false)) return rc;
        d.id_additional = rid;
        return SCN_OK;
    }
    return set_error(SCN_EJSON, "Invalid gaussian_process: '%s'", type.c_str());
}

void fieldDefault(scn_field_desc &f);
void meanDefault(scn_mean_desc &m) {
    memset(&m, 0, sizeof m);
    m.type = SCN_MEAN_SPHERICAL; m.radius = 1.f; m.scale = 1.f; m.dir[0] = 1.; m.min = -FLT_MAX;
    m.inv_transform[0] = m.inv_transform[5] = m.inv_transform[10] = 1.f;
    fieldDefault(m.f);
}
void fieldDefault(scn_field_desc &f) {
    memset(&f, 0, sizeof f);
    f.type = SCN_FIELD_NONE; f.min = 1.; f.max = 500.; f.start = 0.; f.end = 1.; f.min2 = 1.; f.max2 = 500.; f.start2 = 0.; f.end2 = 1.;
    f.grid.interp = SCN_INTERP_LINEAR; // RegularGrid::interp default (math/GPFunctions.hpp:243)
}

} // namespace

int desc_default(scn_medium_desc *d) {
    if (!d) return set_error(SCN_EINVAL, "null descriptor");
    memset(d, 0, sizeof *d);
    d->step_size = 0.01f; d->min_step = 8; d->seed = 0; d->impulse_density = 3.0f;
    d->sampling_scheme_1d = SCN_SCHEME_UNI;
    d->sigma_a = 0.f; d->sigma_s = 0.f; d->density = 1.f;
    d->correlation_context = -1; // "goldfish": invalid until the key is given (media/GaussianProcessMedium.cpp:104-106)
    d->max_bounces = 1024;
    d->gp_id = 0; d->id_additional = 1;
    meanDefault(d->mean); meanDefault(d->mean_additional);
    d->cov_type = SCN_COV_SQUARED_EXPONENTIAL;
    d->sigma = 1.f; d->length_scale = 1.f; d->aniso[0] = d->aniso[1] = d->aniso[2] = 1.f; d->kernel_scale = 3.0f;
    d->aniso_mtx[0] = d->aniso_mtx[4] = d->aniso_mtx[8] = 1.f;
    fieldDefault(d->ls); fieldDefault(d->var); fieldDefault(d->aniso_field);
    d->matern_v = 1.0f; d->gabor_a_inv = 1.f; d->gabor_f_inv = 1.f; d->gabor_omega[2] = 1.f;
    return SCN_OK;
}

void desc_release(scn_medium_desc *d) {
    if (!d) return;
    for (scn_field_desc *f : {&d->ls, &d->var, &d->aniso_field, &d->mean.f, &d->mean_additional.f})
        if (f->grid.owned && f->grid.values) { free((void *)f->grid.values); f->grid.values = nullptr; f->grid.owned = 0; }
}

static int desc_from_json_impl(const char *text, scn_medium_desc *d);
int desc_from_json(const char *text, const char *base_dir, scn_medium_desc *d) {
    g_base_dir = base_dir ? base_dir : "";
    int rc = desc_from_json_impl(text, d);
    if (rc != SCN_OK && d) desc_release(d); // a grid read before the failing key
    return rc;
}

static int desc_from_json_impl(const char *text, scn_medium_desc *d) {
    if (!text || !d) return set_error(SCN_EINVAL, "null argument");
    if (int rc = desc_default(d)) return rc;
    Parser p(text);
    JVal root;
    if (!p.value(root)) return set_error(SCN_EJSON, "JSON parse error: %s", p.err.c_str());
    p.ws();
    if (p.i != p.n) return set_error(SCN_EJSON, "JSON parse error: trailing characters at offset %zu", p.i);
    if (root.kind != JVal::Obj) return set_error(SCN_EJSON, "medium JSON must be an object");
    std::string type;
    // "sparse_conv_noise_gpu" is the factory string the Tungsten-side shim registers (oracle/gpu_medium_shim.cpp)
    if (getS(root, "type", type) && type != "sparse_conv_noise" && type != "sparse_conv_noise_gpu")
        return set_error(SCN_EJSON, "medium type '%s' is not \"sparse_conv_noise\"", type.c_str());

    getI(root, "max_bounces", d->max_bounces);
    getF(root, "sigma_a", d->sigma_a); getF(root, "sigma_s", d->sigma_s); getF(root, "density", d->density);
    std::string ctx = "goldfish";
    getS(root, "correlation_context", ctx);
    if (ctx == "global") d->correlation_context = SCN_CTX_GLOBAL;
    else if (ctx == "renewal+") d->correlation_context = SCN_CTX_RENEWAL_PLUS;
    else if (ctx == "renewal") d->correlation_context = SCN_CTX_RENEWAL;
    else if (ctx == "none") d->correlation_context = SCN_CTX_NONE;
    else return set_error(SCN_EJSON, "Invalid correlation context: '%s'", ctx.c_str());
    std::string im = "gp_discrete", nm = "conditioned_gaussian";
    getS(root, "intersect_method", im); getS(root, "normal_method", nm);
    if (im == "mean") return set_error(SCN_EUNSUPPORTED, "intersect_method \"mean\" (sphere tracing of the mean) is not the sparse-convolution path");
    if (im != "gp_discrete") return set_error(SCN_EJSON, "Invalid intersect method: '%s'", im.c_str());
    if (nm != "finite_differences" && nm != "conditioned_gaussian" && nm != "beckmann" && nm != "ggx")
        return set_error(SCN_EJSON, "Invalid normal sampling method: '%s'", nm.c_str());
    if (const JVal *gp = root.get("gaussian_process")) { if (int rc = parseGP(*gp, *d)) return rc; }

    getF(root, "step_size", d->step_size); getU(root, "min_step", d->min_step); getU(root, "seed", d->seed);
    getF(root, "impulse_density", d->impulse_density);
    getB(root, "single_realization", d->single_realization);
    getB(root, "isotropic_3D_sampling", d->isotropic_3d_sampling);
    getB(root, "1D_sampling", d->sampling_1d);
    std::string scheme = "uni";
    getS(root, "1D_sampling_scheme", scheme);
    if (scheme == "uni" || scheme == "UNI") d->sampling_scheme_1d = SCN_SCHEME_UNI;
    else if (scheme == "nee" || scheme == "NEE") d->sampling_scheme_1d = SCN_SCHEME_NEE;
    else if (scheme == "mis" || scheme == "MIS") d->sampling_scheme_1d = SCN_SCHEME_MIS;
    else return set_error(SCN_EJSON, "Invalid sparse conv sampling scheme: '%s'", scheme.c_str());
    getB(root, "1D_gradient_correlationXY", d->correlation_xy);
    getB(root, "surf_vol_phase_separate", d->surf_vol_phase_separate);
    getF(root, "surf_vol_phase_amp_thresh", d->surf_vol_phase_amp_thresh);
    return SCN_OK;
}

} // namespace scn
