"""Helpers shared by the golden-vector tests (CPU: oracle/hostsim; GPU: the CUDA library)."""
import json
import os

import numpy as np

import support as S

TOL = 1e-4  # north_star: per-query value / gradient / distance within 1e-4 relative in FP32


def grad_tol(name):
    """Gradient tolerance of a QUERY (the point itself differs by an ulp between the implementations).  A ProceduralMean's gradient is a
    forward difference with eps = 1e-3 of a mean held in float (math/GPFunctions.cpp:279-296): the reference's own value carries
    ulp(mean) / eps of rounding noise (1e-4 ... 3e-3), re-drawn when the point moves by an ulp; at a GIVEN point it is bit-identical."""
    return 5e-3 if name.startswith("proc_mean") else TOL


def load(name):
    g = dict(np.load(os.path.join(S.GOLDEN_DIR, name + ".npz"), allow_pickle=False))
    g["medium"] = json.loads(str(g["medium_json"]))
    return g


def rays_state(g, tag):
    rays = {k: g["%s_ray_%s" % (tag, k)] for k in ("o", "d", "tmin", "tmax", "u", "pix")}
    st = {k: g["%s_st_%s" % (tag, k)] for k in ("first_scatter", "last_val", "last_grad", "bounce", "last_gp_id", "info_t")}
    want = {k: g["%s_out_%s" % (tag, k)] for k in ("ok", "exited", "t", "cont_t", "grad", "gp_id", "scheme", "coeffs", "weight")}
    want["state"] = {k: g["%s_out_st_%s" % (tag, k)] for k in ("first_scatter", "last_val", "last_grad", "bounce", "last_gp_id", "info_t")}
    return rays, st, want


def check_against_golden(impl, name, g, exact=False, allow_mismatch=0):
    """Runs every entry point of `impl` on the golden inputs of medium `name` and asserts parity with the
    reference's recorded outputs: bit-exact when `exact` (CPU oracle), else within TOL with decision flips counted."""
    tol = 0.0 if exact else TOL
    is1d = S.IS_1D[name]

    def close(a, b, what, floor=None):
        a = np.asarray(a); b = np.asarray(b)
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin), "%s/%s: finiteness differs" % (name, what)
        if exact:
            assert np.array_equal(a[fin], b[fin]), "%s/%s: not bit-identical (max abs %g)" % (name, what, np.abs(a[fin] - b[fin]).max())
        else:
            err = S.rel_err(a[fin], b[fin], floor)
            assert err.size == 0 or err.max() <= tol, "%s/%s: rel err %.3g > %g" % (name, what, err.max(), tol)

    p, d = g["p"], g["d"]
    if not is1d:
        close(impl.eval_noise(p, d, int(g["noise_seed"])), g["noise"], "noise")
    e = impl.eval(p, d, g["tt"], g["pix"], int(g["scene_seed"]), g["info_t"], g["coeffs"])
    close(e["val"], g["val"], "value"); close(e["grad"], g["grad"], "gradient")
    assert np.array_equal(e["gp_id"], g["gp_id"]), name + "/gp_id"
    close(impl.conditioning(p, d, g["tv"], g["tg"], g["pix"], int(g["scene_seed"]), g["info_t"]), g["cond"], "conditioning")
    if is1d:
        pdf, ng = impl.nee(d, g["nrm"], p, g["tt"], g["pix"], int(g["scene_seed"]), g["info_t"], g["coeffs"])
        close(pdf, g["nee_pdf"], "nee_pdf"); close(ng, g["nee_grad"], "nee_grad")
    for tag in ("pri", "sec"):
        rays, st, want = rays_state(g, tag)
        got = impl.sample_distance(rays, st, 99)
        rep = S.compare_queries(got, want, TOL)
        if exact:
            for k in ("ok", "exited", "t", "cont_t", "gp_id", "scheme", "coeffs"):
                assert np.array_equal(got[k], want[k]), "%s/%s/%s not bit-identical" % (name, tag, k)
            okm = want["ok"] != 0
            assert np.array_equal(got["grad"][okm], want["grad"][okm]), "%s/%s/grad" % (name, tag)
            for k, v in want["state"].items():
                assert np.array_equal(got["state"][k], v), "%s/%s/state.%s" % (name, tag, k)
        else:
            assert rep["decision_mismatch"] <= allow_mismatch, "%s/%s: %d decision mismatches" % (name, tag, rep["decision_mismatch"])
            assert rep["t_rel"] <= tol and rep["grad_rel"] <= grad_tol(name), "%s/%s: %r" % (name, tag, rep)
            same = (got["ok"] == want["ok"]) & (got["exited"] == want["exited"])
            close(got["coeffs"][same], want["coeffs"][same], tag + "/coeffs")
            close(got["state"]["last_val"][same], want["state"]["last_val"][same], tag + "/last_val", 1e-3)
            assert np.array_equal(got["state"]["bounce"][same], want["state"]["bounce"][same])
            assert np.array_equal(got["scheme"][same], want["scheme"][same])
            assert np.array_equal(got["gp_id"][same], want["gp_id"][same])
            close(got["state"]["info_t"][same], want["state"]["info_t"][same], tag + "/info_t", 1e-3)
        T = impl.transmittance(rays, st, 99)
        assert int((T != g["%s_T" % tag]).sum()) <= allow_mismatch, "%s/%s: transmittance mismatches" % (name, tag)
