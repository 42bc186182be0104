"""Developer tool (test infrastructure): compare oracle / hostsim against the compiled reference on every medium
in tests/support.py and print a parity table.  Usage: python tools/compare_cpu.py [orc|sim] [media-substring]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import support as S

kind = sys.argv[1] if len(sys.argv) > 1 else "orc"
flt = sys.argv[2] if len(sys.argv) > 2 else ""
NE, NR = 600, 300
for name, mj in S.MEDIA.items():
    if flt not in name:
        continue
    try:
        ref = S.CpuImpl("ref", mj); tst = S.CpuImpl(kind, mj)
    except Exception as e:
        print(name, "SKIP:", e); continue
    p, d = S.eval_points(NE, lo=-1.5, hi=1.5)
    pix = np.random.default_rng(3).integers(0, 400, (NE, 4)).astype(np.uint32)
    row = [name]
    if not S.IS_1D[name]:
        a, b = tst.eval_noise(p, d, 11), ref.eval_noise(p, d, 11)
        row.append("noise %.1e%s" % (S.rel_err(a, b).max(), "=" if np.array_equal(a, b) else ""))
    # realization-level eval with random conditioning coefficients
    rng = np.random.default_rng(5)
    coeffs = np.concatenate([rng.normal(0, 0.3, (NE, 1)), rng.normal(0, 0.3, (NE, 3)), p + rng.normal(0, 0.05, (NE, 3))], 1).astype(np.float32)
    tt = rng.uniform(0, 0.5, NE).astype(np.float32); tt[::3] = 0
    info_t = rng.uniform(0, 2, NE).astype(np.float32)
    ea = tst.eval(p, d, tt, pix, 77, info_t, coeffs); eb = ref.eval(p, d, tt, pix, 77, info_t, coeffs)
    row.append("val %.1e grad %.1e id %d" % (S.rel_err(ea["val"], eb["val"]).max(), S.rel_err(ea["grad"], eb["grad"]).max(), int((ea["gp_id"] != eb["gp_id"]).sum())))
    tv = rng.normal(0, 0.02, NE).astype(np.float32); tg = rng.normal(0, 1, (NE, 3)).astype(np.float32)
    ca, cb = tst.conditioning(p, d, tv, tg, pix, 77, info_t), ref.conditioning(p, d, tv, tg, pix, 77, info_t)
    row.append("cond %.1e" % S.rel_err(ca, cb).max())
    for sec in (False, True):
        rays, st = S.camera_rays(NR, secondary=sec)
        qa, qb = tst.sample_distance(rays, st, 99), ref.sample_distance(rays, st, 99)
        rep = S.compare_queries(qa, qb)
        cm = float(S.rel_err(qa["coeffs"], qb["coeffs"]).max())
        lv = float(S.rel_err(qa["state"]["last_val"], qb["state"]["last_val"], 1e-3).max())
        row.append("%s mism %d t %.1e g %.1e c %.1e lv %.1e hit %.2f ok %.2f" % ("sec" if sec else "pri", rep["decision_mismatch"], rep["t_rel"], rep["grad_rel"], cm, lv, rep["hit_fraction"], rep["ok_fraction"]))
        ta, tb = tst.transmittance(rays, st, 99), ref.transmittance(rays, st, 99)
        row.append("T mism %d" % int((ta != tb).sum()))
    if S.IS_1D[name]:
        nrm = -d + rng.normal(0, 0.4, (NE, 3)).astype(np.float32); nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
        pa, ga = tst.nee(d, nrm, p, tt, pix, 77, info_t, coeffs); pb, gb = ref.nee(d, nrm, p, tt, pix, 77, info_t, coeffs)
        ok = np.isfinite(pb) & np.isfinite(gb).all(1)
        row.append("nee pdf %.1e grad %.1e (finite %d)" % (S.rel_err(pa[ok], pb[ok]).max(), S.rel_err(ga[ok], gb[ok]).max(), ok.sum()))
    print(" | ".join(row), flush=True)
