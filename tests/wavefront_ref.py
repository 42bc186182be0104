"""Test infrastructure: numpy restatement of the wavefront stage (csrc/scn_render.cu) over the ORACLE's sampleDistance, used by
tests/test_wavefront.py to check the CUDA stage path by path.  It restates this repo's own restricted integrator (include/
scn_render.h says what it covers of integrators/path_tracer/PathTracer.cpp:17-204), with the same counter-based sampler, so the
two must take the same decisions up to float rounding of the direction sampling."""
import numpy as np

import support as S

P2, P3, P4, P5 = np.uint32(2246822519), np.uint32(3266489917), np.uint32(668265263), np.uint32(374761393)


def _rotl17(h):
    return (h << np.uint32(17)) | (h >> np.uint32(15))


def xxhash32_4(x, y, z, w):
    """math/MathUtil.hpp:211-224 on uint32 arrays."""
    with np.errstate(over="ignore"):
        x, y, z, w = (np.asarray(a, np.uint32) for a in (x, y, z, w))
        h = w + P5 + x * P3; h = P4 * _rotl17(h)
        h = h + y * P3; h = P4 * _rotl17(h)
        h = h + z * P3; h = P4 * _rotl17(h)
        h = P2 * (h ^ (h >> np.uint32(15))); h = P3 * (h ^ (h >> np.uint32(13)))
        return h ^ (h >> np.uint32(16))


def wf_rand(px, py, spp, bounce, dim, seed):
    h = xxhash32_4(px | (py << np.uint32(16)), spp, np.uint32(bounce) * np.uint32(8) + np.uint32(dim), np.uint32(seed) ^ np.uint32(0x9e3779b9))
    return ((h >> np.uint32(9)) | np.uint32(0x3F800000)).view(np.float32) - np.float32(1.0)


def cube(sc, o, d):
    with np.errstate(divide="ignore", invalid="ignore"):
        c, h = np.float32(sc["cube_center"]), np.float32(sc["cube_half"])
        inv = np.float32(1.0) / d
        a, b = (c - h - o) * inv, (c + h - o) * inv
        par = d == 0
        inside = np.abs(o - c) <= h
        a = np.where(par, np.where(inside, -np.inf, np.inf), a); b = np.where(par, np.where(inside, np.inf, -np.inf), b)
        tn = np.maximum(0, np.minimum(a, b).max(1)).astype(np.float32); tf = np.maximum(a, b).min(1).astype(np.float32)
    return tn, tf, tf > tn


def make_frame(n):
    """math/TangentFrame.hpp:74-99 (Duff et al.), float32."""
    n = (n / np.sqrt((n * n).sum(1, keepdims=True))).astype(np.float32)
    sign = np.copysign(np.float32(1), n[:, 2]); a = np.float32(-1) / (sign + n[:, 2]); b = n[:, 0] * n[:, 1] * a
    t = np.stack([1 + sign * n[:, 0] * n[:, 0] * a, sign * b, -sign * n[:, 0]], 1).astype(np.float32)
    bt = np.stack([b, sign + n[:, 1] * n[:, 1] * a, -n[:, 1]], 1).astype(np.float32)
    t /= np.sqrt((t * t).sum(1, keepdims=True)); bt /= np.sqrt((bt * bt).sum(1, keepdims=True))
    return t, bt, n


def render(scene, medium_json, tile_first, tile_count, spp_first, spp_count, threads=8):
    """Mean radiance per local pixel ([tile_count * 256]) and per-path radiance, like WavefrontRenderer.render + local_image."""
    sc = scene
    rx, ry, seed = sc["res_x"], sc["res_y"], sc["scene_seed"]
    n = tile_count * 256 * spp_count
    i = np.arange(n)
    s = (i % spp_count).astype(np.uint32); lp = i // spp_count
    tiles_x = (rx + 15) // 16
    tile = tile_first + lp // 256; it = lp % 256
    x = ((tile % tiles_x) * 16 + it % 16).astype(np.uint32); y = ((tile // tiles_x) * 16 + it // 16).astype(np.uint32)
    spp = (spp_first + s).astype(np.uint32)
    radiance = np.zeros(n, np.float32); alive = np.ones(n, bool)
    alive &= (x < rx) & (y < ry)
    f = np.float32(sc["cam_look_at"]) - np.float32(sc["cam_pos"]); f /= np.sqrt((f * f).sum())
    r = np.cross(f, np.float32(sc["cam_up"])); r /= np.sqrt((r * r).sum()); uu = np.cross(r, f)
    pd = np.float32(1.0 / np.tan(np.float32(sc["fov_deg"]) * np.float32(0.017453292519943295) * np.float32(0.5)))
    ratio, pix = np.float32(ry) / np.float32(rx), np.float32(1.0) / np.float32(rx)
    jx = wf_rand(x, y, spp, 0, 6, seed) - np.float32(0.5); jy = wf_rand(x, y, spp, 0, 7, seed) - np.float32(0.5)
    ld = np.stack([-1 + (x.astype(np.float32) + np.float32(0.5) + jx) * 2 * pix, ratio - (y.astype(np.float32) + np.float32(0.5) + jy) * 2 * pix, np.full(n, pd, np.float32)], 1).astype(np.float32)
    ld /= np.sqrt((ld * ld).sum(1, keepdims=True))
    d = (ld[:, :1] * r + ld[:, 1:2] * uu + ld[:, 2:3] * f).astype(np.float32)
    o = np.tile(np.float32(sc["cam_pos"]), (n, 1))
    tn, tf, hit = cube(sc, o, d)
    miss = alive & ~hit
    radiance[miss] = sc["emission"]; alive &= hit
    rays = {"o": o, "d": d, "tmin": np.where(alive, tn, 0).astype(np.float32), "tmax": np.where(alive, tf, 0).astype(np.float32), "u": wf_rand(x, y, spp, 0, 0, seed),
            "pix": np.ascontiguousarray(np.stack([x, y, spp, np.zeros(n, np.uint32)], 1))}
    st = {"first_scatter": np.ones(n, np.uint8), "last_val": np.zeros(n, np.float32), "last_grad": np.zeros((n, 3), np.float64), "bounce": np.zeros(n, np.int32),
          "last_gp_id": np.zeros(n, np.int32), "info_t": np.zeros(n, np.float32)}
    thr = np.ones(n, np.float32)
    orc = S.CpuImpl("orc", medium_json, threads=threads)
    queries = 0
    for bounce in range(sc["max_bounces"]):
        idx = np.nonzero(alive)[0]
        if idx.size == 0:
            break
        queries += idx.size
        sub = {k: np.ascontiguousarray(v[idx]) for k, v in rays.items()}; sst = {k: np.ascontiguousarray(v[idx]) for k, v in st.items()}
        q = orc.sample_distance(sub, sst, seed)
        for k in st:
            st[k][idx] = q["state"][k]
        ok, ex = q["ok"] != 0, q["exited"] != 0
        th = thr[idx] * q["weight"][:, 0]
        dead = ~ok
        out = ok & ex
        radiance[idx[out]] = th[out] * np.float32(sc["emission"])
        sca = ok & ~ex
        th = th * np.float32(sc["albedo"])
        px, py, sp = x[idx], y[idx], spp[idx]
        if bounce > 2:
            rr = sca & (th < np.float32(0.1))
            keep = wf_rand(px, py, sp, bounce, 3, seed) < th
            dead |= rr & ~keep; sca &= ~(rr & ~keep)
            th = np.where(rr & keep, np.float32(1.0), th)
        if bounce + 1 >= sc["max_bounces"]:
            sca[:] = False
        sca &= th > 0
        alive[idx] = sca
        j = idx[sca]
        if j.size:
            t = q["t"][sca]
            p = (rays["o"][j] + t[:, None] * rays["d"][j]).astype(np.float32)
            tt, bt, nn = make_frame(q["grad"][sca].astype(np.float32))
            u1, u2 = wf_rand(px[sca], py[sca], sp[sca], bounce, 1, seed), wf_rand(px[sca], py[sca], sp[sca], bounce, 2, seed)
            phi, rr_ = np.float32(6.283185307179586) * u1, np.sqrt(u2)
            wo = np.stack([np.cos(phi) * rr_, np.sin(phi) * rr_, np.sqrt(np.maximum(1 - u2, 0))], 1).astype(np.float32)
            nd = (wo[:, :1] * tt + wo[:, 1:2] * bt + wo[:, 2:3] * nn).astype(np.float32); nd /= np.sqrt((nd * nd).sum(1, keepdims=True))
            _, t1, inside = cube(sc, p, nd)
            esc = ~inside
            radiance[j[esc]] = th[sca][esc] * np.float32(sc["emission"]); alive[j[esc]] = False
            rays["o"][j] = p; rays["d"][j] = nd; rays["tmin"][j] = 0; rays["tmax"][j] = t1
            rays["u"][j] = wf_rand(px[sca], py[sca], sp[sca], bounce + 1, 0, seed); rays["pix"][j, 3] = bounce + 1
            thr[j] = th[sca]
    orc.close()
    img = radiance.reshape(-1, spp_count).sum(1) / spp_count
    return img, radiance, queries


# ------------------------------------------------------------------------------------------------ next-event estimation stage
# numpy restatement of k_wf_nee_setup / k_wf_nee_gather / k_wf_shade_nee (csrc/scn_render.cu), i.e. of volumeEstimateDirect
# (integrators/TraceBase.cpp:346-420, 444-518) with the Lambert / mirror / conductor phase bsdfs, over the ORACLE's sampleDistance,
# transmittance, neePDF and neeGrad.  Same counter-based streams as the CUDA stage: per-path radiance must agree up to float rounding.
F = np.float32
TWO_PI, INV_PI, INV_4PI = F(6.283185307179586), F(0.3183098861837907), F(0.07957747154594767)


def _n3(a):
    return (a / np.sqrt((a * a).sum(1, keepdims=True))).astype(np.float32)


def _dot(a, b):
    return (a * b).sum(1).astype(np.float32)


def _to_local(fr, v):
    t, b, n = fr
    return np.stack([_dot(t, v), _dot(b, v), _dot(n, v)], 1).astype(np.float32)


def _to_global(fr, v):
    t, b, n = fr
    return (v[:, :1] * t + v[:, 1:2] * b + v[:, 2:3] * n).astype(np.float32)


def _quad(sc):
    base, e0, e1 = F(sc["quad_base"]), F(sc["quad_edge0"]), F(sc["quad_edge1"])
    n = np.cross(e1, e0).astype(np.float32); area = F(np.sqrt((n * n).sum())); n = (n / area).astype(np.float32)
    return {"base": base, "e0": e0, "e1": e1, "n": n, "area": area, "iu0": F(1) / F((e0 * e0).sum()), "iu1": F(1) / F((e1 * e1).sum())}


def _quad_hit(q, o, d):
    """Quad::intersect (primitives/Quad.cpp:71-97) over [0, inf) -> (hit, t, back)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        ndw = _dot(d, q["n"][None]); ok = np.abs(ndw) >= F(1e-6)
        t = (_dot(q["n"][None], q["base"][None] - o) / ndw).astype(np.float32); ok &= t >= 0
        v = (o + t[:, None] * d - q["base"][None]).astype(np.float32)
        l0, l1 = _dot(v, q["e0"][None]) * q["iu0"], _dot(v, q["e1"][None]) * q["iu1"]
        ok &= (l0 >= 0) & (l0 <= 1) & (l1 >= 0) & (l1 <= 1)
    return ok, np.where(ok, t, 0).astype(np.float32), ndw >= 0


def _quad_approx(q, p, emax):
    """Quad::approximateRadiance (primitives/Quad.cpp:256-279)."""
    R0 = (q["base"][None] - p).astype(np.float32); R1 = R0 + q["e0"][None]; R2 = R1 + q["e1"][None]; R3 = R0 + q["e1"][None]
    n0, n1, n2, n3 = _n3(np.cross(R0, R1)), _n3(np.cross(R1, R2)), _n3(np.cross(R2, R3)), _n3(np.cross(R3, R0))
    with np.errstate(invalid="ignore"):
        Q = np.arccos(_dot(n0, n1)) + np.arccos(_dot(n1, n2)) + np.arccos(_dot(n2, n3)) + np.arccos(_dot(n3, n0))
    return np.where(_dot(R0, q["n"][None]) >= 0, F(0), (TWO_PI - np.abs(Q)) * emax).astype(np.float32)


def _conductor(eta, k, c):
    """Fresnel::conductorReflectance (bsdfs/Fresnel.hpp:102-123) for one channel."""
    if eta == 0 and k == 0:
        return np.ones_like(c)
    eta, k = F(eta), F(k)
    c2 = c * c; s2 = np.maximum(F(1) - c2, F(0)); s4 = s2 * s2
    inner = eta * eta - k * k - s2
    ab = np.sqrt(np.maximum(inner * inner + F(4) * eta * eta * k * k, F(0))); a = np.sqrt(np.maximum((ab + inner) * F(0.5), F(0)))
    Rs = ((ab + c2) - (F(2) * a * c)) / ((ab + c2) + (F(2) * a * c))
    Rp = ((c2 * ab + s4) - (F(2) * a * c * s2)) / ((c2 * ab + s4) + (F(2) * a * c * s2))
    return (F(0.5) * (Rs + Rs * Rp)).astype(np.float32)


def _spec_weight(sc, cos_i):
    w = np.tile(F(sc["albedo_rgb"])[None], (len(cos_i), 1))
    if sc["phase_type"] == 2:
        w = w * np.stack([_conductor(sc["eta"][j], sc["k"][j], cos_i) for j in range(3)], 1)
    return w.astype(np.float32)


def _power(a, b):
    with np.errstate(invalid="ignore", divide="ignore", over="ignore"):
        return ((a * a) / (a * a + b * b)).astype(np.float32)


UNI, NEE, MIS = 0, 1, 2


def render_nee(scene, medium_json, tile_first, tile_count, spp_first, spp_count, threads=8):
    """Per-path RGB radiance ([n, 3]) of the nee stage, like WavefrontRenderer.render with scene.nee (after the sample clamp)."""
    sc = scene
    rx, ry, seed, nb_max = sc["res_x"], sc["res_y"], sc["scene_seed"], sc["max_bounces"]
    n = tile_count * 256 * spp_count
    i = np.arange(n); s = (i % spp_count).astype(np.uint32); lp = i // spp_count
    tiles_x = (rx + 15) // 16; tile = tile_first + lp // 256; it = lp % 256
    x = ((tile % tiles_x) * 16 + it % 16).astype(np.uint32); y = ((tile // tiles_x) * 16 + it // 16).astype(np.uint32); spp = (spp_first + s).astype(np.uint32)
    rad = np.zeros((n, 3), np.float32); thr = np.ones((n, 3), np.float32); alive = (x < rx) & (y < ry); scattered = np.zeros(n, bool)
    f = F(sc["cam_look_at"]) - F(sc["cam_pos"]); f /= np.sqrt((f * f).sum()); r = np.cross(f, F(sc["cam_up"])); r /= np.sqrt((r * r).sum()); uu = np.cross(r, f)
    pd = F(1.0 / np.tan(F(sc["fov_deg"]) * F(0.017453292519943295) * F(0.5))); ratio, pix = F(ry) / F(rx), F(1.0) / F(rx)
    jx = wf_rand(x, y, spp, 0, 6, seed) - F(0.5); jy = wf_rand(x, y, spp, 0, 7, seed) - F(0.5)
    ld = _n3(np.stack([-1 + (x.astype(np.float32) + F(0.5) + jx) * 2 * pix, ratio - (y.astype(np.float32) + F(0.5) + jy) * 2 * pix, np.full(n, pd, np.float32)], 1).astype(np.float32))
    d = (ld[:, :1] * r + ld[:, 1:2] * uu + ld[:, 2:3] * f).astype(np.float32); o = np.tile(F(sc["cam_pos"]), (n, 1))
    tn, tf, hit = cube(sc, o, d)
    quad = _quad(sc) if sc["has_quad"] else None
    env, qrgb = F(sc["env_rgb"]), F(sc["quad_rgb"])
    if quad is not None:  # an unscattered path sees the emitters directly
        hq, tq, back = _quad_hit(quad, o, d); first = alive & hq & (~hit | (tq < tn))
        rad[first & ~back] += qrgb; alive &= ~first
    miss = alive & ~hit; rad[miss] += env; alive &= hit
    rays = {"o": o, "d": d, "tmin": np.where(alive, tn, 0).astype(np.float32), "tmax": np.where(alive, tf, 0).astype(np.float32), "u": wf_rand(x, y, spp, 0, 0, seed),
            "pix": np.ascontiguousarray(np.stack([x, y, spp, np.zeros(n, np.uint32)], 1))}
    st = {"first_scatter": np.ones(n, np.uint8), "last_val": np.zeros(n, np.float32), "last_grad": np.zeros((n, 3), np.float64), "bounce": np.zeros(n, np.int32),
          "last_gp_id": np.zeros(n, np.int32), "info_t": np.zeros(n, np.float32)}
    orc = S.CpuImpl("orc", medium_json, threads=threads)
    specular = sc["phase_type"] != 0
    env_max, quad_max = F(max(sc["env_rgb"])), F(max(sc["quad_rgb"]))
    env_light, quad_light = bool(sc["env_sample"]) and env_max > 0, bool(sc["has_quad"]) and quad_max > 0
    sub_of = lambda dct, idx: {k: np.ascontiguousarray(v[idx]) for k, v in dct.items()}
    for bounce in range(nb_max):
        idx = np.nonzero(alive)[0]
        if idx.size == 0:
            break
        q = orc.sample_distance(sub_of(rays, idx), sub_of(st, idx), seed)
        for k in st:
            st[k][idx] = q["state"][k]
        ok, ex = q["ok"] != 0, q["exited"] != 0
        wq = q["weight"][:, 0].astype(np.float32)
        th = (thr[idx] * wq[:, None]).astype(np.float32)
        t = q["t"].astype(np.float32); dirs = rays["d"][idx]; p = (rays["o"][idx] + t[:, None] * dirs).astype(np.float32)
        px, py, sp = x[idx], y[idx], spp[idx]
        sca = ok & ~ex
        if bounce < nb_max - 1 and sca.any():  # ---- volumeEstimateDirect for the paths that scattered at this bounce
            m = np.nonzero(sca)[0]; nm = m.size; nbk = bounce | 0x8000
            P, D, T_, TH = p[m], dirs[m], t[m], th[m]
            g = q["grad"][m].astype(np.float32); fr = make_frame(_n3(g)); nrm = fr[2]
            wi = _n3(_to_local(fr, -D)); sch = q["scheme"][m].astype(int)
            pixm = np.ascontiguousarray(np.stack([px[m], py[m], sp[m], rays["pix"][idx[m], 3]], 1).astype(np.uint32))
            info_t = st["info_t"][idx[m]]; coeffs = np.ascontiguousarray(q["coeffs"][m].astype(np.float32))
            aniso = st["last_grad"][idx[m]].astype(np.float32)
            act = np.ones(nm, bool)
            if env_light and quad_light:
                pe = np.full(nm, TWO_PI * env_max, np.float32); pq = _quad_approx(quad, P, quad_max); total = pe + pq
                act &= total != 0
                tsel = wf_rand(px[m], py[m], sp[m], nbk, 0, seed) * total
                firstp = pq if sc["quad_first"] else pe
                use_quad = np.where(tsel < firstp, bool(sc["quad_first"]), not sc["quad_first"])
                with np.errstate(divide="ignore", invalid="ignore"):
                    lw = (total / np.where(use_quad, pq, pe)).astype(np.float32)
            elif env_light or quad_light:
                use_quad = np.full(nm, quad_light); lw = np.ones(nm, np.float32)
            else:
                act[:] = False; use_quad = np.zeros(nm, bool); lw = np.ones(nm, np.float32)
            Le = np.where(use_quad[:, None], qrgb[None], env[None]).astype(np.float32)
            if quad is not None:
                def vis_far(dd):  # attenuatedEmission: the chosen light must be the first emitter along the ray
                    hq_, tq_, back_ = _quad_hit(quad, P, dd)
                    return hq_, tq_, back_
            shadow = []
            # ---- volumeLightSample
            u1, u2 = wf_rand(px[m], py[m], sp[m], nbk, 1, seed), wf_rand(px[m], py[m], sp[m], nbk, 2, seed)
            okl = act & ~(specular & (sch == UNI))
            dl = np.zeros((nm, 3), np.float32); dist = np.full(nm, np.inf, np.float32); pdf_l = np.zeros(nm, np.float32)
            if quad is not None:
                uq = use_quad
                front = _dot(quad["n"][None], P - quad["base"][None]) > 0
                qp = (quad["base"][None] + u1[:, None] * quad["e0"][None] + u2[:, None] * quad["e1"][None]).astype(np.float32)
                dq = (qp - P).astype(np.float32); r2 = _dot(dq, dq); dsq = np.sqrt(r2).astype(np.float32)
                with np.errstate(divide="ignore", invalid="ignore"):
                    dqn = (dq / dsq[:, None]).astype(np.float32); pq_l = (r2 / (-_dot(quad["n"][None], dqn) * quad["area"])).astype(np.float32)
                dl[uq] = dqn[uq]; dist[uq] = dsq[uq]; pdf_l[uq] = pq_l[uq]; okl &= ~uq | front
            ue = ~use_quad
            phi, z = u1 * TWO_PI, u2 * F(2) - F(1); rr = np.sqrt(np.maximum(F(1) - z * z, F(0)))
            de = np.stack([np.cos(phi) * rr, np.sin(phi) * rr, z], 1).astype(np.float32)
            dl[ue] = de[ue]; pdf_l[ue] = INV_4PI
            wo = _n3(_to_local(fr, np.where(okl[:, None], dl, F([0, 0, 1])[None]).astype(np.float32)))
            fL = np.zeros((nm, 3), np.float32); ppdf = np.zeros(nm, np.float32); an = aniso.copy()
            if not specular:
                lam = (wi[:, 2] > 0) & (wo[:, 2] > 0); kk = INV_PI * wo[:, 2]
                fL[lam] = (F(sc["albedo_rgb"])[None] * kk[:, None])[lam]; ppdf[lam] = (np.abs(wo[:, 2]) * INV_PI)[lam]
            else:
                nw = _n3(_to_global(fr, ((wi + wo) * F(0.5)).astype(np.float32)))
                pn, ng = orc.nee(np.ascontiguousarray(D), np.ascontiguousarray(nw), np.ascontiguousarray(P), np.ascontiguousarray(T_), pixm, seed, info_t, coeffs)
                fL = (_spec_weight(sc, wi[:, 2]) * pn[:, None]).astype(np.float32); ppdf = pn.astype(np.float32); an = ng.astype(np.float32)
            okl &= ~((fL == 0).all(1))
            vis = np.ones(nm, bool); far = np.full(nm, np.inf, np.float32); e = Le.copy()
            if quad is not None:
                hq_, tq_, back_ = _quad_hit(quad, P, dl)
                uqm = use_quad
                vis &= ~uqm | (hq_ & ~(tq_ * (F(1) + F(1e-3)) < dist)); far = np.where(uqm & hq_, tq_, far); e[uqm & back_] = 0
                vis &= uqm | ~hq_
            okl &= vis & ~((e == 0).all(1))
            mis = np.where(specular & (sch == NEE), F(1), _power(pdf_l, ppdf)).astype(np.float32)
            with np.errstate(divide="ignore", invalid="ignore"):
                kL = (lw * mis / pdf_l).astype(np.float32)
            shadow.append((okl, dl, far, an, wf_rand(px[m], py[m], sp[m], nbk, 3, seed), (TH * fL * e * kL[:, None]).astype(np.float32)))
            # ---- volumePhaseSample
            oks = act & ~(specular & (sch == NEE))
            if not specular:
                oks &= wi[:, 2] > 0
                v1, v2 = wf_rand(px[m], py[m], sp[m], nbk, 5, seed), wf_rand(px[m], py[m], sp[m], nbk, 6, seed)
                ph, rr2 = TWO_PI * v1, np.sqrt(v2)
                wo2 = np.stack([np.cos(ph) * rr2, np.sin(ph) * rr2, np.sqrt(np.maximum(F(1) - v2, F(0)))], 1).astype(np.float32)
                w2 = np.tile(F(sc["albedo_rgb"])[None], (nm, 1)); pdf_p = (np.abs(wo2[:, 2]) * INV_PI).astype(np.float32)
            else:
                wo2 = np.stack([-wi[:, 0], -wi[:, 1], wi[:, 2]], 1).astype(np.float32); w2 = _spec_weight(sc, wi[:, 2]); pdf_p = np.ones(nm, np.float32)
                pn2, _ = orc.nee(np.ascontiguousarray(D), np.ascontiguousarray(nrm), np.ascontiguousarray(P), np.ascontiguousarray(T_), pixm, seed, info_t, coeffs)
                pdf_p = np.where(sch != UNI, pn2, pdf_p).astype(np.float32)
            d2 = _n3(_to_global(fr, wo2))
            vis = np.ones(nm, bool); far = np.full(nm, np.inf, np.float32); dpdf = np.full(nm, INV_4PI, np.float32); e = Le.copy()
            if quad is not None:
                hq_, tq_, back_ = _quad_hit(quad, P, d2)
                uqm = use_quad
                vis &= ~uqm | hq_; far = np.where(uqm & hq_, tq_, far); e[uqm & hq_ & back_] = 0
                with np.errstate(divide="ignore", invalid="ignore"):
                    ndw = _dot(quad["n"][None], d2); tt = _dot(quad["n"][None], quad["base"][None] - P) / ndw
                    dpdf = np.where(uqm & hq_, tt * tt / (np.abs(ndw) * quad["area"]), dpdf).astype(np.float32)
                vis &= uqm | ~hq_
            oks &= vis & ~((e == 0).all(1))
            mis2 = np.where((not specular) | (sch != UNI), _power(pdf_p, dpdf), F(1)).astype(np.float32)
            shadow.append((oks, d2, far, aniso, wf_rand(px[m], py[m], sp[m], nbk, 4, seed), (TH * w2 * e * (lw * mis2)[:, None]).astype(np.float32)))
            # ---- transmittance along both shadow rays (MediumState copy, segment index + 1), gather
            for okh, dh, farh, anh, uh, contrib in shadow:
                k = np.nonzero(okh)[0]
                if k.size == 0:
                    continue
                _, t1, inside = cube(sc, P[k], dh[k])
                tmax = np.where(inside, np.minimum(t1, farh[k]), 0).astype(np.float32)
                gi = idx[m[k]]
                srays = {"o": np.ascontiguousarray(P[k]), "d": np.ascontiguousarray(dh[k]), "tmin": np.zeros(k.size, np.float32), "tmax": tmax, "u": np.ascontiguousarray(uh[k]),
                         "pix": np.ascontiguousarray(np.stack([px[m[k]], py[m[k]], sp[m[k]], rays["pix"][gi, 3] + 1], 1).astype(np.uint32))}
                sst = {"first_scatter": st["first_scatter"][gi].copy(), "last_val": st["last_val"][gi].copy(), "last_grad": np.ascontiguousarray(anh[k].astype(np.float64)),
                       "bounce": st["bounce"][gi].copy(), "last_gp_id": st["last_gp_id"][gi].copy(), "info_t": st["info_t"][gi].copy()}
                T = orc.transmittance(srays, sst, seed)
                rad[gi] += contrib[k] * T[:, None]
        # ---- k_wf_shade_nee
        dead = ~ok
        out = ok & ex
        if out.any():
            e = np.zeros((idx.size, 3), np.float32)
            unsc = out & ~scattered[idx]
            e[unsc] = env
            if quad is not None:
                hq_, tq_, back_ = _quad_hit(quad, p, dirs)
                e[unsc & hq_] = qrgb; e[unsc & hq_ & back_] = 0
            rad[idx[out]] += (th * e)[out]
        scattered[idx[sca]] = True
        fr_all = make_frame(_n3(np.where(sca[:, None], q["grad"].astype(np.float32), F([0, 0, 1])[None]).astype(np.float32)))
        wi_all = _n3(_to_local(fr_all, -dirs))
        if not specular:
            sca &= wi_all[:, 2] > 0
            u1, u2 = wf_rand(px, py, sp, bounce, 1, seed), wf_rand(px, py, sp, bounce, 2, seed)
            phi, rr_ = TWO_PI * u1, np.sqrt(u2)
            wo = np.stack([np.cos(phi) * rr_, np.sin(phi) * rr_, np.sqrt(np.maximum(F(1) - u2, F(0)))], 1).astype(np.float32)
            w = np.tile(F(sc["albedo_rgb"])[None], (idx.size, 1))
        else:
            wo = np.stack([-wi_all[:, 0], -wi_all[:, 1], wi_all[:, 2]], 1).astype(np.float32); w = _spec_weight(sc, wi_all[:, 2])
        th = (th * w).astype(np.float32)
        sca &= th.max(1) != 0
        rp = np.abs(th).max(1)
        if bounce > 2:
            rr = sca & (rp < F(0.1))
            keep = wf_rand(px, py, sp, bounce, 3, seed) < rp
            sca &= ~(rr & ~keep)
            with np.errstate(divide="ignore", invalid="ignore"):
                th = np.where((rr & keep)[:, None], th / rp[:, None], th).astype(np.float32)
        if bounce + 1 >= nb_max:
            sca[:] = False
        nd = _n3(_to_global(fr_all, wo))
        _, t1, inside = cube(sc, p, nd)
        sca &= inside
        alive[idx] = sca
        j = idx[sca]
        rays["o"][j] = p[sca]; rays["d"][j] = nd[sca]; rays["tmin"][j] = 0; rays["tmax"][j] = t1[sca]
        rays["u"][j] = wf_rand(px[sca], py[sca], sp[sca], bounce + 1, 0, seed); rays["pix"][j, 3] = bounce + 1
        thr[j] = th[sca]
    orc.close()
    # PathTraceIntegrator::renderTile (PathTraceIntegrator.cpp:148-156) / k_wf_clamp_rgb: a sample with a NaN, an infinite component or a
    # component above 65536 becomes black
    with np.errstate(invalid="ignore"):
        bad = (~np.isfinite(rad)).any(1) | (rad > 65536).any(1)
    rad[bad] = 0
    return rad
