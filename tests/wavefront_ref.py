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
