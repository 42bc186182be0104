"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libscn_ref.so, built by `make -C oracle ref`
from /root/reference with g++ -O2 -msse4.2).  Run in the build container only:  python tests/golden/make_golden.py
Every file stores the seeded inputs next to the reference's outputs, so the GPU box needs neither the reference
nor this script.  The reference has no tests of its own: these vectors are what pins the oracle and the kernels."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import support as S

NE, NR = 192, 96


def main():
    info = open(os.path.join(S.ORACLE_DIR, "_ref", "BUILDINFO.txt")).read().strip()
    if "--only-missing" in sys.argv:
        return media(info)
    prim = S.CpuImpl("ref", S.MEDIA["c1_sphere_3d_none"])
    L = prim.lib
    L.ref_xxhash32_1.restype = L.ref_xxhash32_2.restype = L.ref_xxhash32_3.restype = L.ref_xxhash32_4.restype = __import__("ctypes").c_uint32
    import ctypes as C
    rng = np.random.default_rng(42)
    keys = rng.integers(0, 2**32, (256, 4), dtype=np.uint64).astype(np.uint32)
    keys[:8] = [[0, 0, 0, 0], [1, 0, 0, 0], [0xFFFFFFFF] * 4, [5, 9, 0, 0], [3, 2, 1, 7], [0xFFFFFFFF, 0, 1, 2], [7, 0xFFFFFFFE, 3, 1], [2, 2, 2, 2]]
    h1 = np.array([L.ref_xxhash32_1(C.c_uint32(int(k[0]))) for k in keys], np.uint32)
    h2 = np.array([L.ref_xxhash32_2(C.c_uint32(int(k[0])), C.c_uint32(int(k[1]))) for k in keys], np.uint32)
    h3 = np.array([L.ref_xxhash32_3(*[C.c_uint32(int(x)) for x in k[:3]]) for k in keys], np.uint32)
    h4 = np.array([L.ref_xxhash32_4(*[C.c_uint32(int(x)) for x in k]) for k in keys], np.uint32)
    seeds = np.concatenate([[0, 1, 2393193693, 0xFFFFFFFF], rng.integers(0, 2**32, 60, dtype=np.uint64)]).astype(np.uint32)
    streams = np.zeros((len(seeds), 16), np.float32); n3 = np.zeros((len(seeds), 4), np.float32); bm = np.zeros((len(seeds), 8), np.float64)
    for i, s in enumerate(seeds):
        L.ref_pcg_stream(C.c_uint64(int(s)), 16, streams[i].ctypes.data_as(C.c_void_p))
        L.ref_pcg_next3d(C.c_uint64(int(s)), n3[i].ctypes.data_as(C.c_void_p))
        L.ref_box_muller(C.c_uint64(int(s)), 4, bm[i].ctypes.data_as(C.c_void_p))
    lv = np.zeros((41, 2), np.float64)
    for i, k in enumerate(range(-20, 21)):
        sc = C.c_float(); ad = C.c_int(); L.ref_multires_level(k, C.byref(sc), C.byref(ad)); lv[i] = (sc.value, ad.value)
    np.savez_compressed(os.path.join(S.GOLDEN_DIR, "primitives.npz"), keys=keys, h1=h1, h2=h2, h3=h3, h4=h4, seeds=seeds, streams=streams, next3d=n3,
                        box_muller=bm, levels=lv, buildinfo=info)
    prim.close()
    media(info)


def media(info):
    for name, mj in S.MEDIA.items():
        if "--only-missing" in sys.argv and os.path.exists(os.path.join(S.GOLDEN_DIR, name + ".npz")):
            continue
        ref = S.CpuImpl("ref", mj)
        g = {"medium_json": json.dumps(mj), "buildinfo": info}
        p, d = S.eval_points(NE, lo=-1.5, hi=1.5)
        rng = np.random.default_rng(5)
        pix = np.random.default_rng(3).integers(0, 400, (NE, 4)).astype(np.uint32)
        coeffs = np.concatenate([rng.normal(0, 0.3, (NE, 1)), rng.normal(0, 0.3, (NE, 3)), p + rng.normal(0, 0.05, (NE, 3))], 1).astype(np.float32)
        tt = rng.uniform(0, 0.5, NE).astype(np.float32); tt[::3] = 0
        info_t = rng.uniform(0, 2, NE).astype(np.float32)
        tv = rng.normal(0, 0.02, NE).astype(np.float32); tg = rng.normal(0, 1, (NE, 3)).astype(np.float32)
        nrm = -d + rng.normal(0, 0.4, (NE, 3)).astype(np.float32); nrm = np.ascontiguousarray((nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32))
        g.update(p=p, d=d, pix=pix, coeffs=coeffs, tt=tt, info_t=info_t, tv=tv, tg=tg, nrm=nrm, noise_seed=11, scene_seed=77)
        if not S.IS_1D[name]:
            g["noise"] = ref.eval_noise(p, d, 11)
        e = ref.eval(p, d, tt, pix, 77, info_t, coeffs)
        g.update(val=e["val"], grad=e["grad"], gp_id=e["gp_id"])
        g["cond"] = ref.conditioning(p, d, tv, tg, pix, 77, info_t)
        if S.IS_1D[name]:
            g["nee_pdf"], g["nee_grad"] = ref.nee(d, nrm, p, tt, pix, 77, info_t, coeffs)
        for sec in (False, True):
            tag = "sec" if sec else "pri"
            rays, st = S.camera_rays(NR, secondary=sec)
            q = ref.sample_distance(rays, st, 99)
            for k, v in rays.items():
                g["%s_ray_%s" % (tag, k)] = v
            for k, v in st.items():
                g["%s_st_%s" % (tag, k)] = v
            for k, v in q.items():
                if k == "state":
                    for kk, vv in v.items():
                        g["%s_out_st_%s" % (tag, kk)] = vv
                else:
                    g["%s_out_%s" % (tag, k)] = v
            g["%s_T" % tag] = ref.transmittance(rays, st, 99)
        np.savez_compressed(os.path.join(S.GOLDEN_DIR, name + ".npz"), **g)
        ref.close()
        print("wrote", name)


if __name__ == "__main__":
    main()
