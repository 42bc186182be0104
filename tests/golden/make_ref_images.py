"""Generates tests/golden/render_*.npz: images rendered by the UNMODIFIED reference renderer (Scene::load -> TraceableScene ->
PathTraceIntegrator, driven by oracle/ref_harness.cpp: ref_render) for the SURVEY.md Appendix-A scenes, at a size the CPU finishes
in minutes.  Two renders with different sampler seeds per scene: their mutual relative MSE is the noise floor the GPU image is
judged against.  Build container only (needs oracle/_ref):  python tests/golden/make_ref_images.py [scene ...]
"scene_bvh": false makes TraceableScene intersect its primitives in a plain loop, so embree (not built here) is never entered."""
import ctypes as C
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import support as S

SCENES = {  # name -> (medium, resolution, spp[, quad light])
    "c1": ("c1_sphere_3d_none", 96, 32),
    "c3": ("c3_1d_renewal", 64, 32),
    "c5": ("c5_multires_nonstat_renewalplus", 64, 16),
    "c4": ("c4", 64, 128, True),           # Appendix A.3: conductor (Cu) phase bsdf, NEE + MIS through neePDF / neeGrad, quad light
    "c4_nee": ("c4_nee", 64, 128, True),   # the same with "1D_sampling_scheme": "nee" (light sampling only, no MIS)
    "c3_quad": ("c3_1d_renewal", 64, 64, True),  # Lambert phase bsdf under the quad light + environment: light choice, Quad::sampleDirect, MIS
}
SEEDS = (0xBA5EBA11, 12345)


def main():
    names = sys.argv[1:] or list(SCENES)
    L = C.CDLL(S.REF_LIB)
    L.ref_last_error.restype = C.c_char_p
    info = open(os.path.join(S.ORACLE_DIR, "_ref", "BUILDINFO.txt")).read().strip()
    threads = int(os.environ.get("REF_THREADS", os.cpu_count() or 1))
    for name in names:
        med, res, spp = SCENES[name][:3]
        quad = len(SCENES[name]) > 3 and SCENES[name][3]
        if med.startswith("c4"):
            mj = S.c4_scene_medium()
            if med == "c4_nee":
                mj["1D_sampling_scheme"] = "nee"
        else:
            mj = S.MEDIA[med]
        sc = S.appendix_a_scene(mj, res, spp, quad=quad)
        imgs, secs = [], []
        for seed in SEEDS:
            with tempfile.TemporaryDirectory() as td:
                path = os.path.join(td, "scene.json")
                json.dump(sc, open(path, "w"))
                w, h, sec = C.c_int(), C.c_int(), C.c_double()
                rgb = np.zeros((res, res, 3), np.float32)
                rc = L.ref_render(path.encode(), threads, C.c_uint32(seed), C.byref(w), C.byref(h), rgb.ctypes.data_as(C.c_void_p), C.c_size_t(rgb.size), C.byref(sec))
                if rc != 0:
                    raise RuntimeError(L.ref_last_error().decode())
            imgs.append(rgb.copy() if quad else rgb[:, :, 0].copy()); secs.append(sec.value)  # grey scenes keep one channel
            print(name, hex(seed), "%.1f s" % sec.value, "mean %.4f" % rgb.mean(), flush=True)
        np.savez_compressed(os.path.join(S.GOLDEN_DIR, "render_%s.npz" % name), scene_json=json.dumps(sc), images=np.stack(imgs), seeds=np.array(SEEDS, np.uint32),
                            seconds=np.array(secs), threads=threads, buildinfo=info)


if __name__ == "__main__":
    main()
