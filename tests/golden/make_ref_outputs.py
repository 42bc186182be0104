"""Lets the UNMODIFIED reference write its own result files (Integrator::saveOutputs -> writeBuffers, integrators/Integrator.cpp:40-64:
out.png through lodepng, out.pfm) for a small RGB render, and commits them next to the linear framebuffer the harness read back
(tests/golden/ref_outputs/): the fixture tests/test_outputs.py diffs render.write_pfm / write_png against.  Build container only:
    python tests/golden/make_ref_outputs.py"""
import ctypes as C
import json
import os
import shutil
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import support as S

OUT = os.path.join(S.GOLDEN_DIR, "ref_outputs")


def main():
    os.makedirs(OUT, exist_ok=True)
    L = C.CDLL(S.REF_LIB)
    L.ref_last_error.restype = C.c_char_p
    for tag, tm in (("gamma", "gamma"), ("filmic", "filmic")):
        sc = S.appendix_a_scene(S.c4_scene_medium(), 24, 4, quad=True)  # copper conductor + quad light: a coloured (RGB) image
        sc["camera"]["resolution"] = [28, 20]
        sc["camera"]["tonemap"] = tm
        sc["renderer"].update({"output_file": "out.png", "hdr_output_file": "out.pfm", "overwrite_output_files": True})
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, "scene.json")
            json.dump(sc, open(path, "w"))
            rgb = np.zeros((20, 28, 3), np.float32)
            w, h, sec = C.c_int(), C.c_int(), C.c_double()
            L.ref_render_save_outputs(1)
            rc = L.ref_render(path.encode(), 4, C.c_uint32(7), C.byref(w), C.byref(h), rgb.ctypes.data_as(C.c_void_p), C.c_size_t(rgb.size), C.byref(sec))
            L.ref_render_save_outputs(0)
            if rc != 0:
                raise RuntimeError(L.ref_last_error().decode())
            shutil.copy(os.path.join(td, "out.png"), os.path.join(OUT, "c4_%s.png" % tag))
            shutil.copy(os.path.join(td, "out.pfm"), os.path.join(OUT, "c4_%s.pfm" % tag))
        np.save(os.path.join(OUT, "c4_%s_linear.npy" % tag), rgb)
        json.dump(sc, open(os.path.join(OUT, "c4_%s_scene.json" % tag), "w"))
        print("wrote", tag, rgb.mean(axis=(0, 1)))


if __name__ == "__main__":
    main()
