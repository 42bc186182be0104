"""Developer tool (GPU): the checks compute-sanitizer would make on the cached march kernels, written out (the tool itself is closed on
this pool).  Builds the library a second time with -DSCN_DEBUG_BOUNDS (index / capacity assertions counted on the device, shared-memory
caches poisoned with NaN at kernel start: scn_device.cuh SCN_CHECK / SCN_SMEM_FILL), runs wavefronts of every tuned kernel family through
it and through the product library, and reports (1) the eight assertion counters, which must stay 0, and (2) that every output is
bit-identical to the product library's — a read of a never-written cache slot would surface as a NaN or a changed value.

  python tools/debug_bounds.py [log2 rays, default 16] > gpurun_out/debug_bounds.txt
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
from sparse_conv_gpis_tungsten_b200 import abi, build

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 16)
MEDIA = ["c1_sphere_3d_none", "c5_multires_nonstat_renewalplus", "nonstat_bruteforce_iso", "world_3d_renewalplus", "multires_world", "aniso_mtx_iso_renewalplus",
         "grid_ls_multires_iso", "c3_1d_renewal", "c4_1d_renewalplus_xy_mis", "multires_1d_xy", "nonstat_bruteforce_1d_renewalplus",
         # the kernel variants added late in round 2: full-matrix kernels (MODE 3 / 4), Matern / Gabor (MODE 5), procedural means, 1D with the matrix
         "aniso_mtx_world_renewalplus", "aniso_rot_world_bruteforce", "aniso_rot_iso_multires", "aniso_rot_iso_const_none", "matern15_world", "gabor_aniso_world",
         "proc_mean_knob_iso", "proc_mean_plane_world", "proc_mean_knob_inner_1d", "aniso_mtx_1d_xy_mis"]


def worker(lib_path, out_path):
    abi.LIB_PATH = lib_path
    import support as S
    res = {}
    for name in MEDIA:
        g = S.GpuImpl(S.MEDIA[name])
        for sec in (False, True):
            rays, st = S.camera_rays(n, seed=31, secondary=sec)
            q = g.sample_distance(rays, st, 5)
            tag = name + ("/sec" if sec else "/pri")
            for k in ("ok", "exited", "t", "grad", "coeffs", "n_evals"):
                res[tag + "/" + k] = np.ascontiguousarray(q[k])
            res[tag + "/T"] = g.transmittance(rays, st, 5)
        g.close()
    fails = None
    lib = abi.load()
    if hasattr(lib, "scn_debug_bounds_failures"):
        arr = (C.c_ulonglong * 8)()
        lib.scn_debug_bounds_failures.restype = C.c_int
        if lib.scn_debug_bounds_failures(arr) == 0:
            fails = [int(v) for v in arr]
    np.savez(out_path, fails=np.array(fails if fails is not None else [-1] * 8), **res)


if __name__ == "__main__":
    if len(sys.argv) > 3 and sys.argv[2] == "--worker":
        worker(sys.argv[3], sys.argv[4]); sys.exit(0)
    out_dir = os.path.join(ROOT, "gpurun_out"); os.makedirs(out_dir, exist_ok=True)
    dbg = os.path.join(out_dir, "libscn_b200_dbg.so")
    build.compile_and_link(dbg, os.path.join(out_dir, "dbg_obj"), extra_flags=["-DSCN_DEBUG_BOUNDS"])
    outs = []
    for tag, lib in (("product", abi.LIB_PATH), ("debug", dbg)):
        o = os.path.join(out_dir, "debug_bounds_%s.npz" % tag)
        subprocess.run([sys.executable, os.path.abspath(__file__), str(int(np.log2(n))), "--worker", lib, o], check=True)
        outs.append(dict(np.load(o)))
    prod, deb = outs
    print("# debug-bounds run: %d rays x %d media x {primary, secondary}, sampleDistance + transmittance through the C ABI" % (n, len(MEDIA)))
    print("# assertion counters of the -DSCN_DEBUG_BOUNDS build (0 rebuild slot, 1 iso-ray rebuild slot, 2 padded list <= K, 3 entry loop inside the list,")
    print("#   4 k_march ray index at fetch, 5 at finish, 6 k_march1d ray index at fetch, 7 at finish):", [int(v) for v in deb["fails"]])
    bad = 0
    for k in sorted(prod):
        if k == "fails":
            continue
        a, b = prod[k], deb[k]
        same = np.array_equal(a, b, equal_nan=True)
        nan = int(np.isnan(b.astype(np.float64)).sum()) - int(np.isnan(a.astype(np.float64)).sum())
        if not same or nan:
            bad += 1
            print("DIFF %-60s differing %d  extra NaN %d" % (k, int((a != b).sum()), nan))
    print("# outputs compared: %d arrays, %d differ between the NaN-poisoned debug build and the product build" % (len(prod) - 1, bad))
    for f in ("debug_bounds_product.npz", "debug_bounds_debug.npz", "libscn_b200_dbg.so"):
        os.remove(os.path.join(out_dir, f))
    sys.exit(1 if bad or any(int(v) != 0 for v in deb["fails"]) else 0)
