"""In-tree build of the CUDA library (sm_100a only) and, as a separate artefact, of the test oracle.

`build_library()` compiles csrc/*.cu + csrc/*.cpp (one nvcc -c per file, in parallel) and links libscn_b200.so next to this file with
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo
(nvcc cross-compiles without a GPU).  The CUDA runtime is linked statically, so the .so travels to the GPU box.
`build_oracle()` runs oracle/Makefile (`oracle` always; `ref` only where /root/reference exists) — building the
checker is not using it: nothing in this package imports or loads anything under oracle/.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libscn_b200.so")
SOURCES = ["scn_api.cu", "scn_march_sample.cu", "scn_march_transmit.cu", "scn_march1d.cu", "scn_render.cu", "scn_json.cpp", "scn_params.cpp"]
HEADERS = ["scn_device.cuh", "scn_fast.cuh", "scn_march.cuh", "scn_march1d.cuh", "scn_internal.h", os.path.join("..", "..", "include", "scn.h"), os.path.join("..", "..", "include", "scn_render.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC,-O2"]
OBJDIR = os.path.join(HERE, "build")  # object files (git-ignored; .gpurunignore keeps them off the GPU box)


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the library has no CPU build")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS]
    if not force and not _stale(LIB, deps):
        return LIB
    return compile_and_link(LIB, OBJDIR, extra_flags=(["-Xptxas", "-v"] if verbose else []), verbose=verbose)


def compile_and_link(lib, objdir, extra_flags=(), verbose=False):
    """One nvcc -c per source file, all at once (the kernel families are translation units of their own: scn_march_sample.cu,
    scn_march_transmit.cu, scn_march1d.cu), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()

    def one(src):
        obj = os.path.join(objdir, os.path.splitext(src)[0] + ".o")
        cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + ["-c", "-o", obj, os.path.join(CSRC, src)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as ex:
        done = list(ex.map(one, SOURCES))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib] + [o for o, _ in done]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print("".join(err for _, err in done))
    return lib


def build_oracle(with_reference=None):
    """Compile oracle/scn_oracle.c (always) and oracle/_ref (only when the reference tree is mounted)."""
    odir = os.path.join(ROOT, "oracle")
    subprocess.run(["make", "-s", "-C", odir, "oracle"], check=True)
    if with_reference is None:
        with_reference = os.path.isdir("/root/reference/src/core")
    if with_reference:
        subprocess.run(["make", "-s", "-j", str(os.cpu_count() or 1), "-C", odir, "ref"], check=True)
    return os.path.join(odir, "_build", "libscn_oracle.so")


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
