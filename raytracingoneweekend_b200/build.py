"""Build recipe of libottomarcher_cuda.so (nvcc, sm_100a only, in-tree).

    python -m raytracingoneweekend_b200.build [--force]

Per-file flags matter: the parity kernels (om_kernels_exact.cu) and the framebuffer kernels in
om_api.cu are compiled with -fmad=false; host code with -ffp-contract=off.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libottomarcher_cuda.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-fvisibility=hidden"]
SOURCES = [  # (file, extra flags)
    ("om_kernels_fast.cu", []),
    ("om_kernels_exact.cu", ["-fmad=false"]),
    ("om_api.cu", ["-fmad=false"]),
    ("om_lbvh.cu", []),
    ("om_host.cpp", []),
    ("om_scene.cpp", []),
    ("om_bvh.cpp", []),
]


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "ottomarcher.h")]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    deps = _deps()
    if not force and not _stale(LIB, deps):
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")

    def compile_one(item):
        src, extra = item
        obj = os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o")
        if not force and not _stale(obj, deps):
            return obj
        cmd = [nvcc] + ARCH + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=6) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc] + ARCH + ["-shared", "-o", LIB] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
