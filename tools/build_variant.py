"""Build an experimental variant of the library with extra -D flags (kernels only; host objects are reused):
    python tools/build_variant.py NAME -DOM_PHASE_SYNC=2 ...   ->  raytracingoneweekend_b200/_variants/libom_NAME.so
Run with OM_LIB_PATH=<that file>.  Variants are scratch (git-ignored)."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from raytracingoneweekend_b200 import build as B
name, defs = sys.argv[1], sys.argv[2:]
B.build()
vdir = os.path.join(B.HERE, "_variants"); os.makedirs(vdir, exist_ok=True)
def one(item):
    src, extra = item
    base = src.rsplit(".", 1)[0]
    if not src.startswith("om_kernels"):
        return os.path.join(B.OBJ, base + ".o")
    obj = os.path.join(vdir, "%s_%s.o" % (base, name))
    r = subprocess.run(["nvcc"] + B.ARCH + B.COMMON + extra + defs + ["-c", os.path.join(B.CSRC, src), "-o", obj], capture_output=True, text=True)
    if r.returncode: raise RuntimeError(r.stderr)
    return obj
with ThreadPoolExecutor(4) as ex: objs = list(ex.map(one, B.SOURCES))
lib = os.path.join(vdir, "libom_%s.so" % name)
subprocess.run(["nvcc"] + B.ARCH + ["-shared", "-o", lib] + objs, check=True)
print(lib)
