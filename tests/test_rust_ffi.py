"""The Rust -sys crate cannot be compiled here (no rustc), so it is checked mechanically:
  * rust/ottomarcher-cuda-sys/src/ffi.rs is exactly what tools/gen_rust_ffi.py generates from include/ottomarcher.h;
  * re-parsed on its own, its extern block has every header symbol with the same argument count, and every symbol is exported by
    the built library;
  * every #[repr(C)] struct has the size and field offsets of the ctypes mirror that the GPU tests bind;
  * build.rs compiles the same sources with the same per-file flags as raytracingoneweekend_b200/build.py;
  * the hand-written wrapper only calls functions that exist, with the right number of arguments."""
import ctypes as C
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUST = os.path.join(ROOT, "rust", "ottomarcher-cuda-sys")
sys.path.insert(0, os.path.join(ROOT, "tools"))

PRIM = {"f32": (4, 4), "u32": (4, 4), "i32": (4, 4), "u64": (8, 8), "u8": (1, 1), "usize": (8, 8), "om_status": (4, 4)}


def rust_structs(text):
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\][^\n]*\npub struct (\w+) \{\n(.*?)\n\}", text, flags=re.S):
        fields = []
        for line in m.group(2).splitlines():
            fm = re.match(r"\s*pub (\w+): (?:\[(\w+); (\d+)\]|(\w+)),", line)
            assert fm, line
            fields.append((fm.group(1), fm.group(2) or fm.group(4), int(fm.group(3) or 1)))
        out[m.group(1)] = fields
    return out


def layout(name, structs, cache):
    """size, alignment and field offsets under repr(C)"""
    if name in PRIM:
        return PRIM[name] + ({},)
    if name in cache:
        return cache[name]
    off, align, offs = 0, 1, {}
    for f, t, n in structs[name]:
        sz, al, _ = layout(t, structs, cache)
        off = (off + al - 1) // al * al
        offs[f] = off
        off += sz * n
        align = max(align, al)
    size = (off + align - 1) // align * align
    cache[name] = (size, align, offs)
    return cache[name]


def test_ffi_rs_is_what_the_header_generates():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_rust_ffi.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_extern_block_matches_header_and_library(api):
    import gen_rust_ffi as G
    from raytracingoneweekend_b200 import _lib
    hdr = G.parse(open(G.HEADER).read())
    text = open(os.path.join(RUST, "src", "ffi.rs")).read()
    block = text[text.index('extern "C" {'):]
    rust = {m.group(1): [a for a in m.group(2).split(",") if a.strip()] for m in re.finditer(r"pub fn (\w+)\((.*?)\)(?: -> [^;]+)?;", block)}
    want = {name: params for name, _, params in hdr["funcs"]}
    assert set(rust) == set(want), set(rust) ^ set(want)
    assert len(want) >= 70
    for name, params in want.items():
        assert len(rust[name]) == len(params), name
    lib = C.CDLL(_lib.LIB_PATH)
    assert all(hasattr(lib, n) for n in rust)
    assert set(rust) == set(_lib.PROTOTYPES)
    for name, (ret, args) in _lib.PROTOTYPES.items():   # the ctypes prototypes the tests call through have the same arity
        assert len(args) == len(want[name]), name
    consts = dict(re.findall(r"pub const (\w+): \w+ = (\d+);", text))
    for k, v in (("OM_RENDER_ADAPTIVE", 1), ("OM_RENDER_POOL", 512), ("OM_FREEZE_GPU_BVH", 16), ("OM_CHUNK_PIXELS", 2730), ("OM_ERR_NO_DEVICE", 5),
                 ("OM_SDF_LOAD", 17), ("OM_VIEW_NOALLOC_MAX_PIXELS", 2097152), ("OM_MAT_DIELECTRIC", 2)):
        assert int(consts[k]) == v, k


def test_struct_sizes_and_offsets_equal_the_ctypes_mirrors(A):
    structs = rust_structs(open(os.path.join(RUST, "src", "ffi.rs")).read())
    mirrors = {"om_material": A.Material, "om_sphere": A.Sphere, "om_cube": A.Cube, "om_infinite_plane": A.InfinitePlane, "om_barycentric": A.Barycentric,
               "om_marched_sphere": A.MarchedSphere, "om_marched_box": A.MarchedBox, "om_marched_torus": A.MarchedTorus, "om_dyn_marched": A.DynMarched, "om_dyn_traced": A.DynTraced,
               "om_camera": A.Camera, "om_pixel": A.Pixel, "om_render_params": A.RenderParams}
    assert set(mirrors) == set(structs), set(mirrors) ^ set(structs)
    cache = {}
    for name, T in mirrors.items():
        size, _, offs = layout(name, structs, cache)
        assert size == C.sizeof(T), (name, size, C.sizeof(T))
        cfields = [f[0] for f in T._fields_]
        assert list(offs) == cfields, (name, list(offs), cfields)
        for f in cfields:
            assert offs[f] == getattr(T, f).offset, (name, f)


def test_build_rs_compiles_what_build_py_compiles():
    from raytracingoneweekend_b200 import build as B
    rs = open(os.path.join(RUST, "build.rs")).read()
    got = {m.group(1): re.findall(r'"([^"]+)"', m.group(2)) for m in re.finditer(r'\("(om_\w+\.(?:cu|cpp))", &\[(.*?)\]\)', rs)}
    assert got == {src: extra for src, extra in B.SOURCES}, (got, B.SOURCES)
    for flag in ("arch=compute_100a,code=sm_100a", "-lineinfo", "-ffp-contract=off"):
        assert flag in rs


def test_wrapper_calls_exist_with_the_right_arity():
    import gen_rust_ffi as G
    want = {name: len(params) for name, _, params in G.parse(open(G.HEADER).read())["funcs"]}
    for fn in ("lib.rs",):
        text = open(os.path.join(RUST, "src", fn)).read()
        calls = re.findall(r"\b(om_[a-z0-9_]+)\(([^()]*(?:\([^()]*\)[^()]*)*)\)", text)
        seen = 0
        for name, args in calls:
            if name not in want:
                continue
            n = len([a for a in re.split(r",(?![^()]*\))", args) if a.strip()])
            assert n == want[name], (fn, name, n, want[name])
            seen += 1
        assert seen >= 15
    # the struct literals of convert.rs name exactly the fields of the generated structs
    structs = rust_structs(open(os.path.join(RUST, "src", "ffi.rs")).read())
    conv = open(os.path.join(RUST, "src", "convert.rs")).read()
    code = "\n".join(l for l in conv.splitlines() if not l.lstrip().startswith("//"))
    for m in re.finditer(r"\b(om_\w+) \{([^{}]*(?:\{[^{}]*\}[^{}]*)*)\}", code):
        name, body = m.group(1), m.group(2)
        if name not in structs or ":" not in body:
            continue
        fields = re.findall(r"(?:^|,)\s*(\w+):", re.sub(r"\([^()]*\)", "()", body.replace("\n", " ")))
        assert fields == [f for f, _, _ in structs[name]], (name, fields)
