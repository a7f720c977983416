"""GPU: wavefront pipeline vs persistent kernel, bit for bit, then timings.  python tools/wf_check.py [--time]"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import bench
from raytracingoneweekend_b200 import api, _abi as A

def same(g, o):
    ok = True
    for f in ("sum", "avg_depth"):
        e = (g[f].view(np.uint32) == o[f].view(np.uint32)).all(); ok &= bool(e)
        if not e: print("   differs:", f, int((g[f].view(np.uint32) != o[f].view(np.uint32)).sum()))
    for f in ("n", "bad_avgs", "bloom", "color"):
        e = (g[f] == o[f]).all(); ok &= bool(e)
        if not e: print("   differs:", f, int((g[f] != o[f]).sum()))
    return ok

def run(fz, cam, W, H, spp, flags, passes=None):
    r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=5, flags=flags))
    if passes:
        for p in passes: r.launch(p)
    else:
        r.launch(0)
    r.sync()
    px = r.read_pixels(); d = r.samples_done(); t = r.samples_traced(); r.close()
    return px, d, t

allok = True
for kind, seed in (("random_scene", 1), ("marched", 3), ("basic_scene", 1)):
    fz = api.FrozenHittableList.generate(kind, seed)
    cam = bench.make_camera(api, 1.5)
    for label, fl, spp, passes in (("fixed", 0, 12, None), ("fixed 3 passes", 0, 12, [5, 5, 5]), ("adaptive", A.RENDER_ADAPTIVE, 48, None), ("adaptive passes", A.RENDER_ADAPTIVE, 48, [8] * 6)):
        a = run(fz, cam, 300, 200, spp, fl | A.RENDER_PERSISTENT, passes)
        b = run(fz, cam, 300, 200, spp, fl | A.RENDER_WAVEFRONT, passes)
        ok = same(a[0], b[0]) and a[1] == b[1] and a[2] == b[2]
        allok &= ok
        print(kind, label, "OK" if ok else "MISMATCH", "done", a[1], b[1], "traced", a[2], b[2])
print("ALL OK" if allok else "FAILED")
if "--time" in sys.argv:
    for wl, spp in (("C2", 16), ("C2", 64), ("C2classic", 64), ("C4", 16), ("C3", 8)):
        kind, seed, n_param, W, H, _, aspect = bench.WORKLOADS[wl]
        fz = api.FrozenHittableList.generate(kind, seed, n_param)
        cam = bench.make_camera(api, aspect)
        for fl in (A.RENDER_PERSISTENT, A.RENDER_WAVEFRONT):
            r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=1, flags=fl))
            for i in range(2):
                r.reset(); r.launch(0); r.sync(); ms, _ = r.elapsed_ms()
            print("%s %d spp flags=%d: %.2f ms  %.1f Msamples/s" % (wl, spp, fl, ms, W * H * spp / ms / 1e3))
            r.close()
