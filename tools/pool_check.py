"""k_render_pool (OM_RENDER_POOL) against the persistent kernel: same Vec<Pixel>, same counters.  python tools/pool_check.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from raytracingoneweekend_b200 import api, _abi as A
ok = True
for kind, seed, W, H, spp, extra in [("random_scene", 1, 96, 64, 8, 0), ("random_scene", 1, 333, 222, 5, 0), ("glass_heavy", 4, 160, 90, 6, 0), ("rtiow_classic", 1, 160, 90, 6, 0),
                                     ("marched", 3, 96, 54, 3, 0), ("basic_scene", 1, 64, 48, 4, 0), ("random_scene", 1, 128, 96, 64, A.RENDER_ADAPTIVE)]:
    fz = api.FrozenHittableList.generate(kind, seed)
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    out = []
    for fl in (A.RENDER_PERSISTENT, A.RENDER_POOL):
        r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=17, flags=fl | extra))
        if extra:
            for _ in range(8):
                r.launch(8)
        else:
            r.launch(0)
        r.sync()
        out.append((r.read_pixels().copy(), r.samples_done(), r.samples_traced()))
        r.close()
    (a, ca, ta), (b, cb, tb) = out
    same = all((a[f].view(np.uint32) == b[f].view(np.uint32)).all() if a[f].dtype == np.float32 else (a[f] == b[f]).all() for f in ("sum", "n", "avg_depth", "bad_avgs", "bloom", "color"))
    print(kind, W, H, spp, "adaptive" if extra else "", "pixels equal:", same, "credited", ca, cb, "traced", ta, tb)
    ok &= same and ca == cb and ta == tb
print("ALL OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
