"""Where the end-to-end step of bench.py spends its host time:  python tools/e2e_breakdown.py [WORKLOAD] [SPP]"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import bench
from raytracingoneweekend_b200 import api, dist as D, _abi as A
wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[wl]
if len(sys.argv) > 2: spp = int(sys.argv[2])
fz = api.FrozenHittableList.generate(kind, seed, n_param)
cam = bench.make_camera(api, aspect)
lists = fz.lists()
pinned = torch.empty(W * H * 40, dtype=torch.uint8).pin_memory()
px = pinned.numpy().view(A.PIXEL_DTYPE)
kw = dict(max_depth=bench.MAX_DEPTH, tmin=bench.TMIN, tmax=bench.TMAX, seed=1)
for it in range(3):
    t = [time.perf_counter()]
    world_list = api.FrozenHittableList()
    for k, arr in lists.items():
        if len(arr):
            api._lib.call("om_scene_add_" + k, world_list.h, arr, len(arr))
    t.append(time.perf_counter())
    api._lib.call("om_scene_freeze", world_list.h, None, 0)
    t.append(time.perf_counter())
    d2 = D.DistributedRenderer(world_list, cam, W, H, spp, 0, 1, mode="spp", device=0, **kw)
    t.append(time.perf_counter())
    d2.render(); d2.r.sync()
    t.append(time.perf_counter())
    d2.r.read_pixels(px)
    t.append(time.perf_counter())
    d2.close()
    t.append(time.perf_counter())
    names = ["add", "freeze", "create", "render+sync", "read_pixels", "close"]
    print("  ".join("%s %.1f ms" % (n, 1e3 * (b - a)) for n, a, b in zip(names, t, t[1:])), " total %.1f ms" % (1e3 * (t[-1] - t[0])))
