"""Freeze time and render rate of a workload with the host (binned SAH) and the device (Morton/Karras) BVH build:
    python tools/freeze_time.py C5 [--spp 4]"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from raytracingoneweekend_b200 import api, _lib, _abi as A
ap = argparse.ArgumentParser(); ap.add_argument("workload"); ap.add_argument("--spp", type=int, default=4)
a = ap.parse_args()
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[a.workload]
cam = bench.make_camera(api, aspect)
api.fp32_peak(0)  # CUDA context up before anything is timed
for name, flags in (("host SAH", 0), ("device LBVH", A.FREEZE_GPU_BVH), ("device LBVH (2nd)", A.FREEZE_GPU_BVH)):
    s = api.FrozenHittableList()
    _lib.call("om_scene_generate", s.h, api.SCENE_KINDS[kind], seed, n_param)
    t0 = time.perf_counter(); _lib.call("om_scene_freeze", s.h, None, flags); dt = time.perf_counter() - t0
    st = s.stats()
    r = api.Renderer(s, cam, api.render_params(W, H, a.spp, seed=1))
    r.launch(0); r.sync(); r.reset(); r.elapsed_ms(); r.launch(0); r.sync(); ms, _ = r.elapsed_ms()
    r.close()
    print("%s %-18s freeze %.3f s  depth %d  render %d spp: %.1f ms = %.1f Msamples/s" % (a.workload, name, dt, st["bvh_depth"], a.spp, ms, W * H * a.spp / ms / 1e3))
