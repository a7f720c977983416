"""One render of a named workload, for ncu:  python tools/prof_run.py C2 --spp 8 [--flags N] [--reps R]"""
import argparse, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from raytracingoneweekend_b200 import api

ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--spp", type=int, default=8)
ap.add_argument("--flags", type=int, default=0)
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[a.workload]
fz = api.FrozenHittableList.generate(kind, seed, n_param)
cam = bench.make_camera(api, aspect)
r = api.Renderer(fz, cam, api.render_params(W, H, a.spp, bench.MAX_DEPTH, bench.TMIN, bench.TMAX, seed=1, flags=a.flags))
for i in range(a.reps):
    r.reset(); r.launch(0); r.sync()
    ms, _ = r.elapsed_ms()
    print("%s %dx%d %d spp flags=%d: %.3f ms, %.1f Msamples/s" % (a.workload, W, H, a.spp, a.flags, ms, W * H * a.spp / ms / 1e3))
r.close()
