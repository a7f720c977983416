"""One wavefront render for ncu: python tools/wf_prof.py C2 --spp 16"""
import argparse, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from raytracingoneweekend_b200 import api, _abi as A
ap = argparse.ArgumentParser(); ap.add_argument("workload"); ap.add_argument("--spp", type=int, default=16); ap.add_argument("--flags", type=int, default=A.RENDER_WAVEFRONT); ap.add_argument("--reps", type=int, default=1)
a = ap.parse_args()
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[a.workload]
fz = api.FrozenHittableList.generate(kind, seed, n_param)
cam = bench.make_camera(api, aspect)
r = api.Renderer(fz, cam, api.render_params(W, H, a.spp, seed=1, flags=a.flags))
for i in range(a.reps):
    r.reset(); r.launch(0); r.sync(); ms, _ = r.elapsed_ms()
    print("%s %d spp flags=%d: %.2f ms %.1f Msamples/s" % (a.workload, a.spp, a.flags, ms, W * H * a.spp / ms / 1e3))
r.close()
