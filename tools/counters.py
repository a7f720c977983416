"""Executed-work counters per sample of a workload: python tools/counters.py C3 [--spp 4] [--flags N]"""
import argparse, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from raytracingoneweekend_b200 import api, _abi as A
ap = argparse.ArgumentParser(); ap.add_argument("workload"); ap.add_argument("--spp", type=int, default=4); ap.add_argument("--flags", type=int, default=0); ap.add_argument("--freeze", type=int, default=0)
a = ap.parse_args()
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[a.workload]
fz = api.FrozenHittableList.generate(kind, seed, n_param, flags=a.freeze)
print(fz.stats())
r = api.Renderer(fz, bench.make_camera(api, aspect), api.render_params(W, H, a.spp, seed=1, flags=a.flags | A.RENDER_COUNTERS))
r.launch(0); r.sync()
c = r.perf_counters(); ms, _ = r.elapsed_ms()
print("lane occupancy of the path loop", round(c["lane_iters"] / max(1, 32 * c["warp_iters"]), 4))
print(a.workload, "ms", round(ms, 2), {k: round(v / c["traced"], 3) for k, v in c.items() if k not in ("traced", "credited") and isinstance(v, int)})
r.close()
