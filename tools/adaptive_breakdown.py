"""Where a pass of the adaptive scheduler spends its time (one GPU): python tools/adaptive_breakdown.py [--spp 256] [--pass-spp 16]"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
import bench
from raytracingoneweekend_b200 import api, dist, _abi as A
ap = argparse.ArgumentParser(); ap.add_argument("--spp", type=int, default=256); ap.add_argument("--pass-spp", type=int, default=16); ap.add_argument("--workload", default="C2"); ap.add_argument("--flags", type=int, default=0); ap.add_argument("--emulate-world", type=int, default=1, help="render only the chunks rank 0 of this many ranks would own")
a = ap.parse_args()
kind, seed, n_param, W, H, spp, aspect = bench.WORKLOADS[a.workload]
fz = api.FrozenHittableList.generate(kind, seed, n_param)
cam = bench.make_camera(api, aspect)
for rep in range(2):
    s = dist.AdaptiveScheduler(fz, cam, W, H, a.spp, 0, 1, pass_spp=a.pass_spp, flags=a.flags)
    r = s.r
    if a.emulate_world > 1:
        s.owners = (np.arange(r.n_chunks) % a.emulate_world).astype(np.uint8)
    T = {}
    def timed(name, fn):
        torch.cuda.synchronize(); t = time.perf_counter(); out = fn(); torch.cuda.synchronize(); T[name] = T.get(name, 0.0) + time.perf_counter() - t; return out
    t_all = time.perf_counter()
    live = timed("live_per_chunk", r.live_per_chunk)
    done, passes, traced0 = 0, 0, 0
    per_pass = []
    while done < a.spp and int(live[s.owners == 0].sum()) > 0:
        timed("set_owners", lambda: r.set_chunk_owners(s.owners, 0))
        t0 = T.get("launch", 0.0)
        timed("launch", lambda: (r.launch(a.pass_spp), r.sync()))
        timed("export", lambda: r.export_state(s.state.data_ptr(), only_owned=True))
        timed("import", lambda: r.import_state(s.state.data_ptr()))
        timed("samples_done", r.samples_done)
        n_live = int(live[s.owners == 0].sum())
        live = timed("live_per_chunk", r.live_per_chunk)
        row = (n_live, round((T["launch"] - t0) * 1e3, 2))
        if a.flags & A.RENDER_COUNTERS:
            c = r.perf_counters()
            row += (round((c["lane_iters"] - getattr(s, "_li", 0)) / max(1, 32 * (c["warp_iters"] - getattr(s, "_wi", 0))), 3), c["traced"] - getattr(s, "_tr", 0), c["segments"] - getattr(s, "_sg", 0))
            s._li, s._wi, s._tr, s._sg = c["lane_iters"], c["warp_iters"], c["traced"], c["segments"]
        per_pass.append(row)
        done += a.pass_spp; passes += 1
    total = time.perf_counter() - t_all
    traced = r.samples_traced()
    print("rep", rep, "passes", passes, "total ms", round(total * 1e3, 1), "traced", traced, "Msamples/s", round(traced / total / 1e6, 1))
    print({k: round(v * 1e3, 2) for k, v in T.items()})
    print("live pixels before each pass, launch ms:", per_pass)
    s.close()
