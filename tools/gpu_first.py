"""First GPU session: parity against the oracle + first timings."""
import ctypes as C, json, sys, time
import numpy as np
sys.path.insert(0, ".")
from raytracingoneweekend_b200 import api, _abi as A
from oracle import binding as B

def oracle_scene_from(frozen):
    o = B.OracleScene()
    for k, arr in frozen.lists().items():
        o.add(k, arr)
    return o

out = {}
tf, ms = api.fp32_peak()
print("fp32 peak TFLOP/s", tf, "ms", ms); out["fp32_peak_tflops"] = tf

cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20., 1.5, 0.1, 10.)
fz = api.FrozenHittableList.generate("random_scene", 1)
print(fz.stats())
osc = oracle_scene_from(fz)

# ---- closest-hit parity on random rays
rng = np.random.default_rng(0)
n = 20000
orig = np.tile(np.array([13, 2, 3], np.float32), (n, 1)) + rng.normal(0, 0.5, (n, 3)).astype(np.float32)
tgt = rng.uniform(-6, 6, (n, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 1.5, n)
dirs = tgt - orig
oi, ot, op, on = osc.trace_rays(orig, dirs)
for name, fl in [("exact_bvh", 0), ("exact_brute", A.RENDER_BRUTE_FORCE), ("exact_bvh_global", A.RENDER_NO_SMEM_SCENE), ("fast_bvh", A.RENDER_FAST_MATH)]:
    gi, gt, gp, gn = fz.hit(orig, dirs, flags=fl)
    same_t = (gt.view(np.uint32) == ot.view(np.uint32))
    print(name, "ids equal", (gi == oi).mean(), "t bit-equal", same_t.mean(), "points bit-equal", (gp.view(np.uint32) == op.view(np.uint32)).all(1).mean(),
          "normals bit-equal", (gn.view(np.uint32) == on.view(np.uint32)).all(1).mean(), "hits", (oi != 0).mean())
    if name == "fast_bvh":
        m = (gi == oi) & (oi != 0)
        rel = np.abs(gt[m] - ot[m]) / np.abs(ot[m]); print("  fast: max rel t err", rel.max(), "frac>1e-5", (rel > 1e-5).mean())
        ground = oi[m] == A.obj_id(1, 0)
        print("  fast: ground-sphere hits", ground.mean(), "max rel on ground", rel[ground].max(), "max rel elsewhere", rel[~ground].max(), "frac>1e-5 elsewhere", (rel[~ground] > 1e-5).mean())

# ---- image parity (exact), 96x64x8spp
W, H, SPP = 96, 64, 8
for flags_extra, label in [(0, "fixed"), (A.RENDER_ADAPTIVE, "adaptive")]:
    p = api.render_params(W, H, SPP if not flags_extra else 64, seed=7, flags=flags_extra)
    t0 = time.time(); ro = osc.render(cam, p); t1 = time.time()
    r = api.Renderer(fz, cam, p); r.launch(0); r.sync(); g = r.read_pixels()
    o = ro["pixels"]
    eq = {f: bool((g[f].view(np.uint32 if g[f].dtype == np.float32 else g[f].dtype) == o[f].view(np.uint32 if o[f].dtype == np.float32 else o[f].dtype)).all()) for f in ["sum", "n", "avg_depth", "bad_avgs", "bloom", "color"]}
    print(label, "oracle s", round(t1 - t0, 2), "bit-equal fields:", eq, "credited", r.samples_done(), ro["credited"], "traced", r.samples_traced(), ro["traced"])
    if not all(eq.values()):
        bad = np.nonzero((g["sum"].view(np.uint32) != o["sum"].view(np.uint32)).any(1))[0]
        print("  mismatching pixels", len(bad), bad[:10], g["sum"][bad[:3]], o["sum"][bad[:3]])
    out["exact_" + label] = eq
    r.close()

# ---- primary ids
p = api.render_params(300, 200, 4, seed=3, flags=0)
r = api.Renderer(fz, cam, p)
for mode in (0, 1):
    gi, gt = r.primary(mode, 2); oi2, ot2 = osc.primary(cam, p, mode, 2)
    print("primary mode", mode, "ids equal", (gi == oi2).mean(), "t bit equal", (gt.view(np.uint32) == ot2.view(np.uint32)).mean())
r.close()
p = api.render_params(300, 200, 4, seed=3, flags=0)
r = api.Renderer(fz, cam, p)
gi, gt = r.primary(0, 0); oi2, ot2 = osc.primary(cam, p, 0, 0)
m = (gi == oi2) & (oi2 != 0)
print("fast primary ids equal", (gi == oi2).mean(), "max rel t", (np.abs(gt[m] - ot2[m]) / ot2[m]).max())
r.close()

# ---- timings
def bench(label, fz, cam, W, H, spp, flags=0, reps=2):
    p = api.render_params(W, H, spp, seed=1, flags=flags)
    r = api.Renderer(fz, cam, p)
    best = None
    for i in range(reps + 1):
        r.reset(); r.launch(0); r.sync(); ms, _ = r.elapsed_ms()
        if i > 0: best = ms if best is None else min(best, ms)
    tr = r.samples_traced(); pc = r.perf_counters()
    rate = W * H * spp / (best * 1e-3) / 1e6
    print(label, W, H, spp, "ms", round(best, 2), "Msamples/s", round(rate, 1), {k: v for k, v in pc.items() if k in ("segments", "node_visits", "prim_tests", "sdf_evals", "march_steps", "march_entered")})
    out[label] = dict(ms=best, msamples_s=rate)
    r.close()

cam169 = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20., 16 / 9, 0.1, 10.)
bench("C1_fast", fz, cam, 1200, 800, 10)
bench("C1_fast_counters", fz, cam, 1200, 800, 10, A.RENDER_COUNTERS)
bench("C1_exact", fz, cam, 1200, 800, 10, 0)
bench("C1_fast_global", fz, cam, 1200, 800, 10, A.RENDER_NO_SMEM_SCENE)
bench("C1_fast_brute", fz, cam, 1200, 800, 10, A.RENDER_BRUTE_FORCE, reps=1)
bench("C2_fast_64spp", fz, cam169, 1920, 1080, 64)
fzc = api.FrozenHittableList.generate("rtiow_classic", 1)
bench("C2classic_fast_64spp", fzc, cam169, 1920, 1080, 64)
bench("C2classic_fast_counters", fzc, cam169, 1920, 1080, 16, A.RENDER_COUNTERS)
json.dump(out, open("gpurun_out/first.json", "w"), indent=1, default=str)
rgb = None
p = api.render_params(600, 400, 64, seed=1)
r = api.Renderer(fz, cam, p); r.launch(0); r.sync(); rgb = r.read_rgb(); r.close()
from tools.png import write_png
write_png("gpurun_out/first.png", rgb)
