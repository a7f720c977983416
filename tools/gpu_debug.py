import sys, os
import numpy as np
sys.path.insert(0, ".")
from raytracingoneweekend_b200 import api, _abi as A
from oracle import binding as O
sys.path.insert(0, "tests")
from conftest import oracle_scene_from

# many spheres
fz = api.FrozenHittableList.generate("many_spheres", 5, 20000)
print(fz.stats())
osc = oracle_scene_from(O, fz)
rng = np.random.default_rng(4)
o = np.tile(np.array([60, 12, 14], np.float32), (4000, 1))
tgt = rng.uniform(-40, 40, (4000, 3)).astype(np.float32); tgt[:, 1] = rng.uniform(0, 4, 4000)
d = tgt - o
oi, ot, _, _ = osc.trace_rays(o, d, 0.001, 1000.0)
for name, fl in [("bvh", 0), ("brute", A.RENDER_BRUTE_FORCE)]:
    gi, gt, _, _ = fz.hit(o, d, 0.001, 1000.0, flags=fl)
    bad = np.nonzero(gi != oi)[0]
    print(name, "mismatch", len(bad), "t-bit mismatch", (gt.view(np.uint32) != ot.view(np.uint32)).sum())
    for k in bad[:8]:
        print("   ray", k, "gpu", hex(int(gi[k])), gt[k], "oracle", hex(int(oi[k])), ot[k])

# fast-build t error by type on rtiow
fz = api.FrozenHittableList.generate("random_scene", 1)
osc = oracle_scene_from(O, fz)
rng = np.random.default_rng(6)
n = 60000
orig = np.tile(np.array([13, 2, 3], np.float32), (n, 1)) + rng.normal(0, 0.5, (n, 3)).astype(np.float32)
tg = rng.uniform(-6, 6, (n, 3)).astype(np.float32); tg[:, 1] = rng.uniform(0, 1.5, n)
dd = tg - orig
gi, gt, gp, gn = fz.hit(orig, dd, flags=0)
oi, ot, op, on = osc.trace_rays(orig, dd)
same = (gi == oi) & (oi != 0)
typ = (oi >> np.uint64(32)).astype(int)
for t in range(1, 9):
    m = same & (typ == t)
    if m.any():
        rel = np.abs(gt[m] - ot[m]) / np.abs(ot[m])
        k = np.argmax(rel)
        idx = np.nonzero(m)[0][k]
        print("type", t, "count", m.sum(), "max rel", rel.max(), "frac>1e-5", (rel > 1e-5).mean(), "worst: id", hex(int(oi[idx])), "t", gt[idx], ot[idx])
m = same & (typ == 1) & (oi != A.obj_id(1, 0))
rel = np.abs(gt[m] - ot[m]) / np.abs(ot[m]); print("small spheres: max rel", rel.max(), "frac>1e-5", (rel > 1e-5).mean(), "p99", np.quantile(rel, 0.99))
