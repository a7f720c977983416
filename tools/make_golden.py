"""Generates tests/golden/*.npz from the ORACLE (oracle/, the CPU restatement of the reference).

    python tools/make_golden.py            # ~3 minutes on 8 cores

Fixtures (all for the product generator's random_scene(seed=1), camera of main.rs:136-142):
  (the statistical fixtures of the 3-sigma image test come from tools/make_golden_stat.py)
  rtiow_primary_240x160.npz primary-hit (obj_id, t) maps: pixel-centre rays and sample-3 rays (seed 5)
  rtiow_exact_64x48.npz     full Stats of a 64x48, 12 spp render (seed 11), fixed and adaptive (spp 48)
"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from raytracingoneweekend_b200 import api, _abi as A
from oracle import binding as O

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
os.makedirs(OUT, exist_ok=True)
fz = api.FrozenHittableList.generate("random_scene", 1)
osc = O.OracleScene()
for k, arr in fz.lists().items():
    osc.add(k, arr)

def cam(aspect):
    return api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)

t0 = time.time()
p = api.render_params(240, 160, 8, seed=5)
i0, t_0 = osc.primary(cam(1.5), p, 0, 0)
i1, t_1 = osc.primary(cam(1.5), p, 1, 3)
np.savez_compressed(os.path.join(OUT, "rtiow_primary_240x160.npz"), ids_center=i0, t_center=t_0, ids_sample3=i1, t_sample3=t_1, seed=5, spp_max=8)
p = api.render_params(64, 48, 12, seed=11)
a = osc.render(cam(64 / 48), p)
pa = api.render_params(64, 48, 48, seed=11, flags=A.RENDER_ADAPTIVE)
b = osc.render(cam(64 / 48), pa)
np.savez_compressed(os.path.join(OUT, "rtiow_exact_64x48.npz"), fixed=a["pixels"], adaptive=b["pixels"], adaptive_credited=b["credited"], adaptive_traced=b["traced"])
print("small fixtures", time.time() - t0)
print("done", time.time() - t0)
