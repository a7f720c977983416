"""Oracle fixtures of the statistical image tests (tests/test_gpu_image.py): per-pixel mean and sample variance of N oracle samples.

    python tools/make_golden_stat.py [rtiow|glass|marched ...]      # rtiow: ~15 min on 8 cores, the others a few minutes each

  rtiow_300x200_4096spp.npz    random_scene(1), the north_star's 4096-spp convergence check at 300 x 200
  glass_150x100_2048spp.npz    glass_heavy(4): >= 50 % dielectric spheres (C4's scene)
  marched_96x64_1024spp.npz    marched(3): 65 marched objects (C3's scene)
mean = sum / n (float32), var = unbiased sample variance per channel (float32)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from raytracingoneweekend_b200 import api
from oracle import binding as O

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
JOBS = {"rtiow": ("random_scene", 1, 300, 200, 4096, 2024), "glass": ("glass_heavy", 4, 150, 100, 2048, 2025), "marched": ("marched", 3, 96, 64, 1024, 2026)}
for name in (sys.argv[1:] or list(JOBS)):
    kind, sseed, W, H, spp, seed = JOBS[name]
    fz = api.FrozenHittableList.generate(kind, sseed)
    osc = O.OracleScene()
    for k, arr in fz.lists().items():
        osc.add(k, arr)
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    t0 = time.time()
    r = osc.render(cam, api.render_params(W, H, spp, seed=seed), want_sumsq=True, nthreads=int(os.environ.get("THREADS", "6")))
    s = r["pixels"]["sum"].astype(np.float64)
    mean = s / spp
    var = np.maximum(r["sumsq"] / spp - mean ** 2, 0.0) * spp / (spp - 1.0)
    f = os.path.join(OUT, "%s_%dx%d_%dspp.npz" % (name, W, H, spp))
    np.savez_compressed(f, mean=mean.astype(np.float32), var=var.astype(np.float32), n=spp, seed=seed, width=W, height=H)
    print(name, "done in %.0f s" % (time.time() - t0), os.path.getsize(f), "bytes", flush=True)
