"""GPU tests at image level: the 4096-spp 3-sigma agreement with the oracle's golden image, and
size-independent properties at BASELINE.json's full sizes (the oracle cannot run those)."""
import os

import numpy as np
import pytest

from conftest import bits

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def cam_rtiow(api, aspect=1.5):
    return api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)


@pytest.fixture(scope="module")
def rtiow(api):
    return api.FrozenHittableList.generate("random_scene", 1)


def render(api, fz, cam, W, H, spp, flags=0, seed=1, passes=(0,), **kw):
    r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=seed, flags=flags, **kw))
    for k in passes:
        r.launch(k)
    r.sync()
    out = r.read_pixels(), r.samples_done(), r.samples_traced()
    r.close()
    return out


STAT_SCENES = {  # fixtures of tools/make_golden_stat.py: per-pixel mean and sample variance of n oracle samples
    "rtiow": ("random_scene", 1, "rtiow_300x200_4096spp.npz"),      # the north_star's check: 4096 spp, at 300 x 200
    "glass": ("glass_heavy", 4, "glass_150x100_2048spp.npz"),       # C4's scene: >= 50 % dielectric
    "marched": ("marched", 3, "marched_96x64_1024spp.npz"),         # C3's scene: 65 marched objects
}
# measured on a B200 (round 2; raw = no floors at all; fast / exact build): rtiow 0.0051 / 0.0049, glass 0.0057 / 0.0057, marched 0.0048 / 0.0045
# (floored: 0.0014 / 0.0012, 0.0020 / 0.0020, 0.0012 / 0.0011; RMSE 0.0025, 0.0043, 0.0049) - the table is in DESIGN.md section 2
RAW_EXCEEDANCE_BAR = {"rtiow": 0.006, "glass": 0.008, "marched": 0.006}


@pytest.mark.parametrize("build", ["fast", "exact"])
@pytest.mark.parametrize("scene", ["rtiow", "glass", "marched"])
def test_converged_image_within_3_sigma_of_oracle(api, A, scene, build):
    """north_star: converged images agree per pixel within 3 sigma of the measured per-pixel sample variance; RMSE stated.
    Oracle side: tests/golden/<scene>_*.npz (n samples per pixel, seeds 2024-2026); GPU side: an independent render with as many
    samples and another seed, in both builds.  Two statistics:
      raw      z = |mu_gpu - mu_ref| / sqrt(2 var_ref / n) over the pixel-channels whose MEASURED variance is not zero - no floors;
               a Gaussian gives 0.27 % beyond 3; these are heavy-tailed Monte-Carlo estimates whose variance is itself estimated from
               n samples, so somewhat more is expected, and the bar is what was measured plus a margin;
      floored  the same with sigma floored by the f32 rounding of the running sums (1e-5 relative) and 1/8 of an 8-bit display level,
               over ALL pixel-channels: this one also covers the pixels whose n oracle samples happened to be identical."""
    kind, sseed, fname = STAT_SCENES[scene]
    g = np.load(os.path.join(GOLD, fname))
    n, W, H = int(g["n"]), int(g["width"]), int(g["height"])
    mu_ref, var = g["mean"].astype(np.float64), g["var"].astype(np.float64)
    fz = api.FrozenHittableList.generate(kind, sseed)
    px, _, traced = render(api, fz, cam_rtiow(api, W / H), W, H, n, flags=0 if build == "exact" else A.RENDER_FAST_MATH, seed=77)
    assert traced == W * H * n and (px["n"] == n).all()
    mu = px["sum"].astype(np.float64) / n
    diff = np.abs(mu - mu_ref)
    has_var = var > 0
    raw = diff[has_var] / np.sqrt(2.0 * var[has_var] / n)
    raw3 = float((raw > 3.0).mean())
    sigma = np.sqrt(2.0 * var / n + (1e-5 * mu_ref) ** 2 + 5e-4 ** 2)
    z = diff / sigma
    frac3 = float((z > 3.0).mean())
    rmse = float(np.sqrt(np.mean((mu - mu_ref) ** 2)))
    print("%s / %s build, %dx%d, %d spp: raw frac(|z|>3) = %.4f over %.1f %% of the pixel-channels (0.0027 for a Gaussian), floored frac = %.4f, max floored z = %.2f, "
          "RMSE = %.5f, mean = %.4f vs %.4f" % (scene, build, W, H, n, raw3, 100.0 * has_var.mean(), frac3, z.max(), rmse, mu.mean(), mu_ref.mean()))
    if os.environ.get("OM_STAT_LOG"):
        with open(os.environ["OM_STAT_LOG"], "a") as f:
            f.write("%s %s %dx%d %dspp raw3=%.5f has_var=%.4f floored3=%.5f zmax=%.2f rmse=%.5f mean=%.5f ref=%.5f novar_ok=%.4f\n" % (
                scene, build, W, H, n, raw3, has_var.mean(), frac3, z.max(), rmse, mu.mean(), mu_ref.mean(),
                float((diff[~has_var] <= 1e-4 + 1e-4 * mu_ref[~has_var]).mean()) if (~has_var).any() else 1.0))
    assert raw3 < RAW_EXCEEDANCE_BAR[scene]
    assert has_var.mean() > 0.75   # the rest are low-variance (sky) pixels: the f32 running sum's rounding swallows their measured variance
    # tails: over 180 000 pixel-channels one rare bright path through glass reaches z = 9 (fast build, measured); the tail MASS is what is held
    assert frac3 < 0.01 and float((z > 6.0).mean()) < 1e-4 and z.max() < 12.0
    assert rmse < (0.01 if scene == "rtiow" else 0.02)
    assert abs(mu.mean() - mu_ref.mean()) < 2e-3
    # the pixels without measured variance (every oracle sample identical, e.g. a path that always ends on the same sky colour is not one:
    # the jitter moves it) must agree to the rounding of the f32 sums
    if (~has_var).any():
        assert (diff[~has_var] <= 1e-4 + 1e-4 * mu_ref[~has_var]).mean() > 0.9


def test_fast_and_exact_builds_agree_statistically(api, A, rtiow):
    a, _, _ = render(api, rtiow, cam_rtiow(api), 160, 100, 256, flags=A.RENDER_FAST_MATH, seed=5)
    b, _, _ = render(api, rtiow, cam_rtiow(api), 160, 100, 256, flags=0, seed=5)
    # same seed, same draws: most pixels see identical paths; differences come from rounding only
    close = np.isclose(a["sum"], b["sum"], rtol=1e-3, atol=1e-2).all(axis=1).mean()
    rmse = float(np.sqrt(np.mean((a["sum"] / 256.0 - b["sum"] / 256.0) ** 2)))
    print("fast-math vs default, same seed: %.4f of pixels within 1e-3, RMSE %.5f" % (close, rmse))
    assert close > 0.5 and rmse < 0.02   # a rounding difference flips a few discrete path decisions per pixel
    assert abs(a["sum"].mean() - b["sum"].mean()) / b["sum"].mean() < 2e-3
    assert (a["bloom"] == b["bloom"]).mean() > 0.99


def test_full_size_c1_properties(api, A, rtiow):
    """BASELINE.json configs[0] at full size (1200x800, 10 spp, depth 50), fast build."""
    W, H, spp = 1200, 800, 10
    px, credited, traced = render(api, rtiow, cam_rtiow(api), W, H, spp)
    assert credited == traced == W * H * spp
    assert (px["n"] == spp).all() and np.isfinite(px["sum"]).all() and (px["sum"] >= 0).all()
    assert (px["sum"] <= spp * 1.0001).all()                       # no emitters: a sample never exceeds 1
    sky = np.isinf(px["avg_depth"])
    assert 0.02 < sky.mean() < 0.5 and (px["bloom"][px["avg_depth"] < 0] == 0).all()
    assert (px["bloom"][~sky] != 0).all()                           # every hit pixel recorded an object id
    top = px["color"].reshape(H, W, 3)[:40].astype(int)
    assert (top[..., 2] >= top[..., 0]).mean() > 0.95               # blue-ish sky gradient at the top (render_thread.rs:118-120)


def test_full_size_passes_and_partitions_are_idempotent(api, A, rtiow):
    """Exact build at 1200x800: one pass == many passes == union of chunk partitions == brute force."""
    W, H, spp = 1200, 800, 4
    cam = cam_rtiow(api)
    ex = 0  # default build = parity arithmetic
    whole, _, _ = render(api, rtiow, cam, W, H, spp, flags=ex, seed=3)
    multi, _, _ = render(api, rtiow, cam, W, H, spp, flags=ex, seed=3, passes=(1, 2, 1))
    assert whole.tobytes() == multi.tobytes()
    parts = [render(api, rtiow, cam, W, H, spp, flags=ex, seed=3, part_rank=k, part_world=3)[0] for k in range(3)]
    merged = parts[0].copy()
    from raytracingoneweekend_b200 import dist as D
    for k in (1, 2):
        own = D.owned_pixel_mask(W * H, k, 3)
        merged[own] = parts[k][own]
    assert merged.tobytes() == whole.tobytes()
    nosmem, _, _ = render(api, rtiow, cam, W, H, spp, flags=ex | A.RENDER_NO_SMEM_SCENE, seed=3)
    assert nosmem.tobytes() == whole.tobytes()
    brute, _, _ = render(api, rtiow, cam, 300, 200, spp, flags=ex | A.RENDER_BRUTE_FORCE, seed=3)
    bvh, _, _ = render(api, rtiow, cam, 300, 200, spp, flags=ex, seed=3)
    assert brute.tobytes() == bvh.tobytes()


def test_spp_slices_merge_through_export_import(api, A, rtiow):
    """The multi-GPU exchange on one GPU: two spp-slice handles -> export -> SUM + OR -> import."""
    import torch
    from raytracingoneweekend_b200 import _lib
    W, H, spp = 200, 120, 16
    cam = cam_rtiow(api, W / H)
    whole, _, _ = render(api, rtiow, cam, W, H, spp, flags=0, seed=4)
    accs, blooms, rs = [], [], []
    for base, count in [(0, 8), (8, 8)]:
        r = api.Renderer(rtiow, cam, api.render_params(W, H, spp, seed=4, flags=0, sample_base=base, sample_count=count))
        r.launch(0); r.sync()
        acc = torch.empty((W * H, 8), dtype=torch.float32, device="cuda")
        bl = torch.empty((W * H,), dtype=torch.int64, device="cuda")
        r.export_accum(acc.data_ptr(), bl.data_ptr())
        accs.append(acc); blooms.append(bl); rs.append(r)
    total = accs[0] + accs[1]
    stacked = torch.stack(blooms, 0).contiguous()
    out_bloom = torch.zeros_like(blooms[0])
    _lib.call("om_bloom_or", out_bloom.data_ptr(), stacked.data_ptr(), 2, W * H, 0)
    torch.cuda.synchronize()
    rs[0].import_accum(total.data_ptr(), out_bloom.data_ptr())
    merged = rs[0].read_pixels()
    for r in rs:
        r.close()
    assert (merged["n"] == spp).all() and (merged["bloom"] == whole["bloom"]).all()
    assert np.allclose(merged["sum"], whole["sum"], rtol=1e-5, atol=1e-6)
    assert (np.abs(merged["color"].astype(int) - whole["color"].astype(int)) <= 1).all()
    fin = np.isfinite(whole["avg_depth"])
    assert (np.isfinite(merged["avg_depth"]) == fin).all() and np.allclose(merged["avg_depth"][fin], whole["avg_depth"][fin], rtol=1e-5)


def test_perf_counters_and_progress(api, A, rtiow):
    cam = cam_rtiow(api)
    r = api.Renderer(rtiow, cam, api.render_params(320, 200, 8, seed=1, flags=A.RENDER_COUNTERS))
    r.launch(3); r.sync()
    assert r.samples_done() == 320 * 200 * 3          # the AtomicU64 of main.rs:148 after 3 of 8 passes
    r.launch(0); r.sync()
    c = r.perf_counters()
    assert c["traced"] == c["credited"] == 320 * 200 * 8
    assert 1.0 < c["segments"] / c["traced"] < 8.0 and c["node_visits"] > c["segments"] and c["prim_tests"] > c["segments"]
    assert c["march_entered"] < c["segments"] and c["sdf_evals"] >= c["march_steps"] > 0
    ms, launches = r.elapsed_ms()
    assert launches == 2 and ms > 0
    r.close()


def test_stream_handoff(api, A, rtiow):
    """om_render_set_stream: the handle works on a caller-owned (torch) stream."""
    import torch
    s = torch.cuda.Stream()
    cam = cam_rtiow(api)
    a, _, _ = render(api, rtiow, cam, 100, 60, 4, flags=0, seed=6)
    r = api.Renderer(rtiow, cam, api.render_params(100, 60, 4, seed=6, flags=0))
    r.set_stream(s.cuda_stream)
    r.launch(0)
    s.synchronize()
    assert r.read_pixels().tobytes() == a.tobytes()
    r.set_stream(None)
    r.close()


@pytest.mark.parametrize("kind,seed,n_param,W,H,spp", [("random_scene", 1, 0, 1920, 1080, 6), ("glass_heavy", 4, 0, 3840, 2160, 2),
                                                       ("marched", 3, 0, 1920, 1080, 2), ("many_spheres", 5, 200000, 1920, 1080, 3)])
def test_full_resolution_pipelines_agree_bit_for_bit(api, A, kind, seed, n_param, W, H, spp):
    """BASELINE.json's image sizes (configs 2-5; 2*10^5 of config 5's 10^6 spheres): the persistent kernel and the
    wavefront pipeline must give the same Vec<Pixel>, every pixel must hold exactly spp samples, and the counters
    must agree (a checksum of the whole frame, not a sample of it)."""
    fz = api.FrozenHittableList.generate(kind, seed, n_param)
    cam = api.Camera.new((13, 2, 3) if kind != "many_spheres" else (60, 12, 14), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    out = []
    pipelines = [A.RENDER_PERSISTENT, A.RENDER_WAVEFRONT] + ([A.RENDER_POOL] if kind != "many_spheres" else [])  # the pool needs the scene in shared memory
    for fl in pipelines:
        r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=17, flags=fl))
        r.launch(0); r.sync()
        px = r.read_pixels()
        out.append((px, r.samples_done(), r.samples_traced()))
        r.close()
    a = out[0]
    assert (a[0]["n"] == spp).all() and a[1] == a[2] == W * H * spp
    for b in out[1:]:
        assert a[1:] == b[1:]
        for f in ("sum", "avg_depth"):
            assert (a[0][f].view(np.uint32) == b[0][f].view(np.uint32)).all(), f
        for f in ("n", "bad_avgs", "bloom", "color"):
            assert (a[0][f] == b[0][f]).all(), f
