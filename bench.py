#!/usr/bin/env python
"""Headline benchmark: Msamples/s of the per-pixel path-trace / ray-march hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2] [--adaptive]

One "step" = one full render of the workload (every pixel, every sample) through
render_thread::render's replacement.  Default workload = BASELINE.json configs[1]: the RTIOW
random_scene (main.rs:37-100) at 1920x1080, 1024 spp, depth 50 (fixed spp).  With N > 1 (torchrun,
one rank per GPU) the SAME job is split into spp slices and combined by one NCCL reduce: strong
scaling.  --adaptive runs the reference's own mode instead (a pixel retires when its 8-bit colour
stood still for five samples, render_thread.rs:23-39,194-200; interleaved 2730-pixel chunks over
the ranks, main.rs:172-189).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import pickle
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene kind, seed, n_param, width, height, spp, camera aspect)
    "C1": ("random_scene", 1, 0, 1200, 800, 10, 1.5),
    "C2": ("random_scene", 1, 0, 1920, 1080, 1024, 16.0 / 9.0),
    "C2classic": ("rtiow_classic", 1, 0, 1920, 1080, 1024, 16.0 / 9.0),
    "C3": ("marched", 3, 0, 1920, 1080, 256, 16.0 / 9.0),
    "C4": ("glass_heavy", 4, 0, 3840, 2160, 512, 16.0 / 9.0),
    "C5": ("many_spheres", 5, 1000000, 3840, 2160, 64, 16.0 / 9.0),
}
LOOKFROM = {"many_spheres": (60, 12, 14)}  # C5's camera is pulled back; everything else uses main.rs:136-142
MAX_DEPTH, TMIN, TMAX = 50, 0.001, 100.0  # main.rs:146,205-206
# ncu figures of the dominant kernel per workload, written by tools/ncu_to_json.py from captures of THIS tree (no constants here)
NCU_FILE = os.path.join(ROOT, "profiles", "ncu_current.json")

# flops "as the reference executes it" per executed unit of work.  A node visit of the four-wide BVH tests
# four child boxes (4 x 24 flop slab tests).
FLOPS = dict(raygen=73, accumulate=20, node_visit=96, prim_test=61, sdf_eval=56, normal_eval=210, shade=50)
# bytes a traversal step fetches (scenes larger than shared memory, SURVEY 8(d) "C5"): 112-byte node, 48-byte world->local rows
# per primitive test, and per resolved hit the 48-byte local->world rows + 32-byte material record
BYTES = dict(node_visit=112, prim_test=48, hit=80)


_OUT = sys.stdout


def flops_per_sample(c, samples):
    """Algorithmic flops of one sample from executed-work counters (OM_RENDER_COUNTERS run)."""
    s = float(samples)
    return (FLOPS["raygen"] + FLOPS["accumulate"] + (c["segments"] * FLOPS["shade"] + c["node_visits"] * FLOPS["node_visit"]
            + c["prim_tests"] * FLOPS["prim_test"] + c["sdf_evals"] * FLOPS["sdf_eval"] + c["normal_evals"] * FLOPS["normal_eval"]) / s)


def bytes_per_sample(c, samples, spp):
    s = float(samples)
    return (c["node_visits"] * BYTES["node_visit"] + c["prim_tests"] * BYTES["prim_test"] + c["segments"] * BYTES["hit"]) / s + 64.0 / spp


def load_ncu(workload):
    try:
        with open(NCU_FILE) as f:
            return json.load(f).get(workload)
    except (OSError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def make_camera(api, aspect, kind=None):
    return api.Camera.new(LOOKFROM.get(kind, (13, 2, 3)), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)  # main.rs:136-142


# ---- the reference arm: oracle only ---------------------------------------------------------------------------------------
_DUMP = """
import pickle, sys
sys.path.insert(0, %r)
from raytracingoneweekend_b200 import api
fz = api.FrozenHittableList.generate(%r, %d, %d)
pickle.dump({k: bytes(v) for k, v in fz.lists().items()}, open(sys.argv[1], "wb"))
"""


def oracle_scene_for(kind, seed, n_param):
    """The workload's scene inside the oracle, WITHOUT loading the product library into this process: random_scene comes from
    the oracle's own generator (oracle/ref_render.hpp, restating main.rs:37-100); the synthetic benchmark scenes exist only in
    the product's generator, so a child process makes them and hands them over as data."""
    from oracle import binding as B
    from raytracingoneweekend_b200 import _abi as A  # ctypes struct definitions only: loads nothing
    if kind == "random_scene":
        return B.OracleScene().random_scene(seed)
    with tempfile.NamedTemporaryFile(suffix=".pkl") as tmp:
        subprocess.check_call([sys.executable, "-c", _DUMP % (ROOT, kind, seed, n_param), tmp.name], stdout=subprocess.DEVNULL)
        lists = pickle.load(open(tmp.name, "rb"))
    types = {"spheres": A.Sphere, "cubes": A.Cube, "triangles": A.Barycentric, "infinite_planes": A.InfinitePlane, "parallelograms": A.Barycentric,
             "marched_spheres": A.MarchedSphere, "marched_boxes": A.MarchedBox, "marched_tori": A.MarchedTorus}
    o = B.OracleScene()
    for k, raw in lists.items():
        T = types[k]
        n = len(raw) // __import__("ctypes").sizeof(T)
        if n:
            o.add(k, (T * n).from_buffer_copy(raw))
    return o


def cpu_reference_rate(wl, seconds_target=10.0, threads=0, osc=None):
    """The reference's CPU algorithm (oracle port: brute-force hit(), num_cpus-1 threads, 2730-pixel chunks, main.rs:170-189) on a
    bounded sample of the workload: every `world`-th chunk at 1 spp (samples/s does not depend on spp).  Returns
    (Msamples/s, threads, description, seconds).  `osc`: an OracleScene that already holds the scene."""
    from oracle import binding as B
    from raytracingoneweekend_b200 import _abi as A
    kind, seed, n_param, W, H, spp, aspect = wl
    capped = kind == "many_spheres" and n_param > 10000
    if capped:
        n_param = 10000  # O(N) per segment: the brute-force reference cannot take 1e6
        osc = None
    if osc is None:
        osc = oracle_scene_for(kind, seed, n_param)
    cam = B.camera_new(LOOKFROM.get(kind, (13, 2, 3)), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)
    nthreads = threads or max(1, B.lib().omo_hardware_threads() - 1)
    chunks = (W * H + 2729) // 2730

    def run(world):
        p = A.RenderParams(W, H, 1, MAX_DEPTH, TMIN, TMAX, 1, 0, 0, 0, 0, world, 0)
        t0 = time.perf_counter()
        r = osc.render(cam, p, nthreads=nthreads)
        return r["traced"], time.perf_counter() - t0

    world = max(1, chunks // max(1, nthreads))  # calibration: about one chunk per thread
    n, dt = run(world)
    rate = n / max(dt, 1e-6)
    want = rate * seconds_target
    world2 = int(max(1, min(world, round(W * H / max(want, 1.0)))))
    if world2 < world:
        n, dt = run(world2)
        world = world2
    desc = "%s %dx%d, 1 spp over every %d-th 2730-pixel chunk (%d samples), depth %d" % (kind, W, H, world, n, MAX_DEPTH)
    if capped:
        desc += ", N capped at %d spheres (the reference scans every object per segment)" % n_param
    return n / dt / 1e6, nthreads, desc, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (its C++ restatement in oracle/, since the Rust crate
    cannot be built here) on the box's host cores.  Loads oracle/libom_oracle.so only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    kind, seed, n_param, W, H, spp, aspect = wl
    osc = None if (kind == "many_spheres" and n_param > 10000) else oracle_scene_for(kind, seed, n_param)
    rates, desc, threads = [], "", 0
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    t_all = time.perf_counter()
    for i in range(args.warmup + args.steps):
        rate, threads, desc, dt = cpu_reference_rate(wl, seconds_target=per_step, osc=osc)
        if i >= args.warmup:
            rates.append((rate, dt))
    value = float(np.mean([r for r, _ in rates]))
    ms = float(np.mean([dt for _, dt in rates]) * 1e3)
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "%s: %s %dx%d, %d spp, depth %d" % (args.workload, kind, W, H, spp, MAX_DEPTH), "sampled": desc,
                   "build": "oracle/Makefile: g++ -O2 -ffp-contract=off (rustc release has no target-cpu=native either)"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t_all,
    }
    _OUT.write(json.dumps(line) + "\n"); _OUT.flush()


def pixels_agree(a, b, exact):
    """merge_check: rank 0's merged frame against a single-GPU render of the same job."""
    out = {"n_equal": bool((a["n"] == b["n"]).all()), "bloom_equal": bool((a["bloom"] == b["bloom"]).all()),
           "color_equal_frac": float((a["color"] == b["color"]).all(axis=1).mean())}
    with np.errstate(invalid="ignore", divide="ignore"):
        rel = np.abs(a["sum"].astype(np.float64) - b["sum"]) / np.maximum(np.abs(b["sum"]), 1e-3)
    out["sum_max_rel"] = float(rel.max())
    out["bit_identical"] = bool(all((np.ascontiguousarray(a[f]).view(np.uint8) == np.ascontiguousarray(b[f]).view(np.uint8)).all()
                                    for f in ("sum", "n", "avg_depth", "bad_avgs", "bloom", "color")))
    # spp slices add a pixel's samples as N partial sums instead of one running sum: two orders of the same n f32 terms differ by
    # at most ~n * 2^-24 relative to the sum of their magnitudes (1024 spp: 6.1e-5); the 8-bit colour may flip where sqrt(sum/n)*256
    # sits on an integer
    n_max = int(b["n"].max()) if b["n"].size else 1
    out["sum_rel_bound"] = max(1e-5, n_max * 2.0 ** -24)
    out["ok"] = out["bit_identical"] if exact else (out["n_equal"] and out["bloom_equal"] and out["color_equal_frac"] >= 0.999 and out["sum_max_rel"] <= out["sum_rel_bound"])
    return out


def run_ours(args):
    import torch
    from raytracingoneweekend_b200 import _abi as A
    from raytracingoneweekend_b200 import api
    from raytracingoneweekend_b200 import dist as D

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the renderer has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    kind, seed, n_param, W, H, spp, aspect = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    freeze_flags = A.FREEZE_GPU_BVH if (kind == "many_spheres" and n_param >= 100000) else 0  # device BVH build above ~1e5 primitives
    fz = api.FrozenHittableList.generate(kind, seed, n_param, flags=freeze_flags)
    cam = make_camera(api, aspect, kind)
    kw = dict(max_depth=MAX_DEPTH, tmin=TMIN, tmax=TMAX, seed=1)
    if args.wavefront:
        kw["flags"] = A.RENDER_WAVEFRONT
    if args.pool:
        kw["flags"] = A.RENDER_POOL
    stream = torch.cuda.Stream(device=local)  # a real stream: pointer 0 would mean "the handle's own"
    torch.cuda.set_stream(stream)              # NCCL collectives and the events below use it too

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the job ------------------------------------------------------------------------------------------------------
    phases = {}     # name -> [ms per timed step] (CUDA events on the stream)
    marks = []      # (name, event) of the current step

    def mark(name):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(stream)
        marks.append((name, ev))

    def make_job(frozen):
        if args.adaptive:
            s = D.AdaptiveScheduler(frozen, cam, W, H, spp, rank, world, pass_spp=args.pass_spp, device=local, rebalance=False, refresh_every=args.refresh_every, **kw)
            s.r.set_stream(stream.cuda_stream)
            return s
        d = D.DistributedRenderer(frozen, cam, W, H, spp, rank, world, mode="spp", device=local, **kw)
        d.use_torch_stream(stream)
        return d

    job = make_job(fz)

    def run_job(j, timed=False):
        j.r.reset()
        if args.adaptive:
            j.credited, j.history = 0, []
            j.run()
            return
        if timed:
            mark("start")
        j.render()
        if timed:
            mark("render")
        if world > 1:
            j.exchange(mark if timed else None)

    def step(timed=False):
        del marks[:]
        run_job(job, timed)
        return list(marks)

    for _ in range(args.warmup):
        step()
    barrier()
    job.r.elapsed_ms()  # drop warm-up launch timings
    pc0 = job.r.perf_counters()
    kernels0, graphs0 = pc0["kernels"], pc0["graph_launches"]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    all_marks = []
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        all_marks.append(step(timed=True))
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    for ms_ in all_marks:
        for (_, a), (name, b) in zip(ms_[:-1], ms_[1:]):
            phases.setdefault(name, []).append(a.elapsed_time(b))
    phase_ms = {k: float(np.mean(v)) for k, v in phases.items()}
    if world > 1:
        t = torch.tensor([ms_total, phase_ms.get("render", 0.0)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, slowest_render = float(t[0].item()), float(t[1].item())
        phase_ms["render_slowest_rank"] = slowest_render
    kern_ms, kern_launches = job.r.elapsed_ms()
    pcs = job.r.perf_counters()
    kernels_timed = pcs["kernels"] - kernels0
    graphs_timed = pcs["graph_launches"] - graphs0
    wavefront = bool(pcs["flags"] & A.RENDER_WAVEFRONT)
    pool = bool(pcs["flags"] & A.RENDER_POOL)
    # samples of one step: fixed spp = every pixel x spp; adaptive = what was actually traced (rank 0 holds the whole frame)
    merged = job.r.read_pixels().copy() if (rank == 0 and (world > 1 or args.adaptive)) else None
    samples_step = int(merged["n"].astype(np.int64).sum()) if (rank == 0 and args.adaptive) else W * H * spp
    if world > 1 and args.adaptive:
        t = torch.tensor([samples_step], dtype=torch.int64, device="cuda")
        dist.broadcast(t, src=0)
        samples_step = int(t.item())
    value = samples_step * args.steps / (ms_total * 1e-3) / 1e6
    traced = job.r.samples_traced()

    # ---- merge_check: the N-GPU frame against one GPU rendering the whole job alone -----------------------------------------
    merge_check = None
    if rank == 0 and (world > 1 or args.adaptive):
        flags1 = kw.get("flags", 0) | (A.RENDER_ADAPTIVE if args.adaptive else 0)
        r1 = api.Renderer(fz, cam, api.render_params(W, H, spp, flags=flags1, device=local, **{k: v for k, v in kw.items() if k != "flags"}))
        if args.adaptive:
            for _ in range((spp + args.pass_spp - 1) // args.pass_spp):
                r1.launch(args.pass_spp)
        else:
            r1.launch(0)
        r1.sync()
        alone = r1.read_pixels()
        r1.close()
        # adaptive frames are bit-identical by construction (one owner per pixel); spp slices add float partial sums in another order
        merge_check = pixels_agree(merged, alone, exact=args.adaptive)
        merge_check["against"] = "one GPU rendering the same job alone (%s)" % ("adaptive, same passes" if args.adaptive else "fixed spp")
        if not args.adaptive:
            merge_check["all_pixels_have_spp"] = bool((merged["n"] == spp).all())

    # ---- end to end through the public API with host buffers ------------------------------------
    lists = fz.lists()
    pinned = torch.empty(W * H * 40, dtype=torch.uint8).pin_memory()
    px = pinned.numpy().view(A.PIXEL_DTYPE)
    h2d = 0

    def e2e_step():
        nonlocal h2d
        world_list = api.FrozenHittableList()  # om_scene_create
        for k, arr in lists.items():
            if len(arr):
                api._lib.call("om_scene_add_" + k, world_list.h, arr, len(arr))
        api._lib.call("om_scene_freeze", world_list.h, None, freeze_flags)
        j2 = make_job(world_list)
        run_job(j2)
        if rank == 0:
            j2.r.read_pixels(px)
        else:
            j2.r.sync()
        h2d = world_list.stats()["device_bytes"] + spp
        j2.close()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = samples_step * args.steps / e2e_s / 1e6

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel: executed work x SURVEY §8(d) constants ----
        kwc = dict(kw); kwc["flags"] = kw.get("flags", 0) | A.RENDER_COUNTERS
        spp_c = min(spp, 16)
        rc = api.Renderer(fz, cam, api.render_params(W, H, spp_c, device=local, **kwc))
        rc.launch(0)
        rc.sync()
        c = rc.perf_counters()
        rc.close()
        per_sample = {k: c[k] / c["traced"] for k in ("segments", "node_visits", "prim_tests", "sdf_evals", "march_steps", "normal_evals")}
        in_smem = bool(pcs["scene_in_smem"])
        ncu = load_ncu(args.workload + ("_pool" if pool else ""))
        kernel = ("k_wf_extend + k_wf_shade + k_wf_compact (+ k_render tail)" if wavefront else ("k_render_pool" if pool else "k_render"))
        launch_ms = kern_ms / max(1, kern_launches)
        samples_launch = samples_step / max(1, kern_launches // max(1, args.steps)) if args.adaptive else W * H * (D.spp_slices(spp, world)[0][1])
        traffic = None
        if ncu and ncu.get("dram_bytes") and ncu["dram_bytes"].get("spp") == (spp if world == 1 else -1):
            traffic = ncu["dram_bytes"]["read"] + ncu["dram_bytes"]["write"]
        if in_smem:  # C1-C4: the scene lives in shared memory, the path is FP32 / issue bound
            fps = flops_per_sample(c, c["traced"])
            peak_tf, _ = api.fp32_peak(local)
            achieved = fps * samples_launch / (launch_ms * 1e-3) / 1e12
            roofline = {"bound": "fp32", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf,
                        "peak_source": "FMA microbenchmark measured in this run (MEASURED_PEAKS.json has no FP32 figure)", "flops_per_sample": fps}
        else:        # C5: nodes and primitives are fetched through L2 / HBM: bytes against the measured HBM peak
            bps = bytes_per_sample(c, c["traced"], spp)
            peaks = {}
            try:
                peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            except (OSError, ValueError):
                pass
            peak_gbs = float(peaks.get("hbm_gbs", 6550.4))
            achieved = bps * samples_step / (ms_total / args.steps * 1e-3) / 1e9
            roofline = {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback 6550.4",
                        "bytes_per_sample": bps, "note": "bytes the traversal asks for (112 B per node visit, 48 B per sphere test, 80 B per hit, 64 B framebuffer per pixel); "
                                                         "the 16-byte bounding ball a visit also reads in scenes with per-ray box padding is not counted; the 180 MB scene image is mostly served by the 126 MB L2, so DRAM traffic is lower than this"}
        roofline.update({"traffic": traffic, "traffic_source": (ncu or {}).get("dram_bytes", {}).get("source") if traffic else None, "kernel": kernel,
                         "launch_ms": launch_ms, "per_sample": per_sample,
                         "accumulate_hbm": {"bytes_per_launch": 64 * W * H, "note": "32 B read + 32 B write per pixel per launch"},
                         "ncu": {k: v for k, v in (ncu or {}).items() if k != "dram_bytes"} or None})
        cpu = None
        if world == 1 and not args.no_cpu:
            from oracle import binding as B
            osc = None
            if not (kind == "many_spheres" and n_param > 10000):
                osc = B.OracleScene()
                for k, arr in lists.items():
                    osc.add(k, arr)
            rate, threads, desc, _ = cpu_reference_rate(WORKLOADS[args.workload], seconds_target=12.0, osc=osc)
            cpu = {"value": rate, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": desc}
        mode = ("adaptive (bad_avgs >= 5 retires a pixel), max %d spp, passes of %d, pixel states merged %s" % (spp, args.pass_spp, "after the last pass" if args.refresh_every == 0 else "every %d pass(es)" % args.refresh_every)) if args.adaptive else "fixed spp"
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "%s: %s %dx%d, %d spp, depth %d, %s" % (args.workload, kind, W, H, spp, MAX_DEPTH, mode),
                       "partition": ("interleaved 2730-pixel chunks + 1 all-reduce of the pixel states per pass" if args.adaptive else "spp slices + 1 NCCL reduce") if world > 1 else "single GPU",
                       "pipeline": ("wavefront" if wavefront else ("path pool" if pool else "persistent kernel")) + ("" if (args.wavefront or args.pool) else " (library's choice)"),
                       "l2": "scene image %.1f MB %s; per-step traffic beyond it is the %d MB framebuffer, rewritten by reset() every step"
                             % (fz.stats()["device_bytes"] / 1e6, "with its hot part in shared memory" if in_smem else "read through L1/L2 (larger than shared memory)", W * H * 32 // 1000000),
                       "tile_order": "the timed steps reuse the tile-cost order that the first warm-up step measured with a pilot launch of 1/64 of the samples; the e2e figure creates a fresh handle per step and pays for the pilot every time"},
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": W * H * 40},
            "gpu_launches": int(kernels_timed) + args.steps * (3 if (world > 1 and not args.adaptive) else 0),
            "graph_launches": int(graphs_timed),  # of gpu_launches' kernels, how many CUDA-graph launches carried them (wavefront: 144 kernels each)
            "clocks": clocks, "roofline": roofline, "samples_traced_rank0": traced, "samples_per_step": samples_step,
        }
        if phase_ms:
            line["phase_ms"] = phase_ms  # CUDA events on rank 0's stream, mean over the timed steps; render_slowest_rank = max over ranks
        if merge_check:
            line["merge_check"] = merge_check
        if cpu:
            line["cpu_baseline"] = cpu
    job.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line:
        _OUT.write(json.dumps(line) + "\n"); _OUT.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (a smaller number is NOT the headline config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--wavefront", action="store_true", help="use the global wavefront pipeline (OM_RENDER_WAVEFRONT) instead of the library's choice")
    ap.add_argument("--pool", action="store_true", help="use the CTA-resident wavefront (OM_RENDER_POOL)")
    ap.add_argument("--refresh-every", type=int, default=0, help="--adaptive on N GPUs: merge the pixel states on every rank after every k-th pass (1 = the viewer's refresh); 0 = once, after the last pass")
    ap.add_argument("--adaptive", action="store_true", help="the reference's adaptive mode (OM_RENDER_ADAPTIVE) through dist.AdaptiveScheduler; value counts the samples actually traced")
    ap.add_argument("--pass-spp", type=int, default=16, help="--adaptive: samples per pass between retirement checks / exchanges")
    args = ap.parse_args()
    # ONE JSON line on stdout: native libraries print there too (NCCL's version banner under NCCL_DEBUG), so file
    # descriptor 1 is pointed at stderr for the run and the line goes to a private copy of the real stdout.
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
