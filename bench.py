#!/usr/bin/env python
"""Headline benchmark: Msamples/s of the per-pixel path-trace / ray-march hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2]

One "step" = one full render of the workload (every pixel, every sample) through
render_thread::render's replacement.  Default workload = BASELINE.json configs[1]: the RTIOW
random_scene (main.rs:37-100) at 1920x1080, 1024 spp, depth 50 (fixed spp).  With N > 1 (torchrun,
one rank per GPU) the SAME job is split into spp slices and combined by one NCCL reduce: strong
scaling.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
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
MAX_DEPTH, TMIN, TMAX = 50, 0.001, 100.0  # main.rs:146,205-206
# DRAM bytes per k_render launch measured with ncu, keyed by (workload, spp per launch, GPUs, wavefront)
NCU_DRAM_BYTES = {("C2", 1024, 1, False): 121891584 + 557661952}  # profiles/r1_v1e_c2_1024spp_dram.csv
# ncu --set full of k_render on C2 (16-spp launch, profiles/r1_v1g_c2_k_render.txt): what the FP32 figure does not show
NCU_K_RENDER_C2 = {"source": "profiles/r1_v1g_c2_k_render.txt", "issue_active_pct": 65.1, "pipe_fma_cycles_active_pct": 22.2, "inst_executed_pipe_alu_pct": 53.1,
                   "threads_per_warp_instruction": 13.2, "stall_barrier_per_issue": 3.22, "stall_no_instruction_per_issue": 0.40}

# flops "as the reference executes it" per executed unit of work
FLOPS = dict(raygen=73, accumulate=20, node_visit=48, prim_test=61, sdf_eval=56, normal_eval=210, shade=50)


_OUT = sys.stdout


def flops_per_sample(c, samples):
    """Algorithmic flops of one sample from executed-work counters (OM_RENDER_COUNTERS run)."""
    s = float(samples)
    return (FLOPS["raygen"] + FLOPS["accumulate"] + (c["segments"] * FLOPS["shade"] + c["node_visits"] * FLOPS["node_visit"]
            + c["prim_tests"] * FLOPS["prim_test"] + c["sdf_evals"] * FLOPS["sdf_eval"] + c["normal_evals"] * FLOPS["normal_eval"]) / s)


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


def make_camera(api, aspect):
    return api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)  # main.rs:136-142


def oracle_scene_from(frozen):
    from oracle import binding as B
    o = B.OracleScene()
    for k, arr in frozen.lists().items():
        o.add(k, arr)
    return o


def cpu_reference_rate(api, wl, seconds_target=10.0, threads=0):
    """The reference's CPU algorithm (oracle port: brute-force hit(), num_cpus-1 threads, 2730-pixel
    chunks) on a bounded sample of the workload: every `world`-th chunk at 1 spp.  Returns
    (Msamples/s, threads, description)."""
    from oracle import binding as B
    kind, seed, n_param, W, H, spp, aspect = wl
    if kind == "many_spheres":
        n_param = min(n_param, 10000)  # O(N) per segment: the brute-force reference cannot take 1e6
    fz = api.FrozenHittableList.generate(kind, seed, n_param)
    osc = oracle_scene_from(fz)
    cam = make_camera(api, aspect)
    nthreads = threads or max(1, B.lib().omo_hardware_threads() - 1)
    chunks = (W * H + 2729) // 2730

    def run(world):
        p = api.render_params(W, H, 1, MAX_DEPTH, TMIN, TMAX, seed=1, flags=0, part_rank=0, part_world=world)
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
    if kind == "many_spheres":
        desc += ", N capped at %d spheres" % n_param
    return n / dt / 1e6, nthreads, desc, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (its C++ restatement in
    oracle/, since the Rust crate cannot be built here) on the box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from raytracingoneweekend_b200 import api
    wl = WORKLOADS[args.workload]
    rates, desc, threads = [], "", 0
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    t_all = time.perf_counter()
    for i in range(args.warmup + args.steps):
        rate, threads, desc, dt = cpu_reference_rate(api, wl, seconds_target=per_step)
        if i >= args.warmup:
            rates.append((rate, dt))
    value = float(np.mean([r for r, _ in rates]))
    ms = float(np.mean([dt for _, dt in rates]) * 1e3)
    kind, seed, n_param, W, H, spp, aspect = wl
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "%s: %s %dx%d, %d spp, depth %d" % (args.workload, kind, W, H, spp, MAX_DEPTH), "sampled": desc},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t_all,
    }
    _OUT.write(json.dumps(line) + "\n"); _OUT.flush()


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
    fz = api.FrozenHittableList.generate(kind, seed, n_param)
    cam = make_camera(api, aspect)
    kw = dict(max_depth=MAX_DEPTH, tmin=TMIN, tmax=TMAX, seed=1)
    if args.wavefront:
        kw["flags"] = A.RENDER_WAVEFRONT
    dr = D.DistributedRenderer(fz, cam, W, H, spp, rank, world, mode="spp", device=local, **kw)
    stream = torch.cuda.Stream(device=local)  # a real stream: pointer 0 would mean "the handle's own"
    torch.cuda.set_stream(stream)              # NCCL collectives and the events below use it too
    dr.r.set_stream(stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        dr.r.reset()
        dr.render()
        if world > 1:
            dr.exchange()

    for _ in range(args.warmup):
        step()
    barrier()
    dr.r.elapsed_ms()  # drop warm-up launch timings
    kernels0 = dr.r.perf_counters()["kernels"]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    kern_ms, kern_launches = dr.r.elapsed_ms()
    pcs = dr.r.perf_counters()
    kernels_timed = pcs["kernels"] - kernels0
    wavefront = bool(pcs["flags"] & A.RENDER_WAVEFRONT)
    samples_step = W * H * spp
    value = samples_step * args.steps / (ms_total * 1e-3) / 1e6
    traced = dr.r.samples_traced()

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
        api._lib.call("om_scene_freeze", world_list.h, None, 0)
        d2 = D.DistributedRenderer(world_list, cam, W, H, spp, rank, world, mode="spp", device=local, **kw)
        d2.render()
        if world > 1:
            d2.exchange()
        if rank == 0:
            d2.r.read_pixels(px)
        else:
            d2.r.sync()
        h2d = world_list.stats()["device_bytes"] + spp
        d2.close()

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
        # ---- roofline of the dominant kernel (k_render): executed work x SURVEY §8(d) constants ----
        kwc = dict(kw); kwc["flags"] = kw.get("flags", 0) | A.RENDER_COUNTERS
        pc = api.render_params(W, H, 16, device=local, **kwc)
        rc = api.Renderer(fz, cam, pc)
        rc.launch(0)
        rc.sync()
        c = rc.perf_counters()
        rc.close()
        fps = flops_per_sample(c, c["traced"])
        peak_tf, _ = api.fp32_peak(local)
        launch_ms = kern_ms / max(1, kern_launches)
        samples_launch = W * H * (D.spp_slices(spp, world)[0][1])
        achieved_tf = fps * samples_launch / (launch_ms * 1e-3) / 1e12
        roofline = {"bound": "fp32", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf, "traffic": NCU_DRAM_BYTES.get((args.workload, spp, world, wavefront)),
                    "traffic_source": "profiles/r1_v1e_c2_1024spp_dram.csv: ncu dram__bytes_read.sum + dram__bytes_write.sum of this launch (122 MB read + 558 MB written: the 2 x 66 MB "
                                      "framebuffer pass plus write-backs of the traversal stack); the full section set was captured at 16 spp (profiles/r1_v1g_c2_k_render.txt)" if (args.workload, spp, world, wavefront) in NCU_DRAM_BYTES else None,
                    "peak_source": "FMA microbenchmark measured in this run (MEASURED_PEAKS.json has no FP32 figure)",
                    "flops_per_sample": fps, "kernel": "k_wf_extend + k_wf_shade + k_wf_compact (+ k_render tail)" if wavefront else "k_render", "launch_ms": launch_ms,
                    "per_sample": {k: c[k] / c["traced"] for k in ("segments", "node_visits", "prim_tests", "sdf_evals", "march_steps", "normal_evals")},
                    "accumulate_hbm": {"bytes_per_launch": 64 * W * H, "note": "32 B read + 32 B write per pixel per launch"},
                    "ncu": NCU_K_RENDER_C2 if (args.workload == "C2" and not wavefront) else None}
        cpu = None
        if world == 1 and not args.no_cpu:
            rate, threads, desc, _ = cpu_reference_rate(api, WORKLOADS[args.workload], seconds_target=12.0)
            cpu = {"value": rate, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": desc}
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "%s: %s %dx%d, %d spp, depth %d, fixed spp" % (args.workload, kind, W, H, spp, MAX_DEPTH),
                       "partition": "spp slices + 1 NCCL reduce" if world > 1 else "single GPU", "pipeline": ("wavefront" if wavefront else "persistent kernel") + ("" if args.wavefront else " (library's choice)"),
                       "l2": "scene image %.1f MB %s; per-step traffic beyond it is the %d MB framebuffer, rewritten by reset() every step"
                             % (fz.stats()["device_bytes"] / 1e6, "with its hot part in shared memory" if pcs["scene_in_smem"] else "read through L1/L2 (larger than shared memory)", W * H * 32 // 1000000)},
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": W * H * 40},
            "gpu_launches": int(kernels_timed) + args.steps * (3 if world > 1 else 0),
            "clocks": clocks, "roofline": roofline, "samples_traced_rank0": traced,
        }
        if cpu:
            line["cpu_baseline"] = cpu
    dr.close()
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
    ap.add_argument("--wavefront", action="store_true", help="use the wavefront pipeline (OM_RENDER_WAVEFRONT) instead of the persistent kernel")
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
