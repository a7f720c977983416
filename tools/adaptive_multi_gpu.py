"""Adaptive render over N GPUs with the live pixels re-dealt between passes (dist.AdaptiveScheduler), launched with
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/adaptive_multi_gpu.py [C2] [--spp 256]
Rank 0 also renders the frame alone and checks the scheduled result against it bit for bit; prints one JSON line."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import torch.distributed as dist
import bench
from raytracingoneweekend_b200 import api, _abi as A, dist as D

ap = argparse.ArgumentParser(); ap.add_argument("workload", nargs="?", default="C2"); ap.add_argument("--spp", type=int, default=256); ap.add_argument("--pass-spp", type=int, default=16)
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
kind, seed, n_param, W, H, _, aspect = bench.WORKLOADS[a.workload]
fz = api.FrozenHittableList.generate(kind, seed, n_param)
cam = bench.make_camera(api, aspect)
out = {"workload": a.workload, "gpus": world, "spp_max": a.spp, "pass_spp": a.pass_spp}
for name, rebalance in (("static round-robin chunks", False), ("re-dealt every pass", True)):
    s = D.AdaptiveScheduler(fz, cam, W, H, a.spp, rank, world, pass_spp=a.pass_spp, device=local, rebalance=rebalance, seed=1)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    credited = s.run()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    px = s.r.read_pixels() if rank == 0 else None
    worst = max((max(l) * world / max(1, sum(l))) for _, l in s.history)
    out[name] = {"seconds": dt, "passes": len(s.history), "credited": credited, "traced": int(px["n"].sum()) if rank == 0 else None,
                 "worst live-pixel imbalance (max rank / mean)": worst}
    s.close()
if rank == 0:
    r = api.Renderer(fz, cam, api.render_params(W, H, a.spp, seed=1, flags=A.RENDER_ADAPTIVE, device=local))
    t0 = time.perf_counter()
    for _ in range((a.spp + a.pass_spp - 1) // a.pass_spp):
        r.launch(a.pass_spp)
    ref = r.read_pixels(); dt = time.perf_counter() - t0
    r.close()
    same = all((np.ascontiguousarray(px[f]).view(np.uint8) == np.ascontiguousarray(ref[f]).view(np.uint8)).all() for f in ("sum", "n", "avg_depth", "bad_avgs", "bloom", "color"))
    out["one GPU alone"] = {"seconds": dt, "traced": int(ref["n"].sum())}
    out["bit-identical to one GPU"] = bool(same)
    print(json.dumps(out))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
