"""The N > 1 path on CPU: world_size-2 gloo.  The per-rank renders come from the oracle (this is a
test), the partition / merge / collective plumbing is the product's raytracingoneweekend_b200.dist."""
import os
import socket

import numpy as np
import pytest

from conftest import ROOT  # noqa: F401

W, H, SPP = 40, 30, 6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, q):
    import torch
    import torch.distributed as dist
    from raytracingoneweekend_b200 import _abi as A, api, dist as D
    from oracle import binding as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cam = O.camera_new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    scene = O.OracleScene().random_scene(1)
    if mode == "spp":
        base, count = D.spp_slices(SPP, world)[rank]
        p = api.render_params(W, H, SPP, seed=5, sample_base=base, sample_count=count)
    else:
        p = api.render_params(W, H, SPP, seed=5, flags=A.RENDER_ADAPTIVE, part_rank=rank, part_world=world)
    px = scene.render(cam, p, nthreads=1)["pixels"]
    acc, bloom = D.pixels_to_accum(px)
    acc_t, bloom_t = torch.from_numpy(acc), torch.from_numpy(bloom)
    D.reduce_to_rank0(acc_t, bloom_t)
    if rank == 0:
        q.put((acc_t.numpy().copy(), bloom_t.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["spp", "chunks"])
def test_two_rank_split_and_merge(built, mode):
    import torch.multiprocessing as mp
    from raytracingoneweekend_b200 import _abi as A, api, dist as D
    from oracle import binding as O

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    acc, bloom = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    cam = O.camera_new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    scene = O.OracleScene().random_scene(1)
    flags = A.RENDER_ADAPTIVE if mode == "chunks" else 0
    full = scene.render(cam, api.render_params(W, H, SPP, seed=5, flags=flags), nthreads=2)["pixels"]
    facc, fbloom = D.pixels_to_accum(full)
    assert (bloom == fbloom).all()                       # OR of the per-rank masks
    assert (acc[:, 3] == facc[:, 3]).all()               # n
    if mode == "chunks":                                 # disjoint pixels: the SUM reduce is exact
        assert (acc.view(np.uint32) == facc.view(np.uint32)).all()
    else:                                                # same samples, different summation order
        assert np.allclose(acc[:, :3], facc[:, :3], rtol=1e-5, atol=1e-6)
        fin = np.isfinite(facc[:, 4])
        assert (np.isfinite(acc[:, 4]) == fin).all() and np.allclose(acc[fin, 4], facc[fin, 4], rtol=1e-5)


def test_partition_helpers():
    from raytracingoneweekend_b200 import dist as D
    assert D.spp_slices(1024, 8) == [(128 * k, 128) for k in range(8)]
    s = D.spp_slices(10, 4)
    assert sum(c for _, c in s) == 10 and [b for b, _ in s] == [0, 3, 6, 8]
    m = [D.owned_pixel_mask(10000, r, 3) for r in range(3)]
    assert (np.sum(m, axis=0) == 1).all()                # every pixel has exactly one owner
    assert m[1][2730] and m[2][5460] and m[0][8190] and not m[0][2730]   # main.rs:172-189 round robin
    a = np.zeros(4, dtype=np.dtype([("sum", "<f4", (3,)), ("n", "<u4"), ("avg_depth", "<f4"), ("bad_avgs", "<u4"), ("bloom", "<u8"), ("color", "u1", (3,)), ("_pad", "u1", (5,))]))
    a["n"] = [0, 1, 2, 3]; a["avg_depth"] = [0, np.inf, 2.0, 1.0]; a["bloom"] = [0, 1, 2, 2**63]
    acc, bloom = D.pixels_to_accum(a)
    assert acc[0, 4] == 0 and np.isinf(acc[1, 4]) and acc[2, 4] == 4.0 and bloom.dtype == np.int64
    acc2, bloom2 = D.merge_accum([acc, acc], [bloom, bloom])
    assert (acc2[:, 3] == [0, 2, 4, 6]).all() and (bloom2 == bloom).all()


# ---- adaptive scheduling across ranks (dist.AdaptiveScheduler): the host logic and the exact merge -------------------
def test_deal_chunks_balances_and_is_deterministic():
    from raytracingoneweekend_b200 import dist as D
    rng = np.random.default_rng(3)
    live = rng.integers(0, 2731, 760).astype(np.uint32)
    live[rng.random(760) < 0.6] = 0  # most chunks already retired
    for world in (1, 2, 3, 8):
        owners, load = D.deal_chunks(live, world)
        owners2, _ = D.deal_chunks(live.copy(), world)
        assert (owners == owners2).all() and owners.dtype == np.uint8 and owners.max() < world
        assert load.sum() == live.sum() and load.max() - load.min() <= 2730
        dead = live == 0
        assert (owners[dead] == (np.arange(760) % world)[dead]).all()  # retired chunks keep the round-robin owner
        assert [int(live[owners == k].sum()) for k in range(world)] == load.tolist()


def _state_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from raytracingoneweekend_b200 import dist as D

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(11)  # the same "true" frame on both ranks
    npx = 3 * 2730 + 17
    full = rng.integers(-2**31, 2**31 - 1, (npx, 8), dtype=np.int64).astype(np.int32)
    full[5, 0] = np.array([0x80000000], np.uint32).view(np.int32)[0]  # -0.0
    full[6, 1] = np.array([0x7FC00001], np.uint32).view(np.int32)[0]  # a NaN payload
    live = np.array([100, 0, 2730, 17], np.uint32)
    owners, _ = D.deal_chunks(live, world)
    mine = owners[np.arange(npx) // 2730] == rank
    state = torch.from_numpy(np.where(mine[:, None], full, 0).astype(np.int32))
    dist.all_reduce(state, op=dist.ReduceOp.SUM)  # what AdaptiveScheduler does with the exported states
    if rank == 0:
        q.put((state.numpy().copy(), full))
    dist.barrier()
    dist.destroy_process_group()


def test_state_merge_by_integer_sum_is_bit_exact(built):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_state_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged, full = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
    assert (merged == full).all()
