"""Multi-GPU split of render_thread::render: one process per GPU, torch.distributed for plumbing.

The reference splits pixels over CPU threads in 2730-pixel chunks round-robin (main.rs:172-189)
and never exchanges anything (shared memory).  Here every rank holds the whole (tiny) scene and
renders either
  * an spp slice  — samples [base, base+count) of every pixel (fixed-spp renders; perfectly
    balanced), or
  * its interleaved chunks — the reference's own split, needed for adaptive retirement —
and ONE exchange at the end brings the accumulators to rank 0: a SUM reduce of
{sum.rgb, n, n*avg_depth, bad_avgs} and a gather + bitwise OR of the 64-bit bloom masks (NCCL has
no OR).  The merge arithmetic is plain tensor code so the gloo/CPU tests exercise the same path.
"""
import os

import numpy as np

from . import _abi as A


def spp_slices(spp_total: int, world: int):
    """[(base, count)] per rank; counts differ by at most 1 and sum to spp_total."""
    q, r = divmod(spp_total, world)
    out, base = [], 0
    for k in range(world):
        c = q + (1 if k < r else 0)
        out.append((base, c))
        base += c
    return out


def owned_pixel_mask(npx: int, rank: int, world: int):
    """Interleaved-chunk ownership, main.rs:172-189 with thread id -> rank."""
    idx = np.arange(npx, dtype=np.int64)
    return (idx // A.CHUNK_PIXELS) % max(world, 1) == rank


def pixels_to_accum(px):
    """om_pixel array -> (acc [npx,8] float32, bloom [npx] int64) in the export layout of
    om_render_export_accum: {sum.r,sum.g,sum.b,n, n*avg_depth, bad_avgs, 0, 0}."""
    n = px["n"].astype(np.float32)
    acc = np.zeros((px.shape[0], 8), np.float32)
    acc[:, 0:3] = px["sum"]
    acc[:, 3] = n
    with np.errstate(invalid="ignore"):
        acc[:, 4] = np.where(n > 0, n * px["avg_depth"], 0.0)
    acc[:, 5] = px["bad_avgs"].astype(np.float32)
    return acc, px["bloom"].view(np.int64).copy()


def merge_accum(acc_list, bloom_list):
    """What the collective computes: elementwise SUM of acc, bitwise OR of bloom."""
    acc = np.sum(np.stack(acc_list, 0), axis=0, dtype=np.float32)
    bloom = np.bitwise_or.reduce(np.stack(bloom_list, 0), axis=0)
    return acc, bloom


def reduce_to_rank0(acc_t, bloom_t, group=None, gather_buf=None, mark=None, renderer=None):
    """acc_t: float32 tensor [npx,8]; bloom_t: int64 tensor [npx], both on this rank's device.
    After the call rank 0 holds the merged values in place.  One reduce + one gather.
    gather_buf: rank 0's [world, npx] int64 landing buffer (the gather writes its rows directly; allocated here if None).
    mark(name): optional callback after each step (per-phase timing).
    renderer: an api.Renderer whose stream is torch's current stream; the OR then runs there (om_render_bloom_or)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return acc_t, bloom_t
    if os.environ.get("OM_DIST_ALLREDUCE"):  # A/B switch: an all-reduce can ride NVSwitch's in-switch reduction where a rooted reduce is a tree
        dist.all_reduce(acc_t, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.reduce(acc_t, dst=0, op=dist.ReduceOp.SUM, group=group)
    if mark:
        mark("reduce")
    rows = None
    if rank == 0:
        if gather_buf is None:
            gather_buf = torch.empty((world, bloom_t.numel()), dtype=bloom_t.dtype, device=bloom_t.device)
        rows = [gather_buf[k] for k in range(world)]  # views of one buffer: the ranks' rows land side by side, no stacking copy
    dist.gather(bloom_t, gather_list=rows, dst=0, group=group)
    if rank == 0:
        if gather_buf.is_cuda and renderer is not None:
            renderer.bloom_or(bloom_t.data_ptr(), gather_buf.data_ptr(), world)  # on the handle's stream, in order with the gather
        elif gather_buf.is_cuda:
            from . import _lib
            torch.cuda.synchronize(bloom_t.device)  # om_bloom_or runs on the legacy stream: the gather must have landed first
            _lib.call("om_bloom_or", bloom_t.data_ptr(), gather_buf.data_ptr(), world, bloom_t.numel(), bloom_t.device.index)
        else:
            out = gather_buf[0].clone()
            for k in range(1, world):
                out |= gather_buf[k]
            bloom_t.copy_(out)
    if mark:
        mark("gather_or")
    return acc_t, bloom_t


class DistributedRenderer:
    """One rank's share of a render plus the final exchange.  `mode` is "spp" or "chunks"."""

    def __init__(self, frozen, cam, width, height, spp, rank, world, mode="spp", device=None, **kw):
        import torch
        from . import api

        self.torch = torch
        self.rank, self.world, self.mode = rank, world, mode
        self.device = torch.cuda.current_device() if device is None else device
        flags = kw.pop("flags", 0)
        if mode == "spp":
            base, count = spp_slices(spp, world)[rank]
            if count == 0:
                raise ValueError("fewer samples than ranks")
            p = api.render_params(width, height, spp, flags=flags, sample_base=base, sample_count=count, device=self.device, **kw)
        else:
            p = api.render_params(width, height, spp, flags=flags, part_rank=rank, part_world=world, device=self.device, **kw)
        self.params = p
        self.r = api.Renderer(frozen, cam, p)
        npx = width * height
        dev = torch.device("cuda", self.device)
        self.acc = torch.empty((npx, 8), dtype=torch.float32, device=dev)
        self.bloom = torch.empty((npx,), dtype=torch.int64, device=dev)
        self.gather_buf = torch.empty((world, npx), dtype=torch.int64, device=dev) if (rank == 0 and world > 1) else None
        self.shares_torch_stream = False

    def use_torch_stream(self, stream):
        """Run the handle's kernels on `stream` (a torch.cuda.Stream that is also torch's current stream): the collectives are
        ordered against it by torch, so exchange() needs no device-wide synchronize between the collective and the import."""
        self.r.set_stream(stream.cuda_stream)
        self.shares_torch_stream = True

    def render(self):
        self.r.launch(0)

    def exchange(self, mark=None):
        """export -> collective -> import on rank 0.  Rank 0 holds the full image once its stream has run the import.
        mark(name) is called after each phase (export, reduce, gather_or, import) for per-phase timing."""
        self.r.export_accum(self.acc.data_ptr(), self.bloom.data_ptr())
        if mark:
            mark("export")
        if not self.shares_torch_stream and self.world > 1:
            self.r.sync()  # the export ran on the handle's own stream, which torch knows nothing about
        reduce_to_rank0(self.acc, self.bloom, gather_buf=self.gather_buf, mark=mark, renderer=self.r if self.shares_torch_stream else None)
        if self.rank == 0 and self.world > 1:
            if not self.shares_torch_stream:
                self.torch.cuda.synchronize(self.device)
            self.r.import_accum(self.acc.data_ptr(), self.bloom.data_ptr())
        if mark:
            mark("import")

    def close(self):
        self.r.close()


# ---- adaptive sampling across GPUs -------------------------------------------------------------
def deal_chunks(live_counts, world):
    """Chunk -> rank table for the next pass from the live pixels per chunk (identical on every rank): largest chunk
    first onto the rank with the fewest live pixels so far (ties: lower rank), which balances the counts and scatters
    every rank over the image like the reference's round-robin does (main.rs:172-189).  Chunks without live pixels
    keep the round-robin owner; nobody renders them."""
    counts = np.asarray(live_counts, dtype=np.int64)
    owners = (np.arange(counts.shape[0]) % max(world, 1)).astype(np.uint8)
    load = np.zeros(max(world, 1), np.int64)
    for c in np.argsort(-counts, kind="stable"):
        if counts[c] == 0:
            break
        k = int(np.argmin(load))
        owners[c] = k
        load[k] += counts[c]
    return owners, load


class AdaptiveScheduler:
    """The reference's adaptive render (a pixel retires once its 8-bit colour stood still for five samples,
    render_thread.rs:31-38,194-200) spread over `world` GPUs with the live pixels re-dealt between passes.

    Every rank holds the whole frame.  Per pass: the same chunk->rank table on every rank, `pass_spp` samples for the live
    pixels of the chunks a rank owns, then the raw 32-byte pixel states are combined by an integer SUM all-reduce (every
    pixel has exactly one owner, the others contribute zeros, so the sum is the owner's bits) and loaded into every handle.
    The image is bit-identical to a single-GPU adaptive render; rank 0 can read it after any pass (`on_pass`)."""

    def __init__(self, frozen, cam, width, height, spp, rank, world, pass_spp=8, device=None, rebalance=False, refresh_every=1, **kw):
        """refresh_every: the pixel states are merged on every rank after every k-th pass (1 = the viewer's refresh after each pass,
        the reference's main loop reads the shared pixels whenever it redraws); 0 = only once, after the last pass.  Between merges a
        rank keeps rendering its own chunks from its own state and the ranks exchange two integers per pass (samples credited,
        live pixels).  Re-dealing (`rebalance`) moves pixels between ranks and therefore merges after every pass."""
        import torch
        from . import api

        self.torch = torch
        self.rank, self.world, self.pass_spp, self.rebalance, self.spp = rank, world, max(1, pass_spp), rebalance, spp
        self.refresh_every = 1 if rebalance else max(0, int(refresh_every))
        self.device = torch.cuda.current_device() if device is None else device
        flags = kw.pop("flags", 0) | A.RENDER_ADAPTIVE
        self.r = api.Renderer(frozen, cam, api.render_params(width, height, spp, flags=flags, device=self.device, **kw))
        self.state = torch.empty((width * height, 8), dtype=torch.int32, device=torch.device("cuda", self.device))
        self.owners = (np.arange(self.r.n_chunks) % max(world, 1)).astype(np.uint8)
        self.credited = 0
        self.merges = 0
        self.history = []  # per pass: (live pixels before the pass, live pixels per rank)

    def _all_reduce(self, t):
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

    def merge(self):
        """Every rank's handle gets the whole frame: the raw states of the owned chunks, summed over the ranks (one owner per pixel)."""
        self.r.export_state(self.state.data_ptr(), only_owned=True)
        self._all_reduce(self.state)
        if self.world > 1:
            self.torch.cuda.synchronize(self.device)
        self.r.import_state(self.state.data_ptr())
        self.merges += 1

    def run(self, on_pass=None):
        torch = self.torch
        total = self.r.npx * self.spp
        own = None
        live = self.r.live_per_chunk()
        load = np.array([int(live[self.owners == k].sum()) for k in range(self.world)])  # (every rank starts from the same state)
        done_spp, seen, passes, merged_at = 0, 0, 0, 0
        self.r.set_chunk_owners(self.owners, self.rank)
        while done_spp < self.spp and int(load.sum()) > 0:
            if self.rebalance:
                self.owners, load = deal_chunks(live, self.world)
                self.r.set_chunk_owners(self.owners, self.rank)
                own = None
            if own is None:
                own = self.owners == self.rank
            self.history.append((int(load.sum()), load.tolist()))
            self.r.launch(self.pass_spp)
            passes += 1
            done_spp += self.pass_spp
            merge_now = self.refresh_every > 0 and passes % self.refresh_every == 0
            if merge_now:
                self.merge()
                merged_at = passes
            mine = self.r.samples_done()
            live = self.r.live_per_chunk()
            # two integers per rank and pass: the samples credited (the AtomicU64 of main.rs:148) and the live pixels of its own chunks
            box = torch.zeros((self.world, 2), dtype=torch.int64, device=self.state.device)
            box[self.rank, 0] = mine - seen
            box[self.rank, 1] = int(live[own].sum())
            seen = mine
            self._all_reduce(box)
            got = box.cpu().numpy()
            self.credited += int(got[:, 0].sum())
            load = got[:, 1].astype(np.int64)
            if on_pass is not None:
                on_pass(self)
        if merged_at != passes and self.world > 1:
            self.merge()  # the frame after the last pass
        assert self.credited <= total
        return self.credited

    def close(self):
        self.r.close()
