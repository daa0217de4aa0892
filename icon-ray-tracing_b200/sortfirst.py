"""Sort-first multi-GPU frames: the image is cut into 16x16 tiles, each rank renders a disjoint
tile set against its own full replica of the grid, finished tiles are gathered on rank 0 with NCCL send/recv
over NVLink, and the next frame's tile sets are cut from the per-tile sample counts of this frame.

The reference has no multi-GPU path; there is nothing to mirror here except the frame itself,
which must be bit-identical to the single-GPU frame (no cross-tile blending).

torch / torch.distributed are plumbing only (device buffers, NCCL bootstrap); rendering and the tile scatter are
the library's own kernels behind the C ABI.
"""
import numpy as np

TILE = 16


def morton_order(tiles_x, tiles_y):
    """tile ids (row-major id = ty*tiles_x + tx) sorted along a Z-curve: neighbouring tiles in the list are
    neighbours on screen, so a contiguous slice of the list is a compact screen region (rays that share cells)."""
    ty, tx = np.divmod(np.arange(tiles_x * tiles_y, dtype=np.int64), tiles_x)

    def spread(v):
        v = v & 0xFFFF
        v = (v | (v << 8)) & 0x00FF00FF
        v = (v | (v << 4)) & 0x0F0F0F0F
        v = (v | (v << 2)) & 0x33333333
        v = (v | (v << 1)) & 0x55555555
        return v

    key = spread(tx) | (spread(ty) << 1)
    return np.argsort(key, kind="stable").astype(np.int32)


def assign_tiles(order, costs, n_ranks):
    """Cut the Z-ordered tile list into n_ranks contiguous slices of (nearly) equal cost.

    costs: per-tile sample counts of the previous frame indexed by tile id, or None for the first frame
    (equal-count slices).  Deterministic: every rank computes the same partition from the same all-reduced
    costs.  Returns a list of int32 arrays (tile ids per rank); every tile appears exactly once.
    """
    n = len(order)
    if costs is None:
        w = np.ones(n, np.float64)
    else:
        w = np.asarray(costs, np.float64)[order] + 1.0      # +1: a tile is never free (ray set-up, empty pixels)
    cum = np.cumsum(w)
    total = cum[-1]
    cuts = np.searchsorted(cum, total * np.arange(1, n_ranks) / n_ranks, side="left")
    cuts = np.concatenate([[0], np.minimum(cuts + 1, n), [n]])
    cuts = np.maximum.accumulate(cuts)
    return [order[cuts[r]:cuts[r + 1]].copy() for r in range(n_ranks)]


def assign_tiles_dealt(order, costs, n_ranks, block=16):
    """Deal blocks of `block` Z-consecutive tiles (16 = a 64x64-pixel square) to the ranks: blocks sorted by cost, heaviest
    first, dealt in snake order (0..n-1, n-1..0, ...), so that every rank receives the same mix of heavy (limb), medium and
    empty screen regions.  Unlike contiguous slices this also spreads the frame's longest rays -- serial chains of
    dependent fetches that bound a rank's frame time from below -- over all ranks.  Deterministic (stable sorts); each
    rank's tiles stay in Z order.  Returns a list of int32 arrays; every tile appears exactly once."""
    n = len(order)
    nb = (n + block - 1) // block
    if costs is None:
        w = np.ones(n, np.float64)
    else:
        w = np.asarray(costs, np.float64)[order] + 1.0
    pad = nb * block - n
    bw = np.concatenate([w, np.zeros(pad)]).reshape(nb, block).sum(1)
    rank_of_sorted = np.arange(nb) % (2 * n_ranks)
    rank_of_sorted = np.where(rank_of_sorted < n_ranks, rank_of_sorted, 2 * n_ranks - 1 - rank_of_sorted)
    by_cost = np.argsort(-bw, kind="stable")
    rank_of_block = np.empty(nb, np.int64)
    rank_of_block[by_cost] = rank_of_sorted
    # rank-major tile list, Z order kept inside a rank: sort the BLOCKS by rank (stable), then expand to tiles
    blocks = np.argsort(rank_of_block.astype(np.int16), kind="stable")
    idx = (blocks[:, None] * block + np.arange(block)[None, :]).reshape(-1)
    ranks = np.repeat(rank_of_block[blocks], block)
    keep = idx < n                                               # (the last block may be ragged)
    tiles, ranks = order[idx[keep]], ranks[keep]
    cuts = np.concatenate([[0], np.cumsum(np.bincount(ranks, minlength=n_ranks))])
    return [tiles[cuts[r]:cuts[r + 1]] for r in range(n_ranks)]


class SortFirstFrame:
    """One rank's side of a sort-first frame.  `ctx` is this rank's icb.Context with grid/TF/camera set;
    `dist` is torch.distributed (initialised, nccl for GPUs; gloo works for the host logic with cpu tensors)."""

    def __init__(self, ctx, width, height, rank, world, torch, dist, device):
        self.ctx, self.W, self.H, self.rank, self.world = ctx, width, height, rank, world
        self.torch, self.dist, self.device = torch, dist, device
        self.tx, self.ty = (width + TILE - 1) // TILE, (height + TILE - 1) // TILE
        self.ntiles = self.tx * self.ty
        self.order = morton_order(self.tx, self.ty)
        self.costs = None
        self.parts = assign_tiles(self.order, None, world)
        i32 = torch.int32
        self.costs_dev = torch.zeros(self.ntiles, dtype=i32, device=device)
        # compact tiles of ALL ranks, rank-major, so that rank 0 receives straight into place
        self.compact = torch.zeros(self.ntiles * TILE * TILE, dtype=i32, device=device)
        self.fb = torch.zeros(width * height, dtype=i32, device=device) if rank == 0 else None
        self.tile_ids_dev = torch.zeros(self.ntiles, dtype=i32, device=device)
        self.launches = 0

    def _offsets(self):
        sizes = [len(p) for p in self.parts]
        return np.concatenate([[0], np.cumsum(sizes)])

    def render(self):
        """renders this rank's tiles, gathers all tiles on rank 0, scatters them into rank 0's frame, and
        all-reduces the per-tile costs for the next partition.  Stream-ordered on torch's current stream."""
        torch, dist = self.torch, self.dist
        stream = torch.cuda.current_stream().cuda_stream if self.device != "cpu" else 0
        mine = self.parts[self.rank]
        off = self._offsets()
        px = TILE * TILE
        self.ctx.set_tiles(mine)
        my_slice = self.compact[off[self.rank] * px: off[self.rank + 1] * px]
        self.ctx.render_tiles(my_slice.data_ptr(), 0, self.W, self.H, stream=stream)
        self.ctx.tile_costs(self.costs_dev.data_ptr(), self.ntiles, stream=stream)
        self.launches = 3
        if self.world > 1:
            ops = []
            if self.rank == 0:
                for r in range(1, self.world):
                    if off[r + 1] > off[r]:
                        ops.append(dist.P2POp(dist.irecv, self.compact[off[r] * px: off[r + 1] * px], r))
            elif len(mine) > 0:
                ops.append(dist.P2POp(dist.isend, my_slice, 0))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
            dist.all_reduce(self.costs_dev)                     # each rank contributed its own tiles' counts
        if self.rank == 0:
            all_ids = np.concatenate(self.parts).astype(np.int32)
            self.tile_ids_dev.copy_(torch.from_numpy(all_ids), non_blocking=False)
            self.ctx.scatter_tiles(self.compact.data_ptr(), 0, self.tile_ids_dev.data_ptr(), self.ntiles,
                                   self.fb.data_ptr(), 0, self.W, self.H, stream=stream)
            self.launches += 1

    # ------------------------------------------------------------------------------------------------------
    # pipelined frames (the bench / a viewer loop): nothing below waits for the GPU
    # ------------------------------------------------------------------------------------------------------
    def _init_async(self):
        torch = self.torch
        pin = self.device != "cpu"
        n = self.in_flight = max(1, int(getattr(self, "in_flight", 2)))
        self.tiles_pinned = [torch.zeros(self.ntiles, dtype=torch.int32, pin_memory=pin) for _ in range(n)]
        self.costs_host = [torch.zeros(self.ntiles, dtype=torch.int32, pin_memory=pin) for _ in range(n)]
        self.cost_events = [torch.cuda.Event() for _ in range(n)] if pin else [None] * n
        self.ids_pinned = [torch.zeros(self.ntiles, dtype=torch.int32, pin_memory=pin) for _ in range(n)]
        self.ids_dev = [torch.zeros(self.ntiles, dtype=torch.int32, device=self.device) for _ in range(n)]
        # two frames in flight (icb_set_frame_slot): everything a frame writes exists twice
        self.compact2 = [self.compact] + [torch.zeros_like(self.compact) for _ in range(n - 1)]
        self.costs_dev2 = [self.costs_dev] + [torch.zeros_like(self.costs_dev) for _ in range(n - 1)]
        self.fb2 = ([self.fb] + [torch.zeros_like(self.fb) for _ in range(n - 1)]) if self.rank == 0 else [None] * n
        self.frame_no = 0
        self.t_wait = self.t_host = 0.0
        self.rebalance_every = max(1, int(getattr(self, "rebalance_every", 1)))

    def render_async(self):
        """One frame, enqueued on torch's current stream without a single host synchronisation: this rank's tile list goes up
        by a stream-ordered copy from pinned memory, the all-reduced per-tile costs come back into pinned memory behind an
        event, and the partition of frame f is cut from the costs of frame f-2 (the newest that have certainly arrived),
        so that the host work of a frame overlaps the previous frame's kernel.  The kernel's own heaviest-first order uses
        the all-reduced costs of frame f-1 (icb_set_tile_costs, stream-ordered)."""
        torch, dist = self.torch, self.dist
        if not hasattr(self, "frame_no"):
            self._init_async()
        st = torch.cuda.current_stream()
        stream = st.cuda_stream
        import time
        t_in = time.perf_counter()
        f, b = self.frame_no, self.frame_no % self.in_flight
        if f >= self.in_flight:
            self.cost_events[b].synchronize()          # recorded in_flight frames ago: that frame has left the device
            self.t_wait += time.perf_counter() - t_in  # (time the host spent waiting for the device: 0 = the host is the bottleneck)
            if f % self.rebalance_every == 0:
                self.costs = self.costs_host[b].numpy()
                self.parts = assign_tiles_dealt(self.order, self.costs, self.world)
        elif f == 0:
            self.parts = assign_tiles_dealt(self.order, None, self.world)
        mine = self.parts[self.rank]
        off = self._offsets()
        px = TILE * TILE
        compact, costs_dev = self.compact2[b], self.costs_dev2[b]
        self.ctx.set_frame_slot(b)
        self.tiles_pinned[b][:len(mine)] = torch.from_numpy(mine)
        self.ctx.set_tiles_async(self.tiles_pinned[b].data_ptr(), len(mine), stream)
        my_slice = compact[off[self.rank] * px: off[self.rank + 1] * px]
        self.ctx.render_tiles(my_slice.data_ptr(), 0, self.W, self.H, stream=stream)
        self.ctx.tile_costs(costs_dev.data_ptr(), self.ntiles, stream=stream)
        self.launches = 3
        if self.world > 1:
            ops = []
            if self.rank == 0:
                for r in range(1, self.world):
                    if off[r + 1] > off[r]:
                        ops.append(dist.P2POp(dist.irecv, compact[off[r] * px: off[r + 1] * px], r))
            elif len(mine) > 0:
                ops.append(dist.P2POp(dist.isend, my_slice, 0))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()                                   # (stream-level wait, not a host wait)
            dist.all_reduce(costs_dev)
            self.ctx.set_tile_costs(costs_dev.data_ptr(), self.ntiles, stream)
        self.costs_host[b].copy_(costs_dev, non_blocking=True)
        self.cost_events[b].record(st)
        if self.rank == 0:
            self.ids_pinned[b][:] = torch.from_numpy(np.concatenate(self.parts).astype(np.int32))   # tile id of every compact tile, rank-major
            self.ids_dev[b].copy_(self.ids_pinned[b], non_blocking=True)
            self.ctx.scatter_tiles(compact.data_ptr(), 0, self.ids_dev[b].data_ptr(), self.ntiles,
                                   self.fb2[b].data_ptr(), 0, self.W, self.H, stream=stream)
            self.launches += 1
            self.fb = self.fb2[b]                   # the frame this call produces (valid once `st` has run)
        self.frame_no += 1
        self.t_host += time.perf_counter() - t_in

    def rebalance(self):
        """cut next frame's tile sets from this frame's costs (host side; needs the stream to have finished)"""
        self.costs = self.costs_dev.cpu().numpy()
        self.parts = assign_tiles(self.order, self.costs, self.world)
        return self.costs


class ScalarStream:
    """Time series (BASELINE configs[3]): per-timestep scalars streamed from pinned host memory, double-buffered
    against rendering.  Each rank uploads 1/N of a timestep over its own PCIe link into a device staging buffer, the slabs
    are exchanged over NVLink (one NCCL broadcast per slab, on a communicator of its own so that it runs beside the tile
    gather), and icb_set_scalars_dev scatters the staged values into the grid's internal order (and recomputes the prism
    sampler's vertex table / marks the majorant grid stale).  torch is plumbing: buffers, streams, NCCL."""

    def __init__(self, ctx, n_records, rank, world, torch, dist, device="cuda"):
        self.ctx, self.n, self.rank, self.world, self.torch, self.dist = ctx, int(n_records), rank, world, torch, dist
        self.stage = [torch.empty(self.n * 32, dtype=torch.float32, device=device) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream()
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [None, None]
        self.cuts = [self.n * r // world * 32 for r in range(world + 1)]
        self.group = dist.new_group() if world > 1 else None
        self.bytes_per_rank = (self.cuts[rank + 1] - self.cuts[rank]) * 4

    def upload(self, values_pinned, slot):
        """values_pinned: float32 [n*32] in page-locked host memory (the caller's record order).  Asynchronous."""
        torch, dist = self.torch, self.dist
        with torch.cuda.stream(self.copy_stream):
            if self.free[slot] is not None:
                self.copy_stream.wait_event(self.free[slot])        # the scatter that last read this buffer has run
            a, b = self.cuts[self.rank], self.cuts[self.rank + 1]
            self.stage[slot][a:b].copy_(values_pinned[a:b], non_blocking=True)
            for r in range(self.world if self.world > 1 else 0):
                dist.broadcast(self.stage[slot][self.cuts[r]:self.cuts[r + 1]], src=r, group=self.group)
            self.ready[slot].record(self.copy_stream)

    def apply(self, slot, render_streams):
        """Swap the grid's scalars for the staged timestep on torch's current stream: after every frame still reading the
        old ones, before every frame enqueued afterwards on `render_streams`."""
        torch = self.torch
        st = torch.cuda.current_stream()
        for s in render_streams:
            st.wait_stream(s)
        st.wait_event(self.ready[slot])
        self.ctx.set_scalars_dev(self.stage[slot].data_ptr(), self.n, stream=st.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(st)
        self.free[slot] = ev
        for s in render_streams:
            s.wait_stream(st)
