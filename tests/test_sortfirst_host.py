"""CPU: host logic of the sort-first multi-GPU path (tile order, cost-balanced partition, and the world_size-2
exchange pattern over gloo with a stand-in renderer -- the real renderer needs a GPU and is covered by -m gpu)."""
import os
import socket
import sys

import numpy as np
import pytest

import sortfirst


def test_morton_order_is_a_permutation_with_locality():
    tx, ty = 240, 135                        # 3840x2160 in 16x16 tiles
    o = sortfirst.morton_order(tx, ty)
    assert sorted(o.tolist()) == list(range(tx * ty))
    x, y = o % tx, o // tx
    # consecutive tiles of the curve are close on screen (median step 1 tile)
    step = np.abs(np.diff(x)) + np.abs(np.diff(y))
    assert np.median(step) == 1 and np.mean(step) < 3


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_assign_tiles_covers_and_balances(world):
    tx, ty = 240, 135
    o = sortfirst.morton_order(tx, ty)
    parts = sortfirst.assign_tiles(o, None, world)
    assert len(parts) == world
    assert sorted(np.concatenate(parts).tolist()) == list(range(tx * ty))
    assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 2
    rng = np.random.default_rng(0)
    yy, xx = np.divmod(np.arange(tx * ty), tx)
    costs = (4000 * np.exp(-((xx - 120) ** 2 + (yy - 60) ** 2) / 2000.0)).astype(np.int64) + rng.integers(0, 10, tx * ty)
    parts = sortfirst.assign_tiles(o, costs, world)
    assert sorted(np.concatenate(parts).tolist()) == list(range(tx * ty))
    loads = np.array([costs[p].sum() for p in parts], np.float64)
    assert loads.max() <= loads.mean() * 1.01 + costs.max()
    # deterministic: same inputs, same partition (every rank computes it locally)
    again = sortfirst.assign_tiles(o, costs, world)
    assert all(np.array_equal(a, b) for a, b in zip(parts, again))
    # degenerate: all-zero costs and more ranks than tiles
    z = sortfirst.assign_tiles(sortfirst.morton_order(2, 1), np.zeros(2), 4)
    assert sorted(np.concatenate(z).tolist()) == [0, 1]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_dealt_tiles_cover_and_balance(world):
    order = sortfirst.morton_order(240, 135)
    rng = np.random.default_rng(world)
    costs = np.minimum(rng.pareto(1.5, len(order)) * 1000, 35000).astype(np.int64)   # heavy-tailed, like limb tiles (max / mean ~ 35)
    costs[rng.random(len(order)) < 0.5] = 0                                # half the screen is empty
    for c in (None, costs):
        parts = sortfirst.assign_tiles_dealt(order, c, world)
        allt = np.concatenate(parts)
        assert len(parts) == world and sorted(allt.tolist()) == list(range(len(order)))
        again = sortfirst.assign_tiles_dealt(order, c, world)
        assert all(np.array_equal(a, b) for a, b in zip(parts, again))
        if c is not None and world > 1:
            load = np.array([costs[p].sum() + len(p) for p in parts], np.float64)
            assert load.max() / load.mean() < 1.1
            heavy = np.argsort(-costs)[:64 * world]                        # the heaviest tiles are spread over all ranks
            share = np.array([np.isin(heavy, p).sum() for p in parts])
            assert share.min() > 0


class FakeCtx:
    """stand-in for icb.Context: 'renders' pixel value = tile id * 1000 + pixel index, cost = tile id + 1"""

    def __init__(self, torch, ntiles):
        self.torch, self.ntiles, self.tiles = torch, ntiles, None
        self.last_costs = None

    def set_tiles(self, t):
        self.tiles = np.asarray(t, np.int32)

    def render_tiles(self, ptr, accum_ptr, W, H, stream=0):
        self.pending = ptr

    def fill(self, tensor):
        t = self.torch.from_numpy(self.tiles.astype(np.int64))
        tensor.copy_((t[:, None] * 1000 + self.torch.arange(256)[None, :]).reshape(-1).to(self.torch.int32))
        c = np.zeros(self.ntiles, np.int32)
        c[self.tiles] = self.tiles + 1
        self.last_costs = self.torch.from_numpy(c)

    def tile_costs(self, ptr, n, stream=0):
        pass

    def scatter_tiles(self, *a, **k):
        pass


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H = 160, 96
    ctx = FakeCtx(torch, (W // 16) * (H // 16))
    sf = sortfirst.SortFirstFrame(ctx, W, H, rank, world, torch, dist, "cpu")
    history = []
    for frame in range(3):
        # what SortFirstFrame.render does, with the fake renderer filling this rank's slice
        off = sf._offsets()
        mine = sf.parts[rank]
        ctx.set_tiles(mine)
        my_slice = sf.compact[off[rank] * 256: off[rank + 1] * 256]
        ctx.fill(my_slice)
        sf.costs_dev.copy_(ctx.last_costs)
        ops = []
        if rank == 0:
            for r in range(1, world):
                if off[r + 1] > off[r]:
                    ops.append(dist.P2POp(dist.irecv, sf.compact[off[r] * 256: off[r + 1] * 256], r))
        elif len(mine):
            ops.append(dist.P2POp(dist.isend, my_slice, 0))
        for req in (dist.batch_isend_irecv(ops) if ops else []):
            req.wait()
        dist.all_reduce(sf.costs_dev)
        if rank == 0:
            ids = np.concatenate(sf.parts)
            got = sf.compact.numpy().reshape(-1, 256)
            want = ids[:, None].astype(np.int64) * 1000 + np.arange(256)[None, :]
            history.append(bool((got == want).all()))
        costs = sf.rebalance()
        assert (costs == np.arange(len(costs)) + 1).all()
        history.append([len(p) for p in sf.parts])
    q.put((rank, history))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_and_rebalance_over_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # rank 0 saw every tile of every rank in place, every frame
    assert res[0][0] is True and res[0][2] is True and res[0][4] is True
    # both ranks computed the same partition; after rebalancing by cost (= tile id + 1) it is no longer even
    assert res[0][1] == res[1][0] and res[0][3] == res[1][1] and res[0][5] == res[1][2]
    assert sum(res[0][1]) == 60 and res[0][1][0] != res[0][1][1]
