"""CPU: host logic of the sort-first multi-GPU path.  The partition the library computes in C++ (icb_partition_tiles =
what icb_group_render deals to the ranks, csrc/icb_group.cu) against its numpy statement (tests/partition_ref.py), its
balance properties, and the world_size-2 pattern over gloo: both ranks cut the same partition from the same all-reduced
costs and rank 0 receives every rank's compact tile slice in place (the renderer itself needs a GPU: -m gpu)."""
import os
import socket

import numpy as np
import pytest

import icb
import partition_ref as pref


def test_morton_order_is_a_permutation_with_locality():
    tx, ty = 240, 135                        # 3840x2160 in 16x16 tiles
    o = pref.morton_order(tx, ty)
    assert sorted(o.tolist()) == list(range(tx * ty))
    x, y = o % tx, o // tx
    # consecutive tiles of the curve are close on screen (median step 1 tile)
    step = np.abs(np.diff(x)) + np.abs(np.diff(y))
    assert np.median(step) == 1 and np.mean(step) < 3
    # the library's single-rank "partition" is the curve itself
    assert np.array_equal(icb.partition_tiles(tx, ty, None, 1)[0], o)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("shape", [(240, 135), (13, 7), (2, 1), (1, 1), (120, 68)])
def test_library_partition_equals_the_numpy_statement(world, shape):
    tx, ty = shape
    order = pref.morton_order(tx, ty)
    rng = np.random.default_rng(world * 100 + tx)
    costs = np.minimum(rng.pareto(1.5, len(order)) * 1000, 35000).astype(np.int32)   # heavy-tailed, like limb tiles
    costs[rng.random(len(order)) < 0.5] = 0                                          # half the screen is empty
    ties = np.full(len(order), 7, np.int32)                                          # all blocks tie: position decides
    for c in (None, costs, ties, np.zeros(len(order), np.int32)):
        want = pref.assign_tiles_dealt(order, c, world)
        got = icb.partition_tiles(tx, ty, c, world)
        assert len(got) == world
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        assert sorted(np.concatenate(got).tolist()) == list(range(len(order)))       # every tile exactly once


@pytest.mark.parametrize("world", [2, 4, 8])
def test_dealt_tiles_balance_and_spread_the_heavy_ones(world):
    tx, ty = 240, 135
    rng = np.random.default_rng(world)
    costs = np.minimum(rng.pareto(1.5, tx * ty) * 1000, 35000).astype(np.int32)      # max / mean ~ 35
    costs[rng.random(tx * ty) < 0.5] = 0
    parts = icb.partition_tiles(tx, ty, costs, world)
    again = icb.partition_tiles(tx, ty, costs, world)
    assert all(np.array_equal(a, b) for a, b in zip(parts, again))                   # deterministic
    load = np.array([costs[p].astype(np.int64).sum() + len(p) for p in parts], np.float64)
    assert load.max() / load.mean() < 1.1
    heavy = np.argsort(-costs)[:64 * world]                                          # the heaviest tiles are spread over all ranks
    assert min(np.isin(heavy, p).sum() for p in parts) > 0
    # each rank's tiles keep the Z order of the curve (coherent rays within a rank)
    pos = np.empty(tx * ty, np.int64)
    pos[pref.morton_order(tx, ty)] = np.arange(tx * ty)
    assert all((np.diff(pos[p]) > 0).all() for p in parts)


def test_group_api_is_exported_and_needs_no_nccl_until_used():
    L = icb.lib()
    for name in ("icb_group_unique_id", "icb_group_create", "icb_group_create_local", "icb_group_destroy", "icb_group_set_grid",
                 "icb_group_begin", "icb_group_render", "icb_group_wait", "icb_group_frame", "icb_group_stats",
                 "icb_group_stream_scalars", "icb_group_swap_scalars", "icb_partition_tiles"):
        assert hasattr(L, name), name
    assert L.icb_partition_tiles(0, 1, None, 1, None, None) == -1                    # ICB_ERR_INVALID
    # libicon_b200.so itself does not link NCCL (it is dlopen'ed by icb_group_* only)
    import subprocess
    out = subprocess.run(["ldd", icb.LIB_PATH], capture_output=True, text=True).stdout
    assert "nccl" not in out


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H = 160, 96
    tx, ty = W // 16, H // 16
    nt = tx * ty
    costs = None
    history = []
    for frame in range(3):
        parts = icb.partition_tiles(tx, ty, costs, world)
        off = np.concatenate([[0], np.cumsum([len(p) for p in parts])])
        mine = parts[rank]
        compact = torch.zeros(nt * 256, dtype=torch.int32)
        # a stand-in renderer: pixel = tile id * 1000 + pixel index, cost = tile id + 1 on the tiles this rank rendered
        my_slice = compact[off[rank] * 256: off[rank + 1] * 256]
        t = torch.from_numpy(mine.astype(np.int64))
        my_slice.copy_((t[:, None] * 1000 + torch.arange(256)[None, :]).reshape(-1).to(torch.int32))
        c = np.zeros(nt, np.int32)
        c[mine] = mine + 1
        costs_t = torch.from_numpy(c)
        ops = []
        if rank == 0:
            for r in range(1, world):
                if off[r + 1] > off[r]:
                    ops.append(dist.P2POp(dist.irecv, compact[off[r] * 256: off[r + 1] * 256], r))
        elif len(mine):
            ops.append(dist.P2POp(dist.isend, my_slice, 0))
        for req in (dist.batch_isend_irecv(ops) if ops else []):
            req.wait()
        dist.all_reduce(costs_t)
        if rank == 0:
            ids = np.concatenate(parts)
            got = compact.numpy().reshape(-1, 256)
            want = ids[:, None].astype(np.int64) * 1000 + np.arange(256)[None, :]
            history.append(bool((got == want).all()))
        costs = costs_t.numpy().copy()
        assert (costs == np.arange(nt) + 1).all()
        history.append([len(p) for p in parts])
    q.put((rank, history))
    dist.barrier()
    dist.destroy_process_group()


def _run_two_ranks():
    import queue
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q), daemon=True) for r in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=90) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        return res
    except queue.Empty:
        return None          # (the rendezvous did not come up, e.g. the probed port was taken in between)
    finally:
        for p in procs:      # never leave a worker behind: a lingering child keeps the test run's pipes open
            if p.is_alive():
                p.terminate()
                p.join(timeout=10)


def test_two_rank_gather_and_rebalance_over_gloo():
    res = _run_two_ranks() or _run_two_ranks()
    assert res is not None, "gloo rendezvous failed twice"
    # rank 0 saw every tile of every rank in place, every frame
    assert res[0][0] is True and res[0][2] is True and res[0][4] is True
    # both ranks computed the same partition, every frame; all 60 tiles are dealt
    assert res[0][1] == res[1][0] and res[0][3] == res[1][1] and res[0][5] == res[1][2]
    assert sum(res[0][1]) == 60
