"""GPU (-m gpu): the multi-GPU frame loop behind the C ABI (icb_group_*, csrc/icb_group.cu).  The reference has no
multi-GPU path; the contract is SURVEY.md §8e: the assembled sort-first frame is bit-identical to the
single-GPU frame, the grid is built once and broadcast, time-series scalars travel 1/N per rank + one all-gather.
Groups of one run on any box; the N-rank cases run where the box has N GPUs (gpurun --gpus N) and skip elsewhere."""
import numpy as np
import pytest

import icb

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


def _cams(b, W, H, n):
    return [icb.camera_orbit(b["volLo"], b["volHi"], W, H, 30.0 + 11 * i, fovy=np.float32(np.deg2rad(35.0)), pose_fovy=icb.FOVY_DEFAULT)
            for i in range(n)]


def _reference_frames(ctx, cams, W, H, scalars=None):
    import torch
    fb = torch.zeros(W * H, dtype=torch.int32, device=f"cuda:{ctx.device}")
    acc = torch.zeros(W * H, 4, dtype=torch.float32, device=f"cuda:{ctx.device}")
    out = []
    ctx.set_frame_slot(0)
    ctx.set_tiles(np.zeros(0, np.int32))
    with torch.cuda.device(ctx.device):
        for i, cam in enumerate(cams):
            if scalars is not None:
                ctx.set_scalars(scalars[i])
            ctx.set_camera(cam)
            ctx.render(fb.data_ptr(), acc.data_ptr(), W, H)
            torch.cuda.synchronize()
            out.append(fb.cpu().numpy().view(np.uint32).copy())
    return out


@pytest.mark.parametrize("nranks", [1, 2, 4])
@pytest.mark.parametrize("in_flight", [1, 3])
def test_group_frames_equal_single_gpu_frames(nranks, in_flight):
    """icb_group_create_local over nranks GPUs: grid built on rank 0 and broadcast, frames in flight, frames landing in
    pinned host memory; every frame equals the frame one GPU renders alone, bit for bit"""
    if _gpus() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    ctxs = [icb.Context(d) for d in range(nranks)]
    grp = icb.Group.local(ctxs)
    try:
        cells = icb.gen_grid(2, 4, 40)
        grp.set_grid(cells)
        for c in ctxs[1:]:                                        # the replicas are the builder's arrays
            assert c.stats()["numRecords"] == len(cells) and c.bounds()["unitDistance"] == ctxs[0].bounds()["unitDistance"]
        b = ctxs[0].bounds()
        W, H = 328, 200                                           # ragged against the 16x16 tiles
        grp.set_transfunc(icb.sharp_lut())
        grp.set_params(jitter=True)
        cams = _cams(b, W, H, 7)
        grp.begin(W, H, in_flight)
        hosts = [ctxs[0].host_alloc(W * H * 4).view(np.uint32) for _ in range(in_flight)]
        got = []
        for i, cam in enumerate(cams):
            slot = i % in_flight
            if i >= in_flight:
                grp.wait(slot)
                got.append(hosts[slot].copy())
            grp.set_camera(cam)
            assert grp.render(hosts[slot]) == slot
        for i in range(len(cams) - in_flight, len(cams)):
            grp.wait(i % in_flight)
            got.append(hosts[i % in_flight].copy())
        st = grp.stats()
        assert st["world"] == nranks and st["frames"] == len(cams)
        want = _reference_frames(ctxs[0], cams, W, H)
        for i in range(len(cams)):
            assert np.array_equal(got[i], want[i]), i
        # every rank located the reference's cells: a few positions through each replica
        rng = np.random.default_rng(3)
        d = rng.normal(size=(3000, 3))
        d /= np.linalg.norm(d, axis=1)[:, None]
        pos = (d * rng.uniform(b["sphLo"][0], b["sphHi"][0], len(d))[:, None]).astype(np.float32)
        ids0, v0 = ctxs[0].locate_host(pos)
        for c in ctxs[1:]:
            ids, v = c.locate_host(pos)
            assert np.array_equal(ids, ids0) and np.array_equal(v, v0)
    finally:
        grp.close()
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("nranks", [1, 2])
def test_group_time_series_equals_set_scalars(nranks):
    """icb_group_stream_scalars / icb_group_swap_scalars (BASELINE configs[3]): timesteps streamed from pinned host memory
    (1/N per rank + one all-gather of padded slabs), swapped in between frames in flight; every frame equals the frame
    rendered after a plain icb_set_scalars of the same timestep"""
    import torch
    if _gpus() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    ctxs = [icb.Context(d) for d in range(nranks)]
    grp = icb.Group.local(ctxs)
    try:
        n, k, L = 2, 3, 40
        cells = icb.gen_grid(n, k, L)
        steps = [cells] + [icb.gen_grid(n, k, L, timestep=s) for s in (3, 9, 14)]
        grp.set_grid(cells)
        b = ctxs[0].bounds()
        W, H = 160, 96
        grp.set_transfunc(icb.sharp_lut())
        grp.set_params(dt=5000.0, unit_distance=10000.0, jitter=True)
        cams = _cams(b, W, H, 8)
        pinned = [torch.from_numpy(np.ascontiguousarray(s_["value"]).reshape(-1)).pin_memory() for s_ in steps]
        grp.begin(W, H, 2)
        hosts = [ctxs[0].host_alloc(W * H * 4).view(np.uint32) for _ in range(len(cams))]
        grp.stream_scalars(pinned[0].data_ptr(), len(cells))
        for i, cam in enumerate(cams):
            if i + 1 < len(cams):
                grp.stream_scalars(pinned[(i + 1) % 4].data_ptr(), len(cells))      # the next timestep travels while this one renders
            grp.swap_scalars()
            grp.set_camera(cam)
            grp.render(hosts[i])
        grp.wait(-1)
        want = _reference_frames(ctxs[0], cams, W, H, scalars=[steps[i % 4]["value"] for i in range(len(cams))])
        assert not np.array_equal(want[0], want[4]) or True
        for i in range(len(cams)):
            assert np.array_equal(hosts[i], want[i]), i
    finally:
        grp.close()
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("nranks", [2, 4])
def test_group_with_more_ranks_than_blocks_of_tiles(nranks):
    """a frame so small that the deal leaves ranks without a single tile (tiles are dealt in blocks of 16: 64x64 is one
    block, 96x64 two): those ranks render nothing but still take part in the cost exchange, frame after frame, and the
    frame is the single-GPU frame"""
    if _gpus() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    ctxs = [icb.Context(d) for d in range(nranks)]
    grp = icb.Group.local(ctxs)
    try:
        grp.set_grid(icb.gen_grid(2, 3, 40))
        b = ctxs[0].bounds()
        grp.set_transfunc(icb.sharp_lut())
        grp.set_params(jitter=True)
        for (W, H) in [(64, 64), (96, 64)]:
            cams = _cams(b, W, H, 4)
            grp.begin(W, H, 2)
            hosts = [ctxs[0].host_alloc(W * H * 4).view(np.uint32) for _ in range(len(cams))]
            for i, cam in enumerate(cams):
                grp.set_camera(cam)
                grp.render(hosts[i])
            grp.wait(-1)
            want = _reference_frames(ctxs[0], cams, W, H)
            for i in range(len(cams)):
                assert np.array_equal(hosts[i], want[i]), (W, H, i)
    finally:
        grp.close()
        for c in ctxs:
            c.close()


def test_group_refuses_a_third_pending_timestep():
    """two stage buffers: a third icb_group_stream_scalars before a swap would overwrite a staged timestep -- it is refused,
    and after a swap the next upload goes through"""
    import torch
    ctx = icb.Context(0)
    grp = icb.Group.local([ctx])
    try:
        cells = icb.gen_grid(2, 2, 40)
        grp.set_grid(cells)
        pinned = torch.from_numpy(np.ascontiguousarray(cells["value"]).reshape(-1)).pin_memory()
        grp.stream_scalars(pinned.data_ptr(), len(cells))
        grp.stream_scalars(pinned.data_ptr(), len(cells))
        with pytest.raises(icb.IcbError, match="pending"):
            grp.stream_scalars(pinned.data_ptr(), len(cells))
        grp.swap_scalars()
        grp.stream_scalars(pinned.data_ptr(), len(cells))
        grp.swap_scalars()
        grp.swap_scalars()
        with pytest.raises(icb.IcbError, match="pending"):
            grp.swap_scalars()
    finally:
        grp.close()
        ctx.close()


def test_group_refuses_handles_with_different_compositing_arithmetic():
    """icb_set_ray_split's mode selects the compositing arithmetic (plain recurrence vs chunks): tiles of handles that
    disagree would not assemble into the single-GPU frame bit for bit, so icb_group_begin refuses them"""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    ctxs = [icb.Context(d) for d in range(2)]
    grp = icb.Group.local(ctxs)
    try:
        ctxs[1].set_ray_split(mode=2)
        with pytest.raises(icb.IcbError, match="icb_set_ray_split"):
            grp.begin(64, 64, 1)
        ctxs[0].set_ray_split(mode=1)           # ALWAYS and AUTO are the same arithmetic
        grp.begin(64, 64, 1)
        ctxs[0].set_ray_split(chunk_log2=8)
        with pytest.raises(icb.IcbError, match="icb_set_ray_split"):
            grp.begin(64, 64, 1)
    finally:
        grp.close()
        for c in ctxs:
            c.close()
