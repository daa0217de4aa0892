"""GPU (-m gpu): split rays (icb_set_ray_split; DESIGN.md "split rays").  The marcher composites every ray in chunks of 2^k
sample indices and merges the chunks in order, so the chunks of the rays of a frame's heaviest tiles can be marched by
different lanes (a chunk pass in front of the pixel pass) without changing the result: frames with splitting forced on for
every tile are bit-identical to frames with splitting off, count the same samples, and follow the chunked CPU statement
(oracle march_pixel with chunk > 0) -- which in turn stays within 1e-3 of the plain statement (CPU test)."""
import numpy as np
import pytest

import icb
import oracle_lib as ol

pytestmark = pytest.mark.gpu
TOL = 1e-3


@pytest.fixture()
def ctx():
    c = icb.Context(0)
    yield c
    c.close()


def _render(ctx, W, H, n=2):
    import torch
    fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    acc = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
    for _ in range(n):                      # (the second frame orders and splits by the first one's tile costs)
        ctx.render(fb.data_ptr(), acc.data_ptr(), W, H)
        torch.cuda.synchronize()
    st = ctx.stats()
    return fb.cpu().numpy().view(np.uint32), acc.cpu().numpy(), st


@pytest.mark.parametrize("chunk_log2", [5, 8, 10])
@pytest.mark.parametrize("sampler", ["cell", "prism"])
@pytest.mark.parametrize("tf", ["sharp", "smooth"])
def test_split_frames_are_bit_identical_to_unsplit_frames(ctx, chunk_log2, sampler, tf):
    cells = icb.gen_grid(2, 4, 90)
    ctx.set_grid(cells)
    ctx.set_sampler(sampler)
    b = ctx.bounds()
    W, H = 328, 200
    lut = icb.sharp_lut() if tf == "sharp" else icb.smooth_lut()
    ctx.set_transfunc(lut, opacity_scale=1.0 if tf == "sharp" else 0.02)      # (thin smooth medium: rays run the whole shell)
    ctx.set_params(jitter=True)                                               # unitDistance 1000, dt 500: limb rays ~12 000 samples
    for theta, fovy in ((30.0, 35.0), (200.0, 12.0)):
        ctx.set_camera(icb.camera_orbit(b["volLo"], b["volHi"], W, H, theta, fovy=np.float32(np.deg2rad(fovy)), pose_fovy=icb.FOVY_DEFAULT))
        ctx.set_ray_split(chunk_log2, 0, 0, 1)                                # no ray is split
        fb0, acc0, st0 = _render(ctx, W, H)
        assert st0["kernelLaunches"] == 3
        ctx.set_ray_split(chunk_log2, 384, 0, 1)                                # every tile's rays go through the chunk pass
        fb1, acc1, st1 = _render(ctx, W, H)
        assert st1["kernelLaunches"] == 4
        assert np.array_equal(acc0, acc1) and np.array_equal(fb0, fb1)
        assert st0["steps"] == st1["steps"] and st0["hits"] == st1["hits"] and st0["steps"] > 1e6
        ctx.set_ray_split(chunk_log2, 40, 12, 1)                                # a realistic cut: only tiles above 4096 lane iterations
        fb2, acc2, st2 = _render(ctx, W, H, 3)
        assert np.array_equal(acc0, acc2) and np.array_equal(fb0, fb2) and st0["steps"] == st2["steps"]


def test_split_frame_follows_the_chunked_cpu_statement(ctx):
    """R2B3 x 90, small image: the split GPU frame against the C port's chunked march_pixel (same samples, same hits,
    RGBA within 1e-3), and against the plain statement (the north-star bar)"""
    port = ol.Port()
    cells = icb.gen_grid(2, 3, 90)
    ctx.set_grid(cells)
    b = ctx.bounds()
    W, H = 64, 40
    cam = icb.camera_orient(np.array([0.2e7, 0.9e7, 0.75e7], np.float32), (0.3e7, 0, 0), (0, 0, 1), W, H)
    lut = icb.smooth_lut()
    ctx.set_transfunc(lut, value_range=(0.1, 0.9), opacity_scale=0.05)
    ctx.set_camera(cam)
    ctx.set_params(unit_distance=4000.0, dt=2000.0, jitter=True)
    ctx.set_ray_split(4, 384, 0, 1)
    fb, acc, st = _render(ctx, W, H)
    kw = dict(tf_range=(0.1, 0.9), opacity_scale=0.05, camera=cam, unit_distance=4000.0, dt=2000.0, r_inner=b["sphLo"][0],
              r_outer=b["sphHi"][0], jitter=1)
    want = port.march_frame(ol.Scene(cells, lut, chunk=16, **kw), W, H)
    plain = port.march_frame(ol.Scene(cells, lut, **kw), W, H)
    assert st["steps"] == want["steps"] and st["hits"] == want["hits"] and want["perPixel"][:, 0].max() > 4 * 16
    assert np.abs(acc - want["accum"]).max() <= 1e-5
    assert np.abs(acc - plain["accum"]).max() <= TOL


def test_split_rays_with_tile_lists_and_frames_in_flight(ctx):
    """the sort-first building blocks with splitting on: compact tiles of a tile subset, two slots in flight"""
    import torch
    cells = icb.gen_grid(2, 4, 60)
    ctx.set_grid(cells)
    b = ctx.bounds()
    W, H = 200, 120
    ctx.set_transfunc(icb.sharp_lut())
    ctx.set_params(jitter=True)
    cams = [icb.camera_orbit(b["volLo"], b["volHi"], W, H, 30.0 + 9 * i, fovy=np.float32(np.deg2rad(35.0)), pose_fovy=icb.FOVY_DEFAULT) for i in range(4)]
    ctx.set_ray_split(7, 0, 0, 1)
    want = []
    for cam in cams:
        ctx.set_camera(cam)
        want.append(_render(ctx, W, H, 1)[0])
    ctx.set_ray_split(7, 384, 0, 1)
    tx, ty = (W + 15) // 16, (H + 15) // 16
    rng = np.random.default_rng(4)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    got = [torch.zeros(W * H, dtype=torch.int32, device="cuda") for _ in cams]
    parts = [np.array_split(rng.permutation(tx * ty).astype(np.int32), 2) for _ in cams]
    bufs = []
    for i, cam in enumerate(cams):
        s = i & 1
        ctx.set_frame_slot(s)
        ctx.set_camera(cam)
        for part in parts[i]:
            torch.cuda.synchronize()                 # (icb_set_tiles is the synchronous form)
            ctx.set_tiles(part)
            t_fb = torch.zeros(len(part) * 256, dtype=torch.int32, device="cuda")
            ids = torch.from_numpy(part).cuda()
            bufs.append((t_fb, ids))
            ctx.render_tiles(t_fb.data_ptr(), 0, W, H, stream=streams[s].cuda_stream)
            ctx.scatter_tiles(t_fb.data_ptr(), 0, ids.data_ptr(), len(part), got[i].data_ptr(), 0, W, H, stream=streams[s].cuda_stream)
    torch.cuda.synchronize()
    for s in (0, 1):
        ctx.set_frame_slot(s)
        ctx.set_tiles(np.zeros(0, np.int32))
    ctx.set_frame_slot(0)
    for i in range(len(cams)):
        assert np.array_equal(got[i].cpu().numpy().view(np.uint32), want[i]), i
