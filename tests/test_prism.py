"""Prism sampler (icb_set_sampler(ICB_SAMPLER_PRISM)): barycentric-in-triangle x linear-in-height
interpolation inside the located prism.

Oracle chain:
  reference  intersectWedgeEXT (UElems.h:214-311: Newton inverse map + shape functions), compiled in place (oracle/_ref),
             golden vectors in tests/golden/prism_*.npz (tools/gen_golden_prism.py)
  port       oracle/icon_oracle.c orc_prism_*: the same wedge function in closed form + the vertex-scalar definition
  device     csrc/locator.cuh prismValue / vertexScalarAt: compiled for the host here (tests/emu), on B200 under -m gpu

Bars: record ids and layers bit-exact (location is the constant sampler's); device values bit-identical to the port;
both within PRISM_TOL of the reference's Newton result wherever the reference accepts the located wedge -- the
reference iterates until its pcoords step is below 1e-4 (UElems.h:227), so its own result is only defined to about
1e-4 of the wedge's scalar span; the tolerance is 5e-4 of the span + 1e-6.  Frames: RGBA fp32 within 1e-3 (north_star).
"""
import os

import numpy as np
import pytest

import icb
import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GRIDS = ["r2b1_l40", "r2b5_l20_head1200"]
TOL = 1e-3


def prism_tol(wedges):
    span = wedges[:, :, 3].max(1) - wedges[:, :, 3].min(1)
    return 5e-4 * span + 1e-6


@pytest.fixture(scope="module")
def port():
    return ol.Port()


@pytest.fixture(scope="module")
def emu():
    import emu_lib
    return emu_lib.Emu()


def load(name):
    return np.fromfile(os.path.join(GOLD, f"{name}.ic"), icb.CELL_DTYPE), np.load(os.path.join(GOLD, f"prism_{name}.npz"))


def same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return np.array_equal(a.view(np.uint32)[~np.isnan(a)], b.view(np.uint32)[~np.isnan(a)]) and np.array_equal(np.isnan(a), np.isnan(b))


# ------------------------------------------------------------------------------------------------------------
# CPU: oracle vs the reference's golden vectors, device text on the host vs the oracle
# ------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", GRIDS)
def test_port_prism_vs_reference_golden(port, name):
    cells, g = load(name)
    prism = port.prism_build(cells)
    ids, vals, layers, pc = port.locate_prism(cells, prism, g["pos"])
    assert np.array_equal(ids, g["ids"]) and np.array_equal(layers, g["layers"])
    # the wedges handed to the reference are the ones this build produces (pins the vertex-scalar definition)
    assert np.array_equal(port.prism_wedges(cells, prism, ids, layers), g["wedges"])
    ok = g["ref_accept"] == 1
    assert ok.sum() > (150 if name == "r2b1_l40" else 2500)
    assert (np.abs(vals - g["ref_value"])[ok] <= prism_tol(g["wedges"])[ok]).all()
    assert np.allclose(vals, g["port_value"], rtol=0, atol=1e-6, equal_nan=True)
    # parametric coordinates are inside the wedge (t is clamped; r, s up to fp32 noise)
    hit = ids >= 0
    assert (pc[hit, 2] >= 0).all() and (pc[hit, 2] <= 1).all()
    assert (pc[hit, :2] > -1e-3).all() and (pc[hit, 0] + pc[hit, 1] < 1 + 1e-3).all()


@pytest.mark.skipif(not ol.Ref.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_port_prism_vs_reference_live(port):
    ref = ol.Ref()
    cells = icb.gen_grid(2, 4, 20)
    prism = port.prism_build(cells)
    rng = np.random.default_rng(5)
    d = rng.normal(size=(4000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    b = port.grid_bounds(cells)
    pos = (d * rng.uniform(b["sphLo"][0], b["sphHi"][0], len(d))[:, None]).astype(np.float32)
    ids, vals, layers, _ = port.locate_prism(cells, prism, pos)
    V = port.prism_wedges(cells, prism, ids, layers)
    acc, rv = ref.wedge(pos, V)
    ok = (acc == 1) & (ids >= 0)
    assert ok.sum() > 1000
    assert (np.abs(vals - rv)[ok] <= prism_tol(V)[ok]).all()


def test_vertex_scalars_are_continuous_and_reproduce_linear_fields(port):
    """properties of the definition: (1) columns sharing a vertex see the same scalar there; (2) a field that is the
    same in every column and linear in the layer index is reproduced exactly at the layer mid-heights' interfaces"""
    cells = icb.gen_grid(2, 2, 20).copy()
    for l in range(20):
        cells["value"][:, l] = 0.25 + 0.5 * l / 20
    vtx, geom = port.prism_build(cells)
    I = np.concatenate([[0.25], 0.25 + 0.5 * (np.arange(1, 20) - 0.5) / 20, [0.25 + 0.5 * 19 / 20]]).astype(np.float32)
    assert np.abs(vtx[:, :21, :] - I[None, :, None]).max() < 1e-6
    cells = icb.gen_grid(2, 2, 20)
    vtx, geom = port.prism_build(cells)
    key = {}
    for r in range(len(cells)):
        for k in range(3):
            kk = (cells["lat"][r, k].tobytes(), cells["lon"][r, k].tobytes())
            key.setdefault(kk, []).append(vtx[r, :21, k])
    shared = [v for v in key.values() if len(v) > 1]
    assert len(shared) > 100
    assert all(np.array_equal(v[0], w) for v in shared for w in v[1:])


@pytest.mark.parametrize("n,k,levels", [(2, 2, 20), (2, 3, 90), (3, 1, 35)])
def test_emulated_device_prism_equals_port(emu, port, n, k, levels):
    cells = icb.gen_grid(n, k, levels)
    prism = port.prism_build(cells)
    emu.build(cells)
    emu.set_sampler("prism")
    rng = np.random.default_rng(n * 100 + k)
    b = port.grid_bounds(cells)
    d = rng.normal(size=(6000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos = (d * rng.uniform(b["sphLo"][0] * 0.9999, b["sphHi"][0] * 1.0001, len(d))[:, None]).astype(np.float32)
    ids, vals, _, _ = port.locate_prism(cells, prism, pos)
    for flags, chunk in ((0, 0), (1, 64)):
        eid, ev, _ = emu.locate(pos, flags, chunk)
        assert np.array_equal(eid, ids)
        assert same_bits(vals[ids >= 0], ev[ids >= 0])
    emu.set_sampler("cell")


def test_emulated_prism_on_ragged_grid(emu, port):
    """columns of different depth, a missing neighbour, stacked records with a gap: ids and values still the port's"""
    cells = icb.gen_grid(2, 1, 40).copy()
    cells = np.delete(cells, [5, 77])                       # a column loses its upper / lower record
    cells["numLayers"][10] = 7                              # a shallow record
    cells["height"][20, :] += 30000.0                       # a record lifted off its neighbour below (gap)
    prism = port.prism_build(cells)
    emu.build(cells)
    emu.set_sampler("prism")
    rng = np.random.default_rng(9)
    b = port.grid_bounds(cells)
    d = rng.normal(size=(6000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos = (d * rng.uniform(b["sphLo"][0], b["sphHi"][0], len(d))[:, None]).astype(np.float32)
    ids, vals, _, _ = port.locate_prism(cells, prism, pos)
    eid, ev, _ = emu.locate(pos)
    assert np.array_equal(eid, ids)
    assert same_bits(vals[ids >= 0], ev[ids >= 0])
    emu.set_sampler("cell")


def test_emulated_prism_march_equals_port(emu, port):
    cells = icb.gen_grid(2, 2, 20)
    prism = port.prism_build(cells)
    emu.build(cells)
    emu.set_sampler("prism")
    b = port.grid_bounds(cells)
    W, H = 24, 16
    cam = icb.camera_orient(np.array([0.3e7, 1.1e7, 0.6e7], np.float32), (0, 0, 0), (0, 0, 1), W, H)
    for lut in (icb.smooth_lut(), icb.sharp_lut()):
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=20000.0, dt=9000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0],
                      prism=prism)
        for (x, y) in [(3, 4), (12, 8), (20, 3), (11, 11)]:
            want_ids, wn, wrgba = port.march_trace(sc, W, H, x, y)
            for skip in (False, True):
                ids, n, rgba = emu.march_pixel(sc, W, H, x, y, skip=skip)
                assert n == wn and np.array_equal(ids, want_ids)
                assert np.abs(rgba - wrgba).max() <= 1e-5
    emu.set_sampler("cell")


# ------------------------------------------------------------------------------------------------------------
# GPU: the CUDA path through the C ABI
# ------------------------------------------------------------------------------------------------------------
@pytest.fixture()
def ctx():
    c = icb.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", GRIDS)
def test_gpu_prism_locate_golden(ctx, port, name):
    cells, g = load(name)
    ctx.set_grid(cells)
    ctx.set_sampler("prism")
    prism = port.prism_build(cells)
    _, want, _, _ = port.locate_prism(cells, prism, g["pos"])
    ok = g["ref_accept"] == 1
    for flags in (0, icb.LOCATE_WALK, icb.LOCATE_EXACT):
        ids, vals = ctx.locate_host(g["pos"], flags)
        assert np.array_equal(ids, g["ids"])
        assert same_bits(want[ids >= 0], vals[ids >= 0])                                   # the port, bit for bit
        assert (np.abs(vals - g["ref_value"])[ok] <= prism_tol(g["wedges"])[ok]).all()     # the reference's Newton result
    # back to the reference's shipped sampler: value[layer]
    ctx.set_sampler("cell")
    ids, vals = ctx.locate_host(g["pos"])
    cid, cvals = port.locate(cells, g["pos"])
    assert np.array_equal(ids, cid) and np.array_equal(vals, cvals)


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,levels", [(2, 3, 90), (2, 4, 20), (3, 2, 35)])
def test_gpu_prism_locate_vs_port(ctx, port, n, k, levels):
    cells = icb.gen_grid(n, k, levels)
    ctx.set_sampler("prism")               # before the grid: the table is built by icb_set_grid
    ctx.set_grid(cells)
    prism = port.prism_build(cells)
    rng = np.random.default_rng(k)
    b = port.grid_bounds(cells)
    d = rng.normal(size=(3000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos = (d * rng.uniform(b["sphLo"][0] * 0.9999, b["sphHi"][0] * 1.0001, len(d))[:, None]).astype(np.float32)
    want_ids, want, _, _ = port.locate_prism(cells, prism, pos)
    for flags in (0, icb.LOCATE_WALK):
        ids, vals = ctx.locate_host(pos, flags)
        assert np.array_equal(ids, want_ids)
        assert same_bits(want[ids >= 0], vals[ids >= 0])


@pytest.mark.gpu
@pytest.mark.parametrize("tf", ["smooth", "sharp"])
def test_gpu_prism_frame_vs_port(ctx, port, tf):
    import torch
    cells = icb.gen_grid(2, 3, 90)
    ctx.set_grid(cells)
    ctx.set_sampler("prism")
    prism = port.prism_build(cells)
    b = ctx.bounds()
    W, H = 64, 40
    cam = icb.camera_orient(np.array([0.2e7, 0.9e7, 0.75e7], np.float32), (0.3e7, 0, 0), (0, 0, 1), W, H)
    lut = icb.smooth_lut() if tf == "smooth" else icb.sharp_lut()
    ctx.set_transfunc(lut, value_range=(0.1, 0.9), opacity_scale=0.7)
    ctx.set_camera(cam)
    ctx.set_params(unit_distance=20000.0, dt=9000.0, ambient_color=(0.9, 0.8, 1.0), ambient_radiance=1.1)
    sc = ol.Scene(cells, lut, tf_range=(0.1, 0.9), opacity_scale=0.7, camera=cam, unit_distance=20000.0, dt=9000.0,
                  r_inner=b["sphLo"][0], r_outer=b["sphHi"][0], ambient_color=(0.9, 0.8, 1.0), ambient_radiance=1.1, prism=prism)
    want = port.march_frame(sc, W, H)
    frames = {}
    for skip in (True, False):
        ctx.set_skip_ahead(skip)
        pp = torch.zeros(W * H, 4, dtype=torch.int32, device="cuda")
        fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
        ctx.render_trace(fb.data_ptr(), accum.data_ptr(), pp.data_ptr(), W, H)
        torch.cuda.synchronize()
        assert np.array_equal(pp.cpu().numpy(), want["perPixel"])          # sample counts and id-sequence hashes
        assert np.abs(accum.cpu().numpy() - want["accum"]).max() <= TOL
        frames[skip] = accum.cpu().numpy()
        # the production kernel (no trace) writes the same frame
        accum2 = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
        ctx.render(fb.data_ptr(), accum2.data_ptr(), W, H)
        torch.cuda.synchronize()
        assert np.array_equal(accum2.cpu().numpy(), frames[skip])
    assert np.array_equal(frames[True], frames[False])                     # skip-ahead changes nothing
    assert want["hits"] > 1000
    # and the prism sampler is not the cell sampler
    ctx.set_sampler("cell")
    accum3 = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
    ctx.render(fb.data_ptr(), accum3.data_ptr(), W, H)
    torch.cuda.synchronize()
    assert np.abs(accum3.cpu().numpy() - frames[True]).max() > 1e-3


@pytest.mark.gpu
def test_gpu_prism_follows_scalar_swaps(ctx, port):
    """time series: icb_set_scalars recomputes the vertex scalars on the device"""
    cells = icb.gen_grid(2, 3, 20)
    ctx.set_grid(cells)
    ctx.set_sampler("prism")
    nxt = icb.gen_grid(2, 3, 20, timestep=7)
    assert not np.array_equal(nxt["value"], cells["value"])
    ctx.set_scalars(nxt["value"])
    prism = port.prism_build(nxt)
    rng = np.random.default_rng(3)
    b = port.grid_bounds(cells)
    d = rng.normal(size=(3000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos = (d * rng.uniform(b["sphLo"][0], b["sphHi"][0], len(d))[:, None]).astype(np.float32)
    want_ids, want, _, _ = port.locate_prism(nxt, prism, pos)
    ids, vals = ctx.locate_host(pos)
    assert np.array_equal(ids, want_ids)
    assert same_bits(want[ids >= 0], vals[ids >= 0])


@pytest.mark.gpu
def test_gpu_prism_full_size_properties(ctx, port):
    """C2 (R2B6 x 90, 1920x1080): skip-ahead and the transparent-prism shortcut change nothing, walk == exact query,
    and the frame agrees with the port on a seeded set of pixels"""
    import torch
    cells = icb.gen_grid(2, 6, 90)
    ctx.set_grid(cells)
    ctx.set_sampler("prism")
    b = ctx.bounds()
    W, H = 1920, 1080
    cam = icb.camera_orbit(b["volLo"], b["volHi"], W, H, 30.0, fovy=np.float32(np.deg2rad(35.0)), pose_fovy=icb.FOVY_DEFAULT)
    ctx.set_camera(cam)
    ctx.set_params()
    res = {}
    for tf in ("sharp", "smooth"):
        ctx.set_transfunc(icb.sharp_lut() if tf == "sharp" else icb.smooth_lut())
        for skip in (True, False):
            ctx.set_skip_ahead(skip)
            fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
            accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
            ctx.render(fb.data_ptr(), accum.data_ptr(), W, H)
            torch.cuda.synchronize()
            res[(tf, skip)] = (accum.cpu().numpy(), ctx.stats()["steps"])
        assert np.array_equal(res[(tf, True)][0], res[(tf, False)][0])
        assert res[(tf, True)][1] == res[(tf, False)][1] > 10_000_000
