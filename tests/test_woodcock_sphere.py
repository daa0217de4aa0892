"""The reference's ray-gen woodcockTrackingWithAccel (deviceCode.cu:281-341) in SPHERE_ACCEL_MODE -- the accel mode the
reference STARTS in (hostCode.cu:75 AppState::accelMode, :915) -- as icb_set_integrator(ICB_INTEGRATOR_WOODCOCK_SPHERE):
sdda (ShellAccel.h:83-199) over the (rad, lat, lon) shell accel (hostCode.cu:652-666 buildShellAccel, :298-334
buildShell_ICON, :362-397 computeMaxOpacities(ShellAccel)).

Oracle chain:
  reference  the unmodified ray-gen + sdda compiled in place (oracle/_ref) and the builder kernels cut out of hostCode.cu by
             line range and run on host shims; frames / ranges / max opacities in tests/golden/woodcock_sphere_r2b1_l40.npz
             (tools/gen_golden_woodcock_sphere.py)
  port       oracle/icon_oracle.c: orc_shell_ranges, orc_woodcock_sphere_frame -- pinned to the above bit for bit
  device     csrc/locator.cuh woodcockSpherePixel with csrc/fdlibm_trig.cuh for toSpherical's asinf/atan2f (on the host
             here, tests/emu; tests/test_fdlibm_trig.py sweeps the trig against libm), csrc/kernels.cu shellRangesKernel +
             woodcockSphereKernel on B200

Bar: bit-exact everywhere (RGBA8 frame, fp32 accumBuffer, per-cell value ranges and max opacities).
"""
import os

import numpy as np
import pytest

import icb
import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def port():
    return ol.Port()


@pytest.fixture(scope="module")
def gold():
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    g = np.load(os.path.join(GOLD, "woodcock_sphere_r2b1_l40.npz"))
    w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
    cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
    return cells, g, cam


def scene(port, cells, g, cam, name, acc):
    b = port.grid_bounds(cells)
    return ol.Scene(cells, g[f"lut_{name}"], camera=cam, unit_distance=float(g["unit_distance"]), r_inner=b["sphLo"][0],
                    r_outer=b["sphHi"][0], jitter=1, accum_id=acc)


# ------------------------------------------------------------------------------------------------------------
# CPU
# ------------------------------------------------------------------------------------------------------------
def test_port_shell_builders_reproduce_the_stored_tables(port, gold):
    cells, g, _ = gold
    ranges = port.shell_ranges(cells, g["dims"], g["sphLo"], g["sphHi"])
    assert np.array_equal(ranges, g["ranges"])
    for name in ("smooth", "sharp"):
        assert np.array_equal(port.max_opacities(ranges, g[f"lut_{name}"]), g[f"maxop_{name}"])
    empty = ranges[:, 1] < ranges[:, 0]
    assert empty.mean() < 0.2                                        # a global grid covers (nearly) every (lat, lon) cell
    # lower comes from getValue(height[i]), upper from getValue(height[i+1]) -- not ordered (hostCode.cu:320)
    assert (ranges[~empty, 0] >= cells["value"].min()).all() and (ranges[~empty, 1] <= cells["value"].max()).all()


@pytest.mark.parametrize("name", ["smooth", "sharp"])
def test_port_sphere_frames_golden(gold, port, name):
    cells, g, cam = gold
    W, H = int(g["W"]), int(g["H"])
    for acc in (0, 3):
        fr = port.woodcock_sphere_frame(scene(port, cells, g, cam, name, acc), W, H, g["volLo"], g["volHi"], g["dims"], g["sphLo"], g["sphHi"],
                                        g[f"maxop_{name}"])
        assert np.array_equal(fr["accum"] * np.float32(acc + 1), g[f"sample_{name}_{acc}"])
        if acc == 0:
            assert np.array_equal(fr["fb"], g[f"fb_{name}_{acc}"])
    assert (g[f"sample_{name}_0"][:, 3] > 0).mean() > 0.1


@pytest.mark.parametrize("name", ["smooth", "sharp"])
def test_emulated_device_sphere_frames_golden(gold, port, name):
    """the device text (wsInit / wsStep with the zero-length-callback shortcut, fdlibm asinf/atan2f) on the host"""
    import emu_lib
    emu = emu_lib.Emu()
    cells, g, cam = gold
    emu.build(cells)
    W, H = int(g["W"]), int(g["H"])
    for acc in (0, 3):
        fr = emu.woodcock_sphere_frame(scene(port, cells, g, cam, name, acc), W, H, g["volLo"], g["volHi"], g["dims"], g["sphLo"], g["sphHi"],
                                       g[f"maxop_{name}"])
        assert np.array_equal(fr["accum"] * np.float32(acc + 1), g[f"sample_{name}_{acc}"])
        if acc == 0:
            assert np.array_equal(fr["fb"], g[f"fb_{name}_{acc}"])


@pytest.mark.skipif(not ol.Ref.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("case", [(2, 2, 20, (1, 64, 64)), (2, 3, 90, (1, 1024, 1024)), (3, 1, 35, (3, 40, 17))])
def test_port_shell_builders_equal_the_reference_kernels_live(port, case):
    n, k, levels, dims = case
    ref = ol.Ref()
    cells = icb.gen_grid(n, k, levels)
    b = port.grid_bounds(cells)
    want = ref.shell_ranges(cells, dims, b["sphLo"], b["sphHi"])
    got = port.shell_ranges(cells, dims, b["sphLo"], b["sphHi"])
    assert np.array_equal(got, want)
    for lut, rng in ((icb.smooth_lut(), (0.0, 1.0)), (icb.sharp_lut(64), (0.2, 0.8))):
        assert np.array_equal(port.max_opacities(got, lut, rng), ref.max_opacities(want, lut, rng, shell=True))


@pytest.mark.skipif(not ol.Ref.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("eye", [(0.2e7, 0.9e7, 0.75e7), (-1.1e7, 0.1e7, -0.2e7), (0.05e7, 0.02e7, 1.3e7), (0.0, 0.0, 0.0), (0.3e7, -0.6e7, 0.1e7)])
def test_port_and_device_text_equal_the_reference_live(port, eye):
    """fresh cases against the reference's ray-gen itself: eyes outside the planet (front + back segment), above a pole
    (lon wraps through every cell), at the centre and inside the shell; port and emulated device text, bit for bit"""
    import emu_lib
    ref = ol.Ref()
    cells = icb.gen_grid(2, 2, 20)
    b = port.grid_bounds(cells)
    W, H = 40, 24
    cam = icb.camera_orient(np.array(eye, np.float32), (0.3e7, 0.1e7, 0), (0, 0, 1), W, H)
    lut = icb.sharp_lut(64)
    dims = (1, 200, 333)
    mo = port.max_opacities(port.shell_ranges(cells, dims, b["sphLo"], b["sphHi"]), lut, (0.2, 0.8))
    sc = ol.Scene(cells, lut, tf_range=(0.2, 0.8), opacity_scale=0.9, camera=cam, unit_distance=30000.0, r_inner=b["sphLo"][0],
                  r_outer=b["sphHi"][0], jitter=1, accum_id=2)
    want = ref.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], dims, b["sphLo"], b["sphHi"], mo)
    got = port.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], dims, b["sphLo"], b["sphHi"], mo)
    assert np.array_equal(got["accum"] * np.float32(3), want["accum"])
    emu = emu_lib.Emu()
    emu.build(cells)
    dev = emu.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], dims, b["sphLo"], b["sphHi"], mo)
    assert np.array_equal(dev["accum"] * np.float32(3), want["accum"])
    assert np.array_equal(dev["fb"], got["fb"])


# ------------------------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------------------------
@pytest.fixture()
def ctx():
    c = icb.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["smooth", "sharp"])
def test_gpu_sphere_integrator_is_the_reference_raygen(ctx, gold, name):
    import torch
    cells, g, cam = gold
    W, H = int(g["W"]), int(g["H"])
    ctx.set_grid(cells)
    ctx.set_transfunc(g[f"lut_{name}"])
    ctx.set_accel_shell(*[int(d) for d in g["dims"]])
    ctx.set_integrator("woodcock_sphere")
    mo, vr = ctx.shell_majorants()
    assert np.array_equal(vr, g["ranges"])                        # device builder == the reference's kernels, bit for bit
    assert np.array_equal(mo, g[f"maxop_{name}"])
    ctx.set_camera(cam)
    for acc in (0, 3):
        ctx.set_params(unit_distance=float(g["unit_distance"]), accum_id=acc, jitter=True)
        fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
        ctx.render(fb.data_ptr(), accum.data_ptr(), W, H)
        torch.cuda.synchronize()
        assert np.array_equal(accum.cpu().numpy() * np.float32(acc + 1), g[f"sample_{name}_{acc}"])
        if acc == 0:
            assert np.array_equal(fb.cpu().numpy().view(np.uint32), g[f"fb_{name}_{acc}"])
    st = ctx.stats()
    assert st["raysMarched"] > 0 and st["steps"] > 0


@pytest.mark.gpu
def test_gpu_shell_accel_follows_tf_and_scalar_changes(ctx, port):
    """1 x 1024 x 1024 (the reference's size, the default) on R2B3 x 90, plus a shell with radial cells; the tables are
    re-classified after icb_set_transfunc and rebuilt after icb_set_scalars; frames equal the port's"""
    import torch
    cells = icb.gen_grid(2, 3, 90)
    ctx.set_grid(cells)
    b = ctx.bounds()
    ctx.set_integrator("woodcock_sphere")
    W, H = 48, 32
    cam = icb.camera_orient(np.array([0.2e7, 0.9e7, 0.75e7], np.float32), (0.3e7, 0, 0), (0, 0, 1), W, H)
    ctx.set_camera(cam)
    dims = (1, 1024, 1024)
    cur = cells
    for step, (lut, rng) in enumerate([(icb.smooth_lut(), (0.0, 1.0)), (icb.sharp_lut(), (0.1, 0.9)), (icb.sharp_lut(), (0.1, 0.9)),
                                       (icb.sharp_lut(), (0.1, 0.9))]):
        if step == 2:
            cur = icb.gen_grid(2, 3, 90, timestep=5)
            ctx.set_scalars(cur["value"])
        if step == 3:
            dims = (4, 128, 96)
            ctx.set_accel_shell(*dims)
        ctx.set_transfunc(lut, value_range=rng)
        want_vr = port.shell_ranges(cur, dims, b["sphLo"], b["sphHi"])
        want_mo = port.max_opacities(want_vr, lut, rng)
        mo, vr = ctx.shell_majorants()
        assert np.array_equal(vr, want_vr) and np.array_equal(mo, want_mo)
        ctx.set_params(unit_distance=20000.0, accum_id=step, jitter=True)
        fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
        ctx.render(fb.data_ptr(), accum.data_ptr(), W, H)
        torch.cuda.synchronize()
        sc = ol.Scene(cur, lut, tf_range=rng, camera=cam, unit_distance=20000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0],
                      jitter=1, accum_id=step)
        want = port.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], dims, b["sphLo"], b["sphHi"], want_mo)
        assert np.array_equal(accum.cpu().numpy(), want["accum"])
        assert (want["accum"][:, 3] > 0).mean() > 0.05


@pytest.mark.gpu
@pytest.mark.parametrize("n", [64, 321, 1409])
def test_gpu_sphere_raygen_block_sizes_and_eyes(ctx, port, n):
    """64-, 128- and 256-thread CTAs (by transfer function size) and eyes outside / above a pole / inside the shell"""
    import torch
    cells = icb.gen_grid(2, 2, 30)
    ctx.set_grid(cells)
    b = ctx.bounds()
    ctx.set_integrator("woodcock_sphere")
    dims = (1, 200, 333)
    ctx.set_accel_shell(*dims)
    W, H = 50, 35
    lut = icb.sharp_lut(n)
    ctx.set_transfunc(lut)
    mo, _ = ctx.shell_majorants()
    for i, eye in enumerate([(0.2e7, 0.9e7, 0.75e7), (0.05e7, 0.02e7, 1.3e7), (0.3e7, -0.6e7, 0.1e7)]):
        cam = icb.camera_orient(np.array(eye, np.float32), (0.3e7, 0.1e7, 0), (0, 0, 1), W, H)
        ctx.set_camera(cam)
        ctx.set_params(unit_distance=15000.0, accum_id=i, jitter=True)
        fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
        ctx.render(fb.data_ptr(), accum.data_ptr(), W, H)
        torch.cuda.synchronize()
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=15000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0], jitter=1, accum_id=i)
        want = port.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], dims, b["sphLo"], b["sphHi"], mo)
        assert np.array_equal(accum.cpu().numpy(), want["accum"])
