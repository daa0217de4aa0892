"""GPU (-m gpu): the drop-in executable icon_rt_b200 (host/icon_rt_main.cpp) -- the reference's non-interactive `icon_rt`
(hostCode.cu:703-968 main, pipeline.cu:224-253 flags, :127-169 `.xf`, :733-739 PNG + FPS line) -- against the library it
wraps: the PNG it writes is the frame icb_render_host returns for the same grid, camera and transfer function; `.ic` and
`.xf` files round-trip; --num-cells / --lat-range / --lon-range select the records icb_load_ic selects (hostCode.cu:717-758);
--gpus N assembles the same frame sort-first."""
import os
import re
import struct
import subprocess
import zlib

import numpy as np
import pytest

import icb

pytestmark = pytest.mark.gpu
EXE = os.path.join(os.path.dirname(icb.LIB_PATH), "icon_rt_b200")


def read_png(path):
    d = open(path, "rb").read()
    assert d[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(d):
        n, typ = struct.unpack(">I4s", d[pos:pos + 8])
        body = d[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", d[pos + 8 + n:pos + 12 + n])[0] == (zlib.crc32(typ + body) & 0xffffffff)
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 6)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * 4 + 1)
    assert (raw[:, 0] == 0).all()                                   # filter type none
    return np.ascontiguousarray(raw[:, 1:]).view(np.uint32).reshape(h, w)


def read_xf(path):
    d = open(path, "rb").read()
    opacity, lo, hi, r0, r1, n = struct.unpack("<fffffi", d[:24])
    return np.frombuffer(d[24:24 + 16 * n], np.float32).reshape(n, 4).copy(), (lo, hi), opacity


def run(args, cwd):
    out = subprocess.run([EXE] + [str(a) for a in args], cwd=cwd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    return out.stdout


def library_frame(cells, W, H, lut, rng, opacity, jitter, dt=0.0, integrator="march"):
    ctx = icb.Context(0)
    try:
        ctx.set_grid(cells)
        b = ctx.bounds()
        ctx.set_transfunc(lut, value_range=rng, opacity_scale=opacity)
        ctx.set_camera(icb.camera_view_all(b["volLo"], b["volHi"], W, H))
        ctx.set_integrator(integrator)
        ctx.set_params(unit_distance=b["unitDistance"], dt=dt, jitter=jitter)
        fb, _ = ctx.render_host(W, H)
        return fb.reshape(H, W)
    finally:
        ctx.close()


def test_png_is_the_library_frame_and_files_round_trip(tmp_path):
    W, H = 96, 64
    out = run(["--gen", 2, 3, 40, "--size", W, H, "--sharp", "--jitter", "-o", "a.png", "--save-ic", "g.ic", "--save-xf", "t.xf"], tmp_path)
    cells = icb.gen_grid(2, 3, 40)
    assert int(re.search(r"numCells: (\d+)", out).group(1)) == len(cells)
    assert re.search(r"FPS: [0-9.]+", out) and "Output: a.png" in out
    assert open(tmp_path / "g.ic", "rb").read() == cells.tobytes()            # --save-ic: the records, 284 B each
    lut, rng, opacity = read_xf(tmp_path / "t.xf")
    assert len(lut) == 300 and np.array_equal(lut, icb.sharp_lut())           # default LUT resampled to 300 (pipeline.cu:469-473), --sharp
    png = read_png(tmp_path / "a.png")
    want = library_frame(cells, W, H, lut, rng, opacity, jitter=True)
    assert png.shape == (H, W) and np.array_equal(png, want[::-1])            # rows flipped like stbi_flip_vertically_on_write
    assert (png != 0).mean() > 0.02
    # the same frame from the files written above
    run(["g.ic", "--xf", "t.xf", "--size", W, H, "--jitter", "-o", "b.png"], tmp_path)
    assert np.array_equal(read_png(tmp_path / "b.png"), png)
    lut2, rng2, op2 = read_xf(tmp_path / "t.xf")
    run(["g.ic", "--xf", "t.xf", "--size", W, H, "--save-xf", "u.xf", "-o", "c.png"], tmp_path)
    assert open(tmp_path / "u.xf", "rb").read() == open(tmp_path / "t.xf", "rb").read()


def test_record_filters_are_the_loaders(tmp_path):
    W, H = 80, 48
    cells = icb.gen_grid(2, 2, 20)
    cells.tofile(tmp_path / "g.ic")
    args = ["--num-cells", 2500, "--lat-range", "-35:60", "--lon-range", "-20:140"]
    out = run(["g.ic", "--size", W, H, "--sharp", "-o", "f.png", "--save-xf", "t.xf"] + args, tmp_path)
    sel = icb.load_ic(str(tmp_path / "g.ic"), max_num_cells=2500, lat_range=(-35.0, 60.0), lon_range=(-20.0, 140.0))
    head = cells[:2500]
    rad = lambda d: np.float32(d) * np.float32(np.pi) / np.float32(180.0)            # deg2rad, ICONGrid.h:26-29
    lat, lon = (rad(-35), rad(60)), (rad(-20), rad(140))
    keep = ((head["lat"] >= lat[0]) & (head["lat"] <= lat[1]) & (head["lon"] >= lon[0]) & (head["lon"] <= lon[1])).all(axis=1)
    assert 0 < keep.sum() < 2500 and np.array_equal(sel, head[keep])          # hostCode.cu:717-758
    assert int(re.search(r"numCells: (\d+)", out).group(1)) == len(sel)
    lut, rng, opacity = read_xf(tmp_path / "t.xf")
    want = library_frame(sel, W, H, lut, rng, opacity, jitter=False)
    assert np.array_equal(read_png(tmp_path / "f.png"), want[::-1])


@pytest.mark.parametrize("integrator", ["woodcock", "woodcock_grid", "woodcock_sphere"])
def test_reference_raygens_from_the_command_line(tmp_path, integrator):
    """--sample-limit n frames of the reference's ray-gens (running mean over accumID, pipeline.cu:762-765)"""
    W, H = 64, 40
    run(["--gen", 2, 2, 20, "--size", W, H, "--integrator", integrator, "--sample-limit", 3, "-o", "w.png", "--save-xf", "t.xf"], tmp_path)
    lut, rng, opacity = read_xf(tmp_path / "t.xf")
    ctx = icb.Context(0)
    try:
        cells = icb.gen_grid(2, 2, 20)
        ctx.set_grid(cells)
        b = ctx.bounds()
        ctx.set_transfunc(lut, value_range=rng, opacity_scale=opacity)
        ctx.set_camera(icb.camera_view_all(b["volLo"], b["volHi"], W, H))
        ctx.set_integrator(integrator)
        fb = np.zeros(W * H, np.uint32)
        accum = np.zeros((W * H, 4), np.float32)
        for f in range(3):
            ctx.set_params(unit_distance=b["unitDistance"], accum_id=f, jitter=False)
            ctx.render_host(W, H, fb=fb, accum=accum)
        assert np.array_equal(read_png(tmp_path / "w.png"), fb.reshape(H, W)[::-1])
    finally:
        ctx.close()


def test_gpus_flag_assembles_the_same_frame(tmp_path):
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs 2 GPUs")
    W, H = 200, 120
    run(["--gen", 2, 3, 40, "--size", W, H, "--sharp", "--jitter", "-o", "one.png"], tmp_path)
    out = run(["--gen", 2, 3, 40, "--size", W, H, "--sharp", "--jitter", "--gpus", n, "-o", "many.png"], tmp_path)
    assert np.array_equal(read_png(tmp_path / "many.png"), read_png(tmp_path / "one.png"))
    assert re.search(r"FPS: [0-9.]+", out)
