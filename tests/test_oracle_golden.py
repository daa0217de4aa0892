"""CPU: pins the C restatement (oracle/icon_oracle.c) against golden vectors produced by the REFERENCE's own
sources (oracle/_ref, tools/gen_golden.py) and, when oracle/_ref is present, against the reference directly."""
import os

import numpy as np
import pytest

import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def port():
    return ol.Port()


@pytest.fixture(scope="module")
def cells():
    return np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), ol.CELL_DTYPE)


@pytest.fixture(scope="module")
def blocks():
    return np.load(os.path.join(GOLD, "blocks.npz"))


def test_record_layout(port, cells):
    assert port.sizeof_cell() == 284          # ICONGrid.h:59-76, SURVEY Appendix B
    assert len(cells) == 640 and set(cells["numLayers"]) == {31, 9}


def test_building_blocks_bit_exact(port, cells, blocks):
    g = blocks
    assert np.array_equal(port.to_cartesian(g["sph"]), g["cart"])
    assert np.array_equal(port.cell_planes(cells), g["planes"])
    assert np.array_equal(port.resample_lut(ol.DEFAULT_LUT5, 300), g["lut"])
    assert np.array_equal(port.post_classify(g["lut"], 0.0, 1.0, 1.0, g["v"]), g["cls"])
    assert np.array_equal(port.post_classify(g["lut"], 0.1, 0.9, 2.5, g["v"]), g["cls2"])
    seeds = [(0, 0), (1, 2), (511, 77), (262144 * 3 + 5, 400), (2**32 - 1, 12345)]
    assert np.array_equal(np.stack([port.lcg(a, b, 16) for a, b in seeds]), g["lcg"])
    assert np.array_equal(port.find_height(cells[4:5], g["fh_r"]), g["fh"])
    hit, tn, tf = port.intersect_sphere(g["org"], g["dirs"], float(g["b_sphHi"][0]))
    assert np.array_equal(hit, g["sh_hit"])
    m = hit.astype(bool)
    assert np.array_equal(tn[m], g["sh_tn"][m]) and np.array_equal(tf[m], g["sh_tf"][m])
    for i in range(len(g["org"])):
        s = port.shell_segments(g["org"][i], g["dirs"][i], 0.0, 1e10, float(g["b_sphLo"][0]), float(g["b_sphHi"][0]))
        assert len(s) == g["nseg"][i]
        assert np.array_equal(s.reshape(-1), g["segs"][i, : 2 * len(s)])
    assert np.array_equal(port.srgb_pack(g["rgba"]), g["packed"])
    b = port.grid_bounds(cells)
    for k in b:
        assert np.array_equal(b[k], g["b_" + k]), k
    cam = port.camera_view_all(g["b_volLo"], g["b_volHi"], 48 / 32, 48, 32)
    for k in cam:
        assert np.array_equal(cam[k], g["camva_" + k]), k
    cam = port.camera_orient([1.2e7, 3e6, 2.0e7], [0, 0, 0], [0, 1, 0], ol.FOVY_DEFAULT * 0.6, 1.5, 48, 32)
    for k in cam:
        assert np.array_equal(cam[k], g["camor_" + k]), k


def test_postclassify_quirks(port, blocks):
    """frac weights the LOWER entry, opacityScale multiplies only the upper term, idx=int(v*size)."""
    lut = np.zeros((4, 4), np.float32)
    lut[:, 3] = [0.0, 1.0, 2.0, 3.0]
    out = port.post_classify(lut, 0.0, 1.0, 10.0, np.array([0.3125], np.float32))   # v*size=1.25 -> idx 1, frac .25
    assert out[0, 3] == np.float32(1.0 * 0.25 + 2.0 * 0.75 * 10.0)
    out = port.post_classify(lut, 0.0, 1.0, 1.0, np.array([-0.4, 1.7], np.float32))  # clamped both ends
    assert np.isfinite(out).all()


def test_locate_golden(port, cells):
    g = np.load(os.path.join(GOLD, "locate_r2b1_l40.npz"))
    ids, vals = port.locate(cells, g["pos"])
    assert np.array_equal(ids, g["ids"]) and np.array_equal(vals, g["vals"])
    assert (g["ids"] >= 0).mean() > 0.5 and (g["ids"] < 0).sum() > 100     # both outcomes are covered


def test_disabled_fixture_cell(port):
    """the one known-answer fixture in the reference (hostCode.cu:768-790, #if 0): 2 layers 100/110/120,
    values 0.1/1.0 -> r=105 samples 0.1, r=115 samples 1.0, boundary r=110 takes the lower layer."""
    c = np.zeros(1, ol.CELL_DTYPE)
    c["lon"][0] = np.deg2rad([30.0, 0.0, -30.0]).astype(np.float32)
    c["lat"][0] = np.deg2rad([0.0, 90.0, 0.0]).astype(np.float32)
    c["numLayers"] = 2
    c["height"][0, :3] = [100, 110, 120]
    c["value"][0, :2] = [0.1, 1.0]
    d = np.array([np.cos(np.deg2rad(30.0)), 0.0, np.sin(np.deg2rad(30.0))])
    pos = np.stack([d * 105, d * 115, d * 125, d * 99, -d * 105]).astype(np.float32)
    ids, vals = port.locate(c, pos)
    assert ids.tolist() == [0, 0, -1, -1, -1]
    assert vals[0] == np.float32(0.1) and vals[1] == np.float32(1.0)
    assert port.find_height(c, np.array([110.0], np.float32))[0] == 0


def _scene(cells, lut, g, name, jitter, acc, blocks):
    cam = {k: g[f"{name}_cam_{k}"] for k in ("org", "d00", "du", "dv")}
    return ol.Scene(cells, lut, camera=cam, unit_distance=16000.0, dt=4000.0, r_inner=blocks["b_sphLo"][0],
                    r_outer=blocks["b_sphHi"][0], jitter=jitter, accum_id=acc)


FRAMES = [("smooth", "smooth", 0, 0), ("sharp_jit", "sharp", 1, 0), ("smooth_inside", "smooth", 1, 0), ("smooth_acc3", "smooth", 1, 3)]


def _lut(kind, blocks):
    lut = blocks["lut"].copy()
    if kind == "sharp":
        lut[:, 3] = 0.0
        lut[150:, 3] = 1.0
    return lut


@pytest.mark.parametrize("name,kind,jit,acc", FRAMES)
def test_march_frames_golden(port, cells, blocks, name, kind, jit, acc):
    """A.3 marcher: per-pixel step/hit counts and the hash of the located-id sequence are bit-exact; the image
    agrees to float rounding of expf (glibc on both sides here -> exact)."""
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    sc = _scene(cells, _lut(kind, blocks), g, name, jit, acc, blocks)
    prev = g[name + "_prev"].copy() if acc else None
    out = port.march_frame(sc, W, H, accum=prev)
    assert np.array_equal(out["perPixel"], g[name + "_pp"])
    assert np.array_equal(out["accum"], g[name + "_accum"])
    assert np.array_equal(out["fb"], g[name + "_fb"])


def test_march_traces_golden(port, cells, blocks):
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    sc = _scene(cells, _lut("sharp", blocks), g, "sharp_jit", 1, 0, blocks)
    for (x, y) in [(24, 16), (5, 16), (40, 9), (23, 30)]:
        ids, n, rgba = port.march_trace(sc, W, H, x, y)
        assert np.array_equal(ids, g[f"trace_{x}_{y}_ids"])
        assert np.array_equal(rgba, g[f"trace_{x}_{y}_rgba"])


def test_woodcock_cross_check(port, cells, blocks):
    """The reference's integrator (Woodcock tracking, deviceCode.cu:149-186) estimates the same
    emission-absorption integral the marcher integrates: its 16-frame mean alpha (golden, from the reference)
    must agree with the marched alpha within Monte-Carlo noise (SURVEY §8c acceptance)."""
    w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
    ud = float(w["unit_distance"])
    sc = ol.Scene(cells, blocks["lut"], camera=cam, unit_distance=ud, dt=ud / 8, term_alpha=2.0,
                  r_inner=blocks["b_sphLo"][0], r_outer=blocks["b_sphHi"][0])
    m = port.march_frame(sc, W, H)["accum"]
    n = int(w["n_frames"])
    mean, meansq = w["mean"].astype(np.float64), w["meansq"].astype(np.float64)
    # Woodcock alpha is Bernoulli(p) per frame with p = 1 - transmittance; frame mean over pixels is tight
    assert abs(mean[:, 3].mean() - m[:, 3].mean()) < 0.03
    var = np.maximum(meansq[:, 3] - mean[:, 3] ** 2, 0.0) + 0.25 / n
    z = (mean[:, 3] - m[:, 3]) / np.sqrt(var / n + 1e-4)
    assert np.mean(np.abs(z) < 4.0) > 0.97


@pytest.mark.skipif(not ol.Ref.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_port_equals_reference_live(port, cells, blocks):
    """restatement vs the reference's own compiled sources on fresh random inputs"""
    ref = ol.Ref()
    rng = np.random.default_rng(7)
    assert ref.sizeof_cell() == 284
    sph = np.stack([rng.uniform(6.3e6, 7.2e6, 2000), rng.uniform(-np.pi / 2, np.pi / 2, 2000), rng.uniform(-np.pi, np.pi, 2000)], -1)
    assert np.array_equal(port.to_cartesian(sph), ref.to_cartesian(sph))
    assert np.array_equal(port.cell_planes(cells), ref.cell_planes(cells))
    d = rng.normal(size=(3000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos = (d * rng.uniform(blocks["b_sphLo"][0] - 1e4, blocks["b_sphHi"][0] + 1e4, 3000)[:, None]).astype(np.float32)
    a, b = port.locate(cells, pos), ref.locate(cells, pos)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    v = rng.uniform(-0.5, 1.5, 4000).astype(np.float32)
    assert np.array_equal(port.post_classify(blocks["lut"], 0.2, 0.7, 3.0, v), ref.post_classify(blocks["lut"], 0.2, 0.7, 3.0, v))
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    sc = _scene(cells, blocks["lut"], g, "smooth", 1, 2, blocks)
    xy = np.stack(np.meshgrid(np.arange(W), np.arange(H)), -1).reshape(-1, 2)
    ro, rd = ref.generate_rays(sc, xy, W, H)
    po, pd = port.generate_rays(sc, xy, W, H)
    assert np.array_equal(ro, po) and np.array_equal(rd, pd)
    rect = (10, 8, 26, 20)
    a = port.march_frame(sc, W, H, rect=rect, accum=np.full((W * H, 4), 0.25, np.float32))
    b = ref.march_frame(sc, W, H, rect=rect, accum=np.full((W * H, 4), 0.25, np.float32))
    assert np.array_equal(a["perPixel"], b["perPixel"]) and np.array_equal(a["accum"], b["accum"]) and np.array_equal(a["fb"], b["fb"])


@pytest.mark.skipif(not ol.Ref.available(), reason="oracle/_ref not built")
def test_reference_sdda_is_a_two_segment_clipper(blocks):
    """SURVEY finding 3: the reference's sdda yields exactly one non-empty callback per radial segment; their
    [t0,t1] are the un-shrunk ranges (ShellAccel.h:94-111; sceneEPS only offsets the grid projection, :115-116)
    which the A.3 clipper shrinks by sceneEPS on both ends."""
    ref = ol.Ref()
    g = blocks
    n_checked = 0
    for i in range(0, 256, 5):
        if g["nseg"][i] == 0:
            continue
        t, leaf, n = ref.sdda_callbacks(g["org"][i], g["dirs"][i], 0.0, 1e10, g["b_sphLo"], g["b_sphHi"])
        real = t[t[:, 1] > t[:, 0]]
        segs = g["segs"][i, : 2 * g["nseg"][i]].reshape(-1, 2)
        assert len(real) == len(segs)
        eps = np.float32(g["b_sphLo"][0]) * np.float32(1e-6)
        assert np.array_equal(real[:, 0] + eps, segs[:, 0]) and np.array_equal(real[:, 1] - eps, segs[:, 1])
        n_checked += 1
    assert n_checked > 10


def test_woodcock_frames_golden(port, cells, blocks):
    """the C restatement of woodcockTrackingAE reproduces the reference's own frames (tests/golden, from oracle/_ref
    via tools/gen_golden_woodcock.py) bit for bit -- RNG stream, logf free flights, collisions, accumulation, packing"""
    wf = np.load(os.path.join(GOLD, "woodcock_frames_r2b1_l40.npz"))
    w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
    W, H = int(wf["W"]), int(wf["H"])
    cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
    for acc in (0, 3):
        sc = ol.Scene(cells, blocks["lut"], camera=cam, unit_distance=float(wf["unit_distance"]),
                      r_inner=blocks["b_sphLo"][0], r_outer=blocks["b_sphHi"][0], jitter=1, accum_id=acc)
        got = port.woodcock_frame(sc, W, H, wf["volLo"], wf["volHi"])
        assert np.array_equal(got["fb"], wf[f"fb_{acc}"])
        assert np.allclose(got["accum"] * np.float32(acc + 1), wf[f"sample_{acc}"], rtol=0, atol=1e-6)
