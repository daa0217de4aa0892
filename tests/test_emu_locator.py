"""CPU: the device locator text (csrc/locator.cuh) compiled for the host by tests/emu, plus the host grid
builder, against the oracle: located record ids and values must be bit-exact, whichever path (in-prism fast
path, vertical relocation, neighbour walk, exact candidate-bin query) resolves a sample."""
import os

import numpy as np
import pytest

import emu_lib
import icb
import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def emu():
    return emu_lib.Emu()


@pytest.fixture(scope="module")
def port():
    return ol.Port()


def unit(lat, lon):
    return np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], -1)


def adversarial_points(cells, rng, n):
    """within metres of side planes, exactly at corners, radii exactly on stored heights (layer/record seams)"""
    lat, lon = cells["lat"].astype(np.float64), cells["lon"].astype(np.float64)
    ci = rng.integers(0, len(cells), n)
    k = rng.integers(0, 3, n)
    A, B = unit(lat[ci, k], lon[ci, k]), unit(lat[ci, (k + 1) % 3], lon[ci, (k + 1) % 3])
    s = rng.uniform(-0.02, 1.02, n)[:, None]
    s[: n // 8] = rng.choice([0.0, 1.0], (n // 8, 1))
    Pd = A * (1 - s) + B * s
    Pd /= np.linalg.norm(Pd, axis=1)[:, None]
    nrm = np.cross(A, B)
    nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    d = rng.uniform(-40, 40, n)
    d[n // 8: n // 4] = 0.0
    h, nl = cells["height"], cells["numLayers"]
    j = (rng.uniform(0, 1, n) * (nl[ci] + 1)).astype(int)
    r = np.where(rng.uniform(0, 1, n) < 0.5, h[ci, j].astype(np.float64), rng.uniform(h[ci, 0], h[ci, nl[ci]]))
    return (Pd * r[:, None] + nrm * d[:, None]).astype(np.float32)


def ray_points(b, rng, nrays, nsteps, dt):
    """consecutive samples along straight rays through the shell (what the walk state is built for)"""
    out = []
    for _ in range(nrays):
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        side = np.cross(d, rng.normal(size=3))
        side /= np.linalg.norm(side)
        p0 = d * rng.uniform(b["sphLo"][0] * 0.97, b["sphHi"][0] * 1.01)
        t = (np.arange(nsteps) - nsteps / 2) * dt
        out.append(p0[None, :] + side[None, :] * t[:, None])
    return np.concatenate(out).astype(np.float32)


def test_golden_locate_all_paths(emu):
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    g = np.load(os.path.join(GOLD, "locate_r2b1_l40.npz"))
    info = emu.build(cells)
    assert info["nRec"] == 640 and info["nCol"] == 320 and info["nVert"] == 162 and info["boundaryEdges"] == 0
    for flags in (0, icb.LOCATE_WALK, icb.LOCATE_EXACT):
        ids, vals, _ = emu.locate(g["pos"], flags)
        assert np.array_equal(ids, g["ids"]) and np.array_equal(vals, g["vals"])


@pytest.mark.parametrize("n,k,levels,margin", [(2, 2, 90, 16.0), (2, 3, 20, 16.0), (3, 1, 35, 2.0), (2, 2, 90, 0.0)])
def test_random_adversarial_and_ray_walks_vs_oracle(emu, port, n, k, levels, margin):
    cells = icb.gen_grid(n, k, levels)
    info = emu.build(cells, margin_m=margin)
    # margin 0 = "never trust the walk": every column takes the exact candidate-bin query
    assert info["nGeneric"] == (0 if margin > 0 else info["nCol"]) and info["boundaryEdges"] == 0
    b = port.grid_bounds(cells)
    rng = np.random.default_rng(1000 * n + 10 * k + levels)
    d = rng.normal(size=(3000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    rnd = (d * rng.uniform(b["sphLo"][0] - 2e4, b["sphHi"][0] + 2e4, 3000)[:, None]).astype(np.float32)
    pos = np.concatenate([rnd, adversarial_points(cells, rng, 4000), ray_points(b, rng, 12, 400, 700.0)])
    want_ids, want_vals = port.locate(cells, pos)
    for flags, chunk in ((0, 0), (icb.LOCATE_WALK, 64), (icb.LOCATE_WALK, 0), (icb.LOCATE_EXACT, 0)):
        ids, vals, ctr = emu.locate(pos, flags, chunk)
        bad = np.nonzero(ids != want_ids)[0]
        assert len(bad) == 0, (flags, chunk, bad[:5], ids[bad[:5]], want_ids[bad[:5]])
        assert np.array_equal(vals, want_vals)
    # along rays the fast path must carry most samples (otherwise the kernel's design premise is void)
    rays = ray_points(b, rng, 8, 2000, 500.0)
    ids, vals, ctr = emu.locate(rays, icb.LOCATE_WALK, 0)
    w_ids, w_vals = port.locate(cells, rays)
    assert np.array_equal(ids, w_ids) and np.array_equal(vals, w_vals)
    if margin > 0:
        assert ctr["slow"] < 0.5 * ctr["steps"]


def test_ragged_and_degenerate_grids_vs_oracle(emu, port):
    """what a real `.ic` can contain: records in shuffled order, a lat/lon window cut out (mesh boundary),
    duplicate records (lowest index must win), zero-layer records, a column with overlapping records."""
    rng = np.random.default_rng(5)
    cells = icb.gen_grid(2, 2, 40)
    perm = rng.permutation(len(cells))
    cells = cells[perm]
    keep = ~((cells["lat"] > 0.9).any(1))
    cells = np.ascontiguousarray(cells[keep])
    extra = cells[:40].copy()                        # duplicates of records 0..39 appended at the end
    extra2 = cells[100:110].copy()
    extra2["numLayers"] = 0                          # accept only r == height[0]
    extra3 = cells[200:210].copy()
    extra3["height"] += np.float32(3000.0)           # overlaps its own column vertically
    extra4 = cells[300:306].copy()                   # clockwise corners: the predicate accepts the ANTIPODAL cone
    extra4["lat"] = extra4["lat"][:, ::-1]
    extra4["lon"] = extra4["lon"][:, ::-1]
    extra5 = cells[400:404].copy()                   # repeated corner: a zero plane normal
    extra5["lat"][:, 2] = extra5["lat"][:, 1]
    extra5["lon"][:, 2] = extra5["lon"][:, 1]
    extra6 = cells[500:502].copy()                   # NaN passes every `> 0` rejection
    extra6["lat"][:, 0] = np.nan
    cells = np.concatenate([cells, extra, extra2, extra3, extra4, extra5, extra6])
    info = emu.build(cells)
    assert info["boundaryEdges"] > 0 and info["nGeneric"] > 0 and info["nGlobal"] >= 22
    b = port.grid_bounds(cells)
    d = rng.normal(size=(4000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    rnd = (d * rng.uniform(b["sphLo"][0] - 2e4, b["sphHi"][0] + 2e4, 4000)[:, None]).astype(np.float32)
    good = np.isfinite(cells["lat"]).all(1)
    pos = np.concatenate([rnd, adversarial_points(cells[good], rng, 4000), ray_points(b, rng, 10, 500, 900.0)])
    anti = -adversarial_points(extra4, rng, 300)      # inside / near the antipodal cones of the cw records
    pos = np.concatenate([pos, anti])
    want_ids, want_vals = port.locate(cells, pos)
    assert (want_ids >= len(cells) - 12).sum() > 0   # some answers do come from the degenerate records
    for flags in (0, icb.LOCATE_WALK, icb.LOCATE_EXACT):
        ids, vals, _ = emu.locate(pos, flags)
        bad = np.nonzero(ids != want_ids)[0]
        assert len(bad) == 0, (flags, bad[:5], ids[bad[:5]], want_ids[bad[:5]])
        assert np.array_equal(vals, want_vals)


def test_bad_records_are_rejected(emu):
    cells = icb.gen_grid(2, 0, 10)
    cells["numLayers"][7] = 32
    with pytest.raises(RuntimeError) as e:
        emu.build(cells)
    assert "record 7" in str(e.value)
    with pytest.raises(RuntimeError):
        emu.build(cells[:0])


def test_bounds_match_oracle(emu, port):
    cells = icb.gen_grid(2, 1, 40)
    emu.build(cells)
    import ctypes as C
    b = icb.Bounds()
    emu.lib.emu_bounds(C.byref(b))
    want = port.grid_bounds(cells)
    for k in ("volLo", "volHi", "sphLo", "sphHi", "dataRange"):
        assert np.array_equal(np.array(getattr(b, k)[:], np.float32), want[k]), k
    assert b.unitDistance == 1000.0        # 10^(floor(log10(6.37e6)) - 3), hostCode.cu:838-840


def test_march_pixel_traces_match_golden(emu):
    """the device marcher text, run on the host, reproduces the reference-composed marcher's id sequences"""
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    blocks = np.load(os.path.join(GOLD, "blocks.npz"))
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    emu.build(cells)
    cam = {k: g[f"sharp_jit_cam_{k}"] for k in ("org", "d00", "du", "dv")}
    sc = ol.Scene(cells, icb.sharp_lut(), camera=cam, unit_distance=16000.0, dt=4000.0, r_inner=blocks["b_sphLo"][0],
                  r_outer=blocks["b_sphHi"][0], jitter=1)
    for (x, y) in [(24, 16), (5, 16), (40, 9), (23, 30)]:
        ids, n, rgba = emu.march_pixel(sc, W, H, x, y)
        assert np.array_equal(ids, g[f"trace_{x}_{y}_ids"])
        assert np.abs(rgba - g[f"trace_{x}_{y}_rgba"]).max() <= 1e-6
    # whole frames: steps / hits / id hash per pixel
    for name, lut, jit in (("smooth", blocks["lut"], 0), ("smooth_inside", blocks["lut"], 1)):
        cam = {k: g[f"{name}_cam_{k}"] for k in ("org", "d00", "du", "dv")}
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=16000.0, dt=4000.0, r_inner=blocks["b_sphLo"][0],
                      r_outer=blocks["b_sphHi"][0], jitter=jit)
        pp = g[name + "_pp"].reshape(H, W, 4)
        acc = g[name + "_accum"].reshape(H, W, 4)
        for y in range(0, H, 3):
            for x in range(0, W, 5):
                ids, n, rgba = emu.march_pixel(sc, W, H, x, y)
                assert n == pp[y, x, 0] and int((ids >= 0).sum()) == pp[y, x, 1]
                assert np.abs(rgba - acc[y, x]).max() <= 1e-5


@pytest.mark.parametrize("case", ["far_zoom", "inside", "grazing", "small_margin"])
def test_skip_ahead_changes_nothing(emu, case):
    """The skip-ahead (locator.cuh skipCount) must leave the located-id sequence, the number of samples and the
    composited colour of every ray bit-identical to marching sample by sample (which the tests above pin against
    the oracle) -- at the real scale: Earth radius, dt = 500 m, camera 2.5e7 m away (2 m ulp) or inside the shell."""
    cells = icb.gen_grid(2, 4, 90)
    margin = 4.0 if case == "small_margin" else 16.0
    emu.build(cells, margin_m=margin)
    import ctypes as C
    bb = icb.Bounds()
    emu.lib.emu_bounds(C.byref(bb))
    b = {k: np.array(getattr(bb, k)[:], np.float32) for k in ("volLo", "volHi", "sphLo", "sphHi")}
    W, H = 256, 144
    if case == "inside":
        eye = np.array([0, 0, float(b["sphLo"][0]) + 2.0e5], np.float32)
        cam = icb.camera_orient(eye, [5e6, 1e6, float(b["sphLo"][0]) + 2.5e5], [0, 0, 1], W, H)
    elif case == "grazing":
        eye = np.array([0.0, float(b["sphHi"][0]) - 1.0e5, 2.4e7], np.float32)
        cam = icb.camera_orient(eye, [0.0, float(b["sphHi"][0]) - 1.0e5, 0.0], [0, 1, 0], W, H, fovy=np.float32(0.02))
    else:
        cam = icb.camera_orbit(b["volLo"], b["volHi"], W, H, 30.0, fovy=np.float32(np.deg2rad(35.0)), pose_fovy=icb.FOVY_DEFAULT)
    rng = np.random.default_rng(abs(hash(case)) % 1000)
    total, skipped, rays = 0, 0, 0
    for lut in (icb.sharp_lut(), icb.smooth_lut()):
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=1000.0, dt=500.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0],
                      jitter=1, term_alpha=0.99 if lut[0, 3] == 0 else 2.0)
        for _ in range(60):
            x, y = int(rng.integers(W)), int(rng.integers(H))
            ids0, n0, c0 = emu.march_pixel(sc, W, H, x, y)
            ids1, n1, c1 = emu.march_pixel(sc, W, H, x, y, skip=True, margin_m=margin)
            assert n0 == n1 and np.array_equal(ids0, ids1), (case, x, y, n0, n1)
            assert np.array_equal(c0, c1)
            total += n1
            skipped += emu.lib.emu_last_skipped()
            rays += n1 > 0
    assert rays > 20 and total > 20000
    assert skipped > 0.5 * total, (skipped, total)       # and it must be worth having


def test_woodcock_integrator_matches_the_reference_bit_for_bit(emu, port):
    """The reference's own integrator (woodcockTrackingAE + woodcockTracking, deviceCode.cu:149-186, 239-275) run by
    the device text (woodcockPixel: locator state machine, host-tabulated logf) against the oracle: the frame --
    every collision position, located record, accepted sample, accumulated colour and packed pixel -- is identical."""
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    blocks = np.load(os.path.join(GOLD, "blocks.npz"))
    W, H = int(g["W"]), int(g["H"])
    cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
    emu.build(cells)
    b = port.grid_bounds(cells)
    oracles = [port] + ([ol.Ref()] if ol.Ref.available() else [])
    for acc in (0, 3):
        sc = ol.Scene(cells, blocks["lut"], camera=cam, unit_distance=float(w["unit_distance"]), r_inner=b["sphLo"][0],
                      r_outer=b["sphHi"][0], jitter=1, accum_id=acc)
        got = emu.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
        assert got["rays"] > 0.3 * W * H and got["hits"] > 0
        for o in oracles:
            if isinstance(o, ol.Ref):
                want = o.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
                want_accum = want["accum"] / np.float32(acc + 1)
            else:
                want = o.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
                want_accum = want["accum"]
            assert np.array_equal(got["fb"], want["fb"]), type(o).__name__
            if not isinstance(o, ol.Ref):
                assert np.array_equal(got["accum"], want_accum), type(o).__name__
            else:
                assert np.allclose(got["accum"], want_accum, rtol=0, atol=1e-7)


def test_woodcock_frames_on_degenerate_grids(emu, port):
    """The Woodcock ray-gens cull samples outside [min height[0], max height[numLayers]] before locating
    (ICONGrid.h:184 rejects them for every record).  Records that bend that rule must not: NaN-cornered records (accepted
    wherever their radius range holds), zero-layer and shifted records (they move the range), and a record whose
    height[0] is NaN (its radius test passes everything below -- the builder switches the cull off)."""
    base = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
    blocks = np.load(os.path.join(GOLD, "blocks.npz"))
    cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
    W, H = 24, 16
    cam = dict(cam)
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    cam["du"] = cam["du"] * np.float32(int(g["W"]) / W)
    cam["dv"] = cam["dv"] * np.float32(int(g["H"]) / H)
    variants = {}
    e1 = base[10:12].copy(); e1["lat"][:, 0] = np.nan                     # accepted wherever r is in its range
    e2 = base[20:24].copy(); e2["height"] += np.float32(40000.0)           # raises the outer bound
    e3 = base[30:32].copy(); e3["numLayers"] = 0
    variants["nan corners, shifted, zero-layer"] = np.concatenate([base, e1, e2, e3])
    e4 = base[40:41].copy(); e4["height"][:, 0] = np.nan                  # r < NaN is false: accepted all the way down
    variants["nan height[0]"] = np.concatenate([base, e4])
    for name, cells in variants.items():
        try:
            emu.build(cells)
        except RuntimeError as err:                                        # (a builder that refuses the record is also fine)
            assert "record" in str(err), name
            continue
        b = port.grid_bounds(cells)
        sc = ol.Scene(cells, blocks["lut"], camera=cam, unit_distance=float(w["unit_distance"]) * 4, r_inner=b["sphLo"][0],
                      r_outer=b["sphHi"][0], jitter=1, accum_id=1)
        got = emu.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
        want = port.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
        assert np.array_equal(got["accum"], want["accum"]), name
        assert (want["accum"][:, 3] > 0).any(), name


@pytest.mark.parametrize("chunk_log2", [4, 6, 10])
def test_chunked_compositing_is_the_plain_marcher_within_rounding(emu, port, chunk_log2):
    """Split rays: the CUDA marcher composites a ray in chunks of 2^k consecutive sample indices per shell
    segment and merges the chunks in order with `over`, so that the chunks of a long limb ray can be marched by different
    lanes.  The CPU statement of that order (march_pixel with chunk > 0, in the reference-function marcher oracle/_ref and in
    the C port) must stay within the north-star tolerance 1e-3 of the plain sample-by-sample statement (measured: a few
    ulps), with the same samples and the same located ids; and the device text follows the chunked statement."""
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    blocks = np.load(os.path.join(GOLD, "blocks.npz"))
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    W, H = int(g["W"]), int(g["H"])
    emu.build(cells)
    chunk = 1 << chunk_log2
    worst = 0.0
    for name, lut, jit in (("smooth", blocks["lut"], 0), ("sharp_jit", icb.sharp_lut(), 1), ("smooth_inside", blocks["lut"], 1)):
        cam = {k: g[f"{name}_cam_{k}"] for k in ("org", "d00", "du", "dv")}
        kw = dict(camera=cam, unit_distance=16000.0, dt=1000.0, r_inner=blocks["b_sphLo"][0], r_outer=blocks["b_sphHi"][0], jitter=jit)
        plain = port.march_frame(ol.Scene(cells, lut, **kw), W, H)
        chunked = port.march_frame(ol.Scene(cells, lut, chunk=chunk, **kw), W, H)
        assert np.array_equal(plain["perPixel"], chunked["perPixel"])                  # same samples, same ids, same stop
        d = float(np.abs(plain["accum"] - chunked["accum"]).max())
        worst = max(worst, d)
        assert d <= 1e-3
        assert plain["perPixel"][:, 0].max() > chunk or chunk_log2 == 10                # rays really span several chunks
        if ol.Ref.available():                                                           # the reference-function marcher, same order
            r = ol.Ref().march_frame(ol.Scene(cells, lut, chunk=chunk, **kw), W, H, rect=(10, 8, 30, 20))
            m = np.zeros((H, W), bool); m[8:20, 10:30] = True
            assert np.array_equal(r["accum"][m.reshape(-1)], chunked["accum"][m.reshape(-1)])
        # the device text (marchPixel) against the chunked statement: same samples; colours to the last bits (expf differs)
        sc = ol.Scene(cells, lut, chunk=chunk, **kw)
        acc = chunked["accum"].reshape(H, W, 4)
        pp = chunked["perPixel"].reshape(H, W, 4)
        for y in range(1, H, 5):
            for x in range(2, W, 7):
                for skip in (False, True):
                    ids, n, rgba = emu.march_pixel(sc, W, H, x, y, skip=skip, chunk_log2=chunk_log2)
                    assert n == pp[y, x, 0] and int((ids >= 0).sum()) == pp[y, x, 1]
                    assert np.abs(rgba - acc[y, x]).max() <= 1e-5
    assert worst < 1e-5


def test_band_samples_from_both_sides_of_an_edge(emu, port):
    """The band test of locateAction (a sample within the margin of exactly one side plane: this record and the one across the
    edge, each with its own planes, lower caller index wins) entered with the walk state in EITHER of the two records -- so
    that both outcomes occur: the record just fetched wins (it stays), the record left wins (the lane returns through the
    shared edge).  Pairs (a point 2 km inside a column, a point within +-12 m of one of its edges), located as a walk of two."""
    n, k, levels = 2, 3, 40
    cells = icb.gen_grid(n, k, levels)
    emu.build(cells)
    rng = np.random.default_rng(77)
    m = 6000
    lat, lon = cells["lat"].astype(np.float64), cells["lon"].astype(np.float64)
    ci = rng.integers(0, len(cells), m)
    kk = rng.integers(0, 3, m)
    A, B = unit(lat[ci, kk], lon[ci, kk]), unit(lat[ci, (kk + 1) % 3], lon[ci, (kk + 1) % 3])
    Cn = unit(lat[ci, 0], lon[ci, 0]) + unit(lat[ci, 1], lon[ci, 1]) + unit(lat[ci, 2], lon[ci, 2])
    s = rng.uniform(0.25, 0.75, m)[:, None]
    Pd = A * (1 - s) + B * s
    Pd /= np.linalg.norm(Pd, axis=1)[:, None]
    nrm = np.cross(A, B)
    nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    inward = np.sign(np.einsum("ij,ij->i", nrm, Cn))[:, None]
    h, nl = cells["height"].astype(np.float64), cells["numLayers"]
    j = (rng.uniform(0, 1, m) * nl[ci]).astype(int)
    r = (0.5 * (h[ci, j] + h[ci, j + 1]))[:, None]
    d = rng.uniform(-12.0, 12.0, m)[:, None]                       # signed distance from the edge plane, inside the 16 m margin
    side = np.where(rng.uniform(0, 1, m) < 0.5, 1.0, -1.0)[:, None]   # start in this column or in the one across the edge
    p0 = Pd * r + nrm * inward * side * 2000.0
    p1 = Pd * r + nrm * inward * d
    pos = np.empty((2 * m, 3), np.float32)
    pos[0::2], pos[1::2] = p0, p1
    want_ids, want_vals = port.locate(cells, pos)
    ids, vals, ctr = emu.locate(pos, icb.LOCATE_WALK, 2)
    assert np.array_equal(ids, want_ids) and np.array_equal(vals, want_vals)
    # both outcomes occurred: band points answered by the record of the start point and by the other one
    start, band = want_ids[0::2], want_ids[1::2]
    hit = (start >= 0) & (band >= 0)
    assert (band[hit] == start[hit]).sum() > m // 8 and (band[hit] != start[hit]).sum() > m // 8
    # and they were resolved by the walk, not handed to the exact query (the start points are: a walk of two has no history)
    assert ctr["exact"] < m + m // 4, ctr
