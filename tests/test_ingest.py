"""the converter's array -> `.ic` half (icb_ingest_columns) against the REFERENCE converter.
tests/golden/convert_icon.npz holds inputs and the records tools/convert_icon/convert_icon.cpp wrote for them (compiled where
it lies against a stand-in netcdf.h: oracle/ref_convert.cpp, tools/gen_golden_convert.py); where oracle/_ref/convert_icon_ref
exists the reference is also run live on fresh inputs.  Bit-exact."""
import os
import sys
import tempfile

import numpy as np
import pytest

import icb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_golden_convert as gg  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "convert_icon.npz"))
CASES = sorted({k.rsplit("_records", 1)[0] for k in G.files if k.endswith("_records")})


def ours(inp, cap, mode=icb.INGEST_REFERENCE, order_seed=5):
    rng = np.random.default_rng(order_seed)                    # level fields handed over in an arbitrary file order
    hp, dp = rng.permutation(len(inp["hhl_height"])), rng.permutation(len(inp["data_height"]))
    data = [icb.ingest_normalize(inp["data"][i]) for i in dp]
    hhl = [inp["hhl"][i].astype(np.float32) for i in hp]
    return icb.ingest_columns(inp["clon"], inp["clat"], inp["hsurf"], inp["hhl_height"][hp], hhl, inp["data_height"][dp], data,
                              max_layers=cap, mode=mode)


def case_inputs(name):
    return {k: G[f"{name}_{k}"] for k in ("clat", "clon", "hsurf", "hhl", "hhl_height", "data", "data_height")}, int(G[f"{name}_cap"])


@pytest.mark.parametrize("name", CASES)
def test_records_equal_the_reference_converters(name):
    inp, cap = case_inputs(name)
    rec = ours(inp, cap)
    want = G[f"{name}_records"].view(icb.CELL_DTYPE).reshape(-1)
    assert len(rec) == len(want)
    assert rec.tobytes() == want.tobytes()


def test_reference_defects_are_reproduced_and_fixed_mode_repairs_them():
    inp, cap = case_inputs("l90")
    cells = len(inp["hsurf"])
    ref = ours(inp, cap)
    fix = ours(inp, cap, mode=icb.INGEST_FIXED)
    assert ref["numLayers"][:3].tolist() == [31, 31, 25]          # (90 % 32) - 1, convert_icon.cpp:365
    assert fix["numLayers"][:3].tolist() == [31, 31, 28]
    assert ref[0::3].tobytes() == fix[0::3].tobytes()      # the full records are the same records
    assert np.array_equal(fix["value"][2::3, :25], ref["value"][2::3, :25])
    # every data level lands in exactly one layer, bottom-up (descending `height`), normalised per level
    col = np.concatenate([fix["value"][0, :31], fix["value"][1, :31], fix["value"][2, :28]])
    want = np.array([icb.ingest_normalize(inp["data"][i])[0] for i in np.argsort(-inp["data_height"])])
    assert np.array_equal(col, want)
    # a column's records are stacked: H[0] of a record is the last height of the one below it
    assert np.array_equal(fix["height"][1::3, 0], fix["height"][0::3, 31]) and np.array_equal(fix["height"][2::3, 0], fix["height"][1::3, 31])
    inp32, _ = case_inputs("l32")
    assert ours(inp32, 1000)["numLayers"][:2].tolist() == [31, -1]                 # the reference writes a record with -1 layers
    assert ours(inp32, 1000, mode=icb.INGEST_FIXED)["numLayers"][:2].tolist() == [31, 1]
    assert len(ours(inp, 5)) == cells and ours(inp, 5)["numLayers"][0] == 4        # maxLayers{5} caps silently (:24, :347-349)


def test_too_few_level_fields_is_an_error_not_a_read_past_the_end():
    inp, _ = case_inputs("l20")
    inp = dict(inp, hhl=inp["hhl"][:10], hhl_height=inp["hhl_height"][:10])
    with pytest.raises(icb.IcbError):
        ours(inp, 1000)


@pytest.mark.skipif(not os.path.exists(gg.REF_BIN), reason="oracle/_ref/convert_icon_ref not built (needs /root/reference)")
@pytest.mark.parametrize("cells,levels,cap", [(41, 7, 5), (13, 45, 45), (7, 93, 1000), (5000, 31, 31)])
def test_live_reference_on_fresh_inputs(cells, levels, cap):
    inp = gg.make_inputs(cells, levels, seed=cells * 131 + levels)
    with tempfile.TemporaryDirectory() as d:
        want = gg.run_reference(inp, cap, d)
    rec = ours(inp, cap)
    assert rec.tobytes() == want.tobytes()


def _atmosphere(n, k, levels, land):
    """converter inputs for the generator's icosahedral mesh: corner angles of every column, a surface height, `levels` + 1
    half levels (ICON numbering: 1 = model top) and `levels` data levels"""
    cols = icb.gen_grid(n, k, 1)                                   # one record per column: the mesh
    ncol = len(cols)
    clat, clon = cols["lat"].astype(np.float64), cols["lon"].astype(np.float64)
    rng = np.random.default_rng(3)
    hsurf = rng.uniform(0.0, 3000.0, ncol) * (rng.random(ncol) < 0.5) if land else np.zeros(ncol)
    kk = np.arange(1, levels + 2)
    hhl = hsurf[None, :] + (75000.0 - hsurf[None, :]) * (((levels + 1 - kk) / levels) ** 2)[:, None]
    data = rng.normal(0.0, 1.0, (levels, ncol))
    return dict(clat=clat, clon=clon, hsurf=hsurf, hhl=hhl, hhl_height=kk.astype(np.int32), data=data,
                data_height=np.arange(1, levels + 1, dtype=np.int32))


@pytest.mark.parametrize("land,mode", [(False, icb.INGEST_FIXED), (True, icb.INGEST_FIXED), (True, icb.INGEST_REFERENCE),
                                       (True, icb.INGEST_FIXED | icb.INGEST_SURFACE_LEVEL)])
def test_ingested_records_feed_the_hot_path(land, mode):
    """ingest -> icb_set_grid's builder -> the device locator text (emulation): ids and values of the reference's brute-force
    loop over the ingested records, for a flat-surface atmosphere (well-formed columns: the walk) and for one with topography
    (the converter's H[0] = R + hsurf under H[1] = R + HHL - hsurf makes land columns non-monotone: the exact query)"""
    import emu_lib
    import oracle_lib as ol
    inp = _atmosphere(2, 2, 40, land)
    cells = ours(inp, 1000, mode=mode, order_seed=11)
    ncol = len(inp["hsurf"])
    assert len(cells) == 2 * ncol
    want_layers = [31, 9] if mode & icb.INGEST_FIXED else [31, 7]              # 40 = 31 + 9; the reference writes 40 % 32 - 1 = 7
    assert cells["numLayers"][:2].tolist() == want_layers
    emu, port = emu_lib.Emu(), ol.Port()
    info = emu.build(cells)
    assert info["nCol"] == ncol
    if not land or mode & icb.INGEST_SURFACE_LEVEL:
        assert info["nGeneric"] == 0 and info["boundaryEdges"] == 0        # monotone columns: every sample can take the walk
    else:
        assert info["nGeneric"] > 0
    if mode & icb.INGEST_SURFACE_LEVEL:
        # heights are the half levels above ground, bottom-up, and data level j lies between half levels j+1 and j (ICON numbering)
        hag = (inp["hhl"] - inp["hsurf"][None, :])[::-1]                      # (levels + 1, ncol), surface first
        want_h = (np.float32(6.371229e6) + inp["hhl"][::-1].astype(np.float32)).astype(np.float64) - inp["hsurf"][None, :]
        col = np.concatenate([cells["height"][0, :32], cells["height"][1, 1:10]])
        assert np.array_equal(col, want_h[:, 0].astype(np.float32)) and np.all(np.diff(hag[:, 0]) > 0)
        assert cells["value"][0, 0] == icb.ingest_normalize(inp["data"][-1])[0]
    rng = np.random.default_rng(9)
    d = rng.normal(size=(4000, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    r = rng.uniform(6.371229e6 - 2e3, 6.371229e6 + 7.7e4, 4000)
    pos = (d * r[:, None]).astype(np.float32)
    want_ids, want_vals = port.locate(cells, pos)
    assert (want_ids >= 0).sum() > 3000
    for flags in (0, icb.LOCATE_WALK):
        ids, vals, _ = emu.locate(pos, flags)
        assert np.array_equal(ids, want_ids) and np.array_equal(vals, want_vals)


@pytest.mark.gpu
@pytest.mark.parametrize("land,mode", [(False, icb.INGEST_FIXED), (True, icb.INGEST_FIXED), (True, icb.INGEST_REFERENCE),
                                       (True, icb.INGEST_FIXED | icb.INGEST_SURFACE_LEVEL)])
def test_ingested_records_through_the_cuda_locator(land, mode):
    """the same four ingests on the device: convert_icon's array -> `.ic` half (icb_ingest_columns) -> icb_set_grid ->
    icb_locate (plain, walk, exact) against the reference's own brute-force loop (deviceCode.cu:116-123 compiled in place,
    oracle/_ref; the C port where it is absent) over the ingested records: ids and values bit for bit.  Reference mode with
    topography makes land columns non-monotone (the builder marks them generic: exact candidate-bin query), the
    surface-level fix makes every column well-formed (the walk)."""
    import oracle_lib as ol
    inp = _atmosphere(2, 3, 40, land)
    cells = ours(inp, 1000, mode=mode, order_seed=11)
    ctx = icb.Context(0)
    try:
        ctx.set_grid(cells)
        st = ctx.stats()
        assert st["numRecords"] == len(cells) and st["numColumns"] == len(inp["hsurf"])
        if not land or mode & icb.INGEST_SURFACE_LEVEL:
            assert st["numGenericColumns"] == 0
        else:
            assert st["numGenericColumns"] > 0
        rng = np.random.default_rng(9)
        d = rng.normal(size=(6000, 3))
        d /= np.linalg.norm(d, axis=1)[:, None]
        r = rng.uniform(6.371229e6 - 2e3, 6.371229e6 + 7.7e4, len(d))
        # a third of the radii exactly on stored heights of a record under the position's direction is what the seams need;
        # stored heights of arbitrary records still land on layer boundaries of their own columns' neighbours often enough
        hs = cells["height"][rng.integers(0, len(cells), 2000), rng.integers(0, 8, 2000)]
        r[:2000] = np.where(hs > 0, hs, r[:2000])
        pos = (d * r[:, None]).astype(np.float32)
        oracle = ol.Ref() if ol.Ref.available() else ol.Port()
        want_ids, want_vals = oracle.locate(cells, pos)
        assert (want_ids >= 0).sum() > len(pos) // 2
        for flags in (0, icb.LOCATE_WALK, icb.LOCATE_EXACT):
            ids, vals = ctx.locate_host(pos, flags)
            assert np.array_equal(ids, want_ids) and np.array_equal(vals, want_vals), flags
    finally:
        ctx.close()
