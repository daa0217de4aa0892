"""CPU: the C-ABI library loads and exports every symbol include/icon_b200.h declares; the host mirror
(camera, LUT resampling, .ic loader, synthetic grid) agrees with the oracle / golden vectors; no CPU fallback."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import icb
import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "icon_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(icb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 30
    L = icb.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert L.icb_version().decode().startswith("icon-b200")


def test_struct_sizes_match_header():
    assert C.sizeof(icb.Bounds) == 15 * 4
    assert C.sizeof(icb.Stats) == 9 * 8 + 7 * 8 + 8      # 8 int64 + double + 7 uint64 + int32 + float


@pytest.mark.skipif(_have_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback_without_device():
    with pytest.raises(icb.IcbError) as e:
        icb.Context(0)
    assert "no CPU path" in str(e.value) or "-6" in str(e.value)


def test_argument_errors_do_not_crash():
    L = icb.lib()
    assert L.icb_create(0, None) == -1
    assert L.icb_destroy(None) == -1
    assert L.icb_stats(None, None) == -1
    assert L.icb_set_grid(None, None, 0) == -1
    assert L.icb_gen_num_records(0, 0, 10) == -1


def test_generator_counts_and_format():
    assert icb.lib().icb_gen_num_columns(2, 4) == 20480         # R2B4 (BASELINE configs[0])
    assert icb.lib().icb_gen_num_columns(2, 6) == 327680        # R2B6
    assert icb.lib().icb_gen_num_records(2, 8, 90) == 15728640  # R2B8 x 90 -> 3 records per column
    cells = icb.gen_grid(2, 1, 40)
    gold = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    assert cells.tobytes() == gold.tobytes()                    # the fixture is reproducible
    c90 = icb.gen_grid(2, 0, 90)
    assert len(c90) == 240 and c90["numLayers"].reshape(-1, 3).tolist()[0] == [31, 31, 28]
    # stacked records share lat/lon and meet exactly at the seam
    c3 = c90.reshape(-1, 3)
    assert np.array_equal(c3["lat"][:, 0], c3["lat"][:, 1]) and np.array_equal(c3["lon"][:, 0], c3["lon"][:, 2])
    assert np.array_equal(c3["height"][:, 0, 31], c3["height"][:, 1, 0])
    assert np.array_equal(c3["height"][:, 1, 31], c3["height"][:, 2, 0])
    assert (np.diff(c90["height"][0, :32]) > 0).all()
    assert 0.0 <= c90["value"].min() and c90["value"].max() <= 1.0
    # corners counter-clockwise seen from outside (fixture hostCode.cu:771-777)
    lat, lon = c90["lat"].astype(np.float64), c90["lon"].astype(np.float64)
    v = np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], -1)
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    assert ((n * v.sum(1)).sum(-1) > 0).all()
    # time steps shift the scalar field only
    t5 = icb.gen_grid(2, 0, 90, timestep=5)
    assert np.array_equal(t5["height"], c90["height"]) and not np.array_equal(t5["value"], c90["value"])


def test_host_mirror_camera_and_lut_match_reference_vectors():
    g = np.load(os.path.join(GOLD, "blocks.npz"))
    cam = icb.camera_view_all(g["b_volLo"], g["b_volHi"], 48, 32, aspect=48 / 32)
    for k in cam:
        assert np.array_equal(cam[k], g["camva_" + k]), k
    cam = icb.camera_orient([1.2e7, 3e6, 2.0e7], [0, 0, 0], [0, 1, 0], 48, 32, aspect=1.5, fovy=icb.FOVY_DEFAULT * np.float32(0.6))
    for k in cam:
        assert np.array_equal(cam[k], g["camor_" + k]), k
    assert np.array_equal(icb.smooth_lut(), g["lut"])
    s = icb.sharp_lut()
    assert (s[:150, 3] == 0).all() and (s[150:, 3] == 1).all() and np.array_equal(s[:, :3], g["lut"][:, :3])


def test_ic_loader_filters_like_the_reference(tmp_path):
    """hostCode.cu:717-758: --num-cells truncation first, then every record with a corner outside the window goes;
    survivors keep file order (record id == index after filtering)."""
    path = os.path.join(GOLD, "r2b1_l40.ic")
    cells = np.fromfile(path, icb.CELL_DTYPE)
    L = icb.lib()
    L.icb_load_ic.restype = C.c_int64
    L.icb_load_ic.argtypes = [C.c_char_p, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int64]
    out = np.zeros(len(cells), icb.CELL_DTYPE)
    n = L.icb_load_ic(path.encode(), -1, -90, 90, -180, 180, out.ctypes.data, len(out))
    assert n == 640 and out.tobytes() == cells.tobytes()
    n = L.icb_load_ic(path.encode(), 100, -90, 90, -180, 180, out.ctypes.data, len(out))
    assert n == 100
    n = L.icb_load_ic(path.encode(), -1, -30, 60, -90, 120, out.ctypes.data, len(out))
    py = icb.read_ic(path, lat_range=(-30, 60), lon_range=(-90, 120))
    assert 0 < n < 640 and n == len(py) and out[:n].tobytes() == py.tobytes()
    assert L.icb_load_ic(b"/nonexistent.ic", -1, -90, 90, -180, 180, None, 0) < 0


def test_oracle_is_not_reachable_from_the_product():
    """the product library must not link or open anything under oracle/ (no CPU fallback)"""
    import subprocess
    so = icb.LIB_PATH
    deps = subprocess.run(["ldd", so], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "icon_ref" not in deps
    strings = subprocess.run(["strings", so], capture_output=True, text=True).stdout
    assert "libicon_oracle" not in strings and "libicon_ref" not in strings
    for fn in ("icb.py", "sortfirst.py"):
        p = os.path.join(ROOT, "icon-ray-tracing_b200", fn)
        if os.path.exists(p):
            assert "import oracle_lib" not in open(p).read()


def test_gen_values_is_the_generators_scalar_field_for_any_slab():
    """icb_gen_values (a time-series rank's slab of every timestep) against the value arrays of icb_gen_grid"""
    n, k, L = 2, 3, 40
    grids = {t: icb.gen_grid(n, k, L, timestep=t) for t in (0, 1, 2, 7)}
    nrec = len(grids[0])
    for first, count in ((0, nrec), (1, 7), (nrec - 5, 5), (333, 1001)):
        out = np.full((3, count, 32), np.nan, np.float32)
        icb.gen_values(n, k, L, 0, 3, first, count, out.ctypes.data)
        for t in range(3):
            assert np.array_equal(out[t], grids[t]["value"][first:first + count]), (first, count, t)
    one = np.empty((1, 64, 32), np.float32)
    icb.gen_values(n, k, L, 7, 1, 100, 64, one.ctypes.data)
    assert np.array_equal(one[0], grids[7]["value"][100:164])
    with pytest.raises(icb.IcbError):
        icb.gen_values(n, k, L, 0, 1, nrec - 1, 2, one.ctypes.data)
