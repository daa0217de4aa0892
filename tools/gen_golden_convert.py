"""Golden vectors for the converter's array -> .ic half: runs the REFERENCE converter
(tools/convert_icon/convert_icon.cpp, compiled where it lies against oracle/stub_nc/netcdf.h -> oracle/_ref/convert_icon_ref)
on small synthetic inputs and stores inputs + the records it wrote in tests/golden/convert_icon.npz.
Entries of H[] / value[] beyond numLayers are uninitialised stack in the reference's output; they are zeroed in the fixture.
Run in the build container (needs /root/reference): python tools/gen_golden_convert.py"""
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "convert_icon_ref")
CELL = np.dtype([("lat", "<f4", 3), ("lon", "<f4", 3), ("numLayers", "<i4"), ("height", "<f4", 32), ("value", "<f4", 32)])

# (name, cells, data levels, maxLayers handed to the converter)
CASES = [("cap5", 37, 12, 5), ("l20", 37, 20, 1000), ("l31", 19, 31, 1000), ("l32", 11, 32, 1000), ("l62", 9, 62, 62),
         ("l90", 23, 90, 1000), ("cap12of40", 9, 40, 12)]


def write_nc(path, dims, variables):
    with open(path, "wb") as f:
        for name, n in dims.items():
            f.write(struct.pack("<BI", 0, len(name)) + name.encode() + struct.pack("<Q", n))
        for name, arr in variables.items():
            a = np.ascontiguousarray(arr, np.float64).reshape(-1)
            f.write(struct.pack("<BI", 1, len(name)) + name.encode() + struct.pack("<Q", a.size) + a.tobytes())


def make_inputs(cells, levels, seed):
    rng = np.random.default_rng(seed)
    clat = rng.uniform(-1.5, 1.5, (cells, 3))
    clon = rng.uniform(-3.1, 3.1, (cells, 3))
    hsurf = np.where(rng.random(cells) < 0.3, 0.0, rng.uniform(0.0, 3000.0, cells))
    k = np.arange(1, levels + 2)                                   # ICON half levels, 1 = model top
    frac = ((levels + 1 - k) / levels) ** 2
    hhl = hsurf[None, :] + (75000.0 - hsurf[None, :]) * frac[:, None] + rng.uniform(0, 1e-3, (levels + 1, cells))
    data = rng.normal(250.0, 40.0, (levels, cells)) * (1.0 + np.arange(levels)[:, None])
    return dict(clat=clat, clon=clon, hsurf=hsurf, hhl=hhl, hhl_height=k.astype(np.int32), data=data,
                data_height=np.arange(1, levels + 1, dtype=np.int32))


def run_reference(inp, max_layers, workdir):
    cells = inp["hsurf"].shape[0]
    write_nc(os.path.join(workdir, "hgrid.nc"), {"cell": cells, "vertex": 3 * cells},
             {"clon_vertices": inp["clon"], "clat_vertices": inp["clat"]})
    write_nc(os.path.join(workdir, "hsurf.nc"), {}, {"HSURF": inp["hsurf"]})
    rng = np.random.default_rng(99)
    hhl_files, data_files = [], []
    for i in rng.permutation(len(inp["hhl_height"])):                 # file order is NOT level order: the converter sorts
        p = os.path.join(workdir, f"hhl_{i}.nc")
        write_nc(p, {}, {"height": [float(inp["hhl_height"][i])], "HHL": inp["hhl"][i]})
        hhl_files.append(p)
    for i in rng.permutation(len(inp["data_height"])):
        p = os.path.join(workdir, f"data_{i}.nc")
        write_nc(p, {"ncells": cells}, {"height": [float(inp["data_height"][i])], "pres": inp["data"][i]})
        data_files.append(p)
    out = os.path.join(workdir, "out")
    cmd = [REF_BIN, "-hgrid", os.path.join(workdir, "hgrid.nc"), "-hsurf", os.path.join(workdir, "hsurf.nc"),
           "-hhl", *hhl_files, "-data", *data_files, "-o", out]
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, env=dict(os.environ, REF_CONVERT_MAX_LAYERS=str(max_layers)))
    rec = np.fromfile(out + ".ic", CELL)
    for r in rec:                                                       # uninitialised stack beyond numLayers -> zero
        nl = max(int(r["numLayers"]), 0)
        r["height"][nl + 1:] = 0
        r["value"][nl:] = 0
    return rec


def main():
    if not os.path.exists(REF_BIN):
        sys.exit("oracle/_ref/convert_icon_ref missing: make -C oracle ref (needs /root/reference)")
    out = {}
    for idx, (name, cells, levels, cap) in enumerate(CASES):
        inp = make_inputs(cells, levels, 1000 + idx)
        with tempfile.TemporaryDirectory() as d:
            rec = run_reference(inp, cap, d)
        for k_, v in inp.items():
            out[f"{name}_{k_}"] = v
        out[f"{name}_cap"] = np.int32(cap)
        out[f"{name}_records"] = rec.view(np.uint8).reshape(-1, 284)
        print(name, "cells", cells, "levels", levels, "cap", cap, "->", len(rec), "records, numLayers of a column:",
              rec["numLayers"][:len(rec) // cells].tolist())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "convert_icon.npz"), **out)


if __name__ == "__main__":
    main()
