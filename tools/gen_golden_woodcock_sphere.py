#!/usr/bin/env python
"""tests/golden/woodcock_sphere_r2b1_l40.npz: single frames of the REFERENCE's ray-gen woodcockTrackingWithAccel
(deviceCode.cu:281-341) in SPHERE_ACCEL_MODE -- the accel mode the reference starts in (hostCode.cu:75, :915): sdda
(ShellAccel.h:83-199) over the (rad, lat, lon) shell accel, through Pipeline::launch (oracle/_ref) -- on the golden
grid/camera for two transfer functions and accumID 0 and 3.  The shell's value ranges and max opacities come from the
reference's own builder kernels (hostCode.cu:216-225 initGrid, :298-334 buildShell_ICON, :362-397 computeMaxOpacities on
host shims) and are stored alongside.  Run in the build container (needs oracle/_ref):
    python tools/gen_golden_woodcock_sphere.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "icon-ray-tracing_b200"))
import icb  # noqa: E402
import oracle_lib as ol  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
assert ol.Ref.available(), "build oracle/_ref first: make -C oracle ref"
ref = ol.Ref()
cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
w = np.load(os.path.join(GOLD, "woodcock_r2b1_l40.npz"))
g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
blocks = np.load(os.path.join(GOLD, "blocks.npz"))
W, H = int(g["W"]), int(g["H"])
cam = {k: w[f"cam_{k}"] for k in ("org", "d00", "du", "dv")}
b = ref.grid_bounds(cells)
DIMS = (1, 96, 160)
ranges = ref.shell_ranges(cells, DIMS, b["sphLo"], b["sphHi"])
out = {"ranges": ranges}
for name, lut in (("smooth", blocks["lut"]), ("sharp", icb.sharp_lut())):
    mo = ref.max_opacities(ranges, lut, shell=True)
    out[f"lut_{name}"] = np.asarray(lut, np.float32)
    out[f"maxop_{name}"] = mo
    for acc in (0, 3):
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=float(w["unit_distance"]), r_inner=b["sphLo"][0],
                      r_outer=b["sphHi"][0], jitter=1, accum_id=acc)
        fr = ref.woodcock_sphere_frame(sc, W, H, b["volLo"], b["volHi"], DIMS, b["sphLo"], b["sphHi"], mo)
        out[f"fb_{name}_{acc}"] = fr["fb"]
        out[f"sample_{name}_{acc}"] = fr["accum"]
        print(name, "accumID", acc, "covered", float((fr["accum"][:, 3] > 0).mean()), "majorant>0", float((mo > 0).mean()))
np.savez_compressed(os.path.join(GOLD, "woodcock_sphere_r2b1_l40.npz"), W=W, H=H, dims=np.asarray(DIMS, np.int32),
                    unit_distance=float(w["unit_distance"]), volLo=b["volLo"], volHi=b["volHi"], sphLo=b["sphLo"], sphHi=b["sphHi"], **out)
