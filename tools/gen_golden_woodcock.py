#!/usr/bin/env python
"""tests/golden/woodcock_frames_r2b1_l40.npz: single frames of the REFERENCE's own integrator (woodcockTrackingAE through
Pipeline::launch, oracle/_ref) for accumID 0 and 3 on the golden grid/camera -- the vectors the CUDA Woodcock integrator
and the C port are compared with bit for bit.  Run in the build container (needs oracle/_ref):
    python tools/gen_golden_woodcock.py"""
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
out = {}
for acc in (0, 3):
    sc = ol.Scene(cells, blocks["lut"], camera=cam, unit_distance=float(w["unit_distance"]), r_inner=b["sphLo"][0],
                  r_outer=b["sphHi"][0], jitter=1, accum_id=acc)
    fr = ref.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
    out[f"fb_{acc}"] = fr["fb"]
    out[f"sample_{acc}"] = fr["accum"]          # accumBuffer * (accumID + 1): this frame's collision per pixel
    print("accumID", acc, "covered", float((fr["accum"][:, 3] > 0).mean()))
np.savez_compressed(os.path.join(GOLD, "woodcock_frames_r2b1_l40.npz"), W=W, H=H, unit_distance=float(w["unit_distance"]),
                    volLo=b["volLo"], volHi=b["volHi"], **out)
