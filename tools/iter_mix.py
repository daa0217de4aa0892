#!/usr/bin/env python
"""Developer tool: render bench frames of a workload with the TRACE kernel and print the iteration mix of the
march loop (icb_debug_counters).  GPU only.   python tools/iter_mix.py [C3|C2|C1] [frames]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "icon-ray-tracing_b200"))
import torch

import bench
import icb

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 2
n, k, levels, W, H, tf, desc = bench.WORKLOADS[wl]
ctx = icb.Context(0)
ctx.set_grid(icb.gen_grid(n, k, levels))
b = ctx.bounds()
ctx.set_transfunc(icb.sharp_lut() if tf == "sharp" else icb.smooth_lut())
ctx.set_params(jitter=True)
cams = bench.orbit_cameras(icb, b, W, H, frames)
fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
pp = torch.zeros(W * H, 4, dtype=torch.int32, device="cuda")
for i in range(frames):
    ctx.set_camera(cams[i])
    ctx.render_trace(fb.data_ptr(), accum.data_ptr(), pp.data_ptr(), W, H)
    torch.cuda.synchronize()
    st, d = ctx.stats(), ctx.debug_counters()
    li = max(d["lane_iters"], 1)
    print(f"frame {i}: {st['lastKernelMs']:.2f} ms (trace kernel), samples {st['steps']/1e6:.1f}M rays {st['raysMarched']/1e6:.2f}M exact {st['exactQueries']/1e6:.2f}M")
    print(f"  warp iters {d['warp_iters']/1e6:.2f}M, lane iters {li/1e6:.1f}M ({li/max(d['warp_iters'],1):.1f} lanes/iter), samples/lane-iter {st['steps']/li:.2f}")
    print("  mix: " + ", ".join(f"{k} {d[k]/li*100:.1f}%" for k in ("fast", "layer", "fetch", "retry", "no_skip")))
    steps = pp[:, 0].to(torch.float32)
    print(f"  per-pixel samples: mean {steps.mean().item():.1f} max {steps.max().item():.0f}")
