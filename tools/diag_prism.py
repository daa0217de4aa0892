import sys, os, numpy as np, torch
sys.path.insert(0,'tests'); sys.path.insert(0,'icon-ray-tracing_b200')
import icb, oracle_lib as ol
port = ol.Port()
cells = icb.gen_grid(2, 3, 90)
ctx = icb.Context(0)
ctx.set_grid(cells); ctx.set_sampler("prism")
prism = port.prism_build(cells)
b = ctx.bounds()
W, H = 64, 40
cam = icb.camera_orient(np.array([0.2e7, 0.9e7, 0.75e7], np.float32), (0.3e7, 0, 0), (0, 0, 1), W, H)
lut = icb.sharp_lut()
ctx.set_transfunc(lut, value_range=(0.1, 0.9), opacity_scale=0.7)
ctx.set_camera(cam)
ctx.set_params(unit_distance=20000.0, dt=9000.0, ambient_color=(0.9, 0.8, 1.0), ambient_radiance=1.1)
sc = ol.Scene(cells, lut, tf_range=(0.1, 0.9), opacity_scale=0.7, camera=cam, unit_distance=20000.0, dt=9000.0,
              r_inner=b["sphLo"][0], r_outer=b["sphHi"][0], ambient_color=(0.9, 0.8, 1.0), ambient_radiance=1.1, prism=prism)
want = port.march_frame(sc, W, H)
for skip in (True, False):
    ctx.set_skip_ahead(skip)
    pp = torch.zeros(W * H, 4, dtype=torch.int32, device="cuda")
    fb = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    accum = torch.zeros(W * H, 4, dtype=torch.float32, device="cuda")
    ctx.render_trace(fb.data_ptr(), accum.data_ptr(), pp.data_ptr(), W, H)
    torch.cuda.synchronize()
    g = pp.cpu().numpy(); a = accum.cpu().numpy()
    bad = np.where((g != want["perPixel"]).any(1))[0]
    print("skip", skip, "env", os.environ.get("ICB_PRISM_NO_EMPTY_SKIP"), "mismatching pixels", len(bad), "max accum diff", np.abs(a - want["accum"]).max())
    for p in bad[:6]:
        x, y = p % W, p // W
        print("  px", x, y, "gpu", g[p][:2], "port", want["perPixel"][p][:2], "rgba gpu", a[p], "port", want["accum"][p])
        ids, n, rgba = ctx.trace_pixel(x, y, W, H)
        wids, wn, wrgba = port.march_trace(sc, W, H, x, y)
        print("     trace_pixel n", n, "port n", wn, "ids equal", np.array_equal(ids, wids), rgba, wrgba)
