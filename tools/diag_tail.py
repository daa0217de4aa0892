"""critical path of a C3 frame: per-tile costs (lane iterations) and the kernel time of the heaviest tiles rendered alone"""
import sys, os, numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests'); sys.path.insert(0, 'icon-ray-tracing_b200')
import icb, bench
n, k, levels, W, H, tf, desc = bench.WORKLOADS["C3"]
cells = icb.gen_grid(n, k, levels)
ctx = icb.Context(0); ctx.set_grid(cells); ctx.set_transfunc(icb.sharp_lut()); ctx.set_params(jitter=True)
b = ctx.bounds()
cams = bench.orbit_cameras(icb, b, W, H, 4)
npx = W * H
fb = torch.zeros(npx, dtype=torch.int32, device="cuda"); accum = torch.zeros(npx, 4, dtype=torch.float32, device="cuda")
tx, ty = (W + 15) // 16, (H + 15) // 16
for i in range(3):
    ctx.set_camera(cams[i]); ctx.render(fb.data_ptr(), accum.data_ptr(), W, H); torch.cuda.synchronize()
print("full frame ms", ctx.stats()["lastKernelMs"], "steps", ctx.stats()["steps"])
costs = torch.zeros(tx * ty, dtype=torch.int32, device="cuda")
ctx.tile_costs(costs.data_ptr(), tx * ty); torch.cuda.synchronize()
c = costs.cpu().numpy().astype(np.int64)
print("tile cost: sum %d mean %.0f p50 %d p99 %d max %d" % (c.sum(), c.mean(), np.percentile(c, 50), np.percentile(c, 99), c.max()))
pp = torch.zeros(npx, 4, dtype=torch.int32, device="cuda")
top = np.argsort(-c)[:6]
t_fb = torch.zeros(256 * 4096, dtype=torch.int32, device="cuda")
for t in list(top) + [int(np.argsort(c)[len(c) // 2])]:
    ctx.set_tiles(np.array([t], np.int32))
    for _ in range(2):
        ctx.render_tiles(t_fb.data_ptr(), 0, W, H); torch.cuda.synchronize()
    st = ctx.stats()
    print("tile", t, "(x,y)=", (t % tx) * 16, (t // tx) * 16, "cost", c[t], "alone ms", round(st["lastKernelMs"], 3), "steps", st["steps"], "slow", st["slowSteps"], "exact", st["exactQueries"])
# groups: heaviest 1%, random 1%
for name, sel in (("heaviest 324 tiles", np.argsort(-c)[:324]), ("every 100th tile", np.arange(0, len(c), 100))):
    ctx.set_tiles(sel.astype(np.int32))
    for _ in range(2):
        ctx.render_tiles(t_fb.data_ptr(), 0, W, H); torch.cuda.synchronize()
    st = ctx.stats()
    print(name, "ms", round(st["lastKernelMs"], 3), "steps", st["steps"], "cost", c[sel].sum())
ctx.set_tiles(np.zeros(0, np.int32))
ctx.render_trace(fb.data_ptr(), accum.data_ptr(), pp.data_ptr(), W, H); torch.cuda.synchronize()
p = pp.cpu().numpy()
print("per-pixel steps: max", p[:, 0].max(), "p99.9", np.percentile(p[:, 0], 99.9), "mean", p[:, 0].mean())
