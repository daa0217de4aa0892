#!/usr/bin/env python
"""Generate tests/golden/*.npz + the small .ic fixture from the REFERENCE's own code
(oracle/_ref/libicon_ref.so, built from /root/reference in place by `make -C oracle ref`).

Run in the build container only (needs /root/reference at build time of the .so):
    python tools/gen_golden.py
The fixtures are what pins the C restatement (oracle/icon_oracle.c) and the CUDA path on machines
where /root/reference does not exist.  Seeds are fixed; the synthetic grid comes from icb_gen_grid.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "icon-ray-tracing_b200"))
import icb  # noqa: E402
import oracle_lib as ol  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def unit(lat, lon):
    return np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], -1)


def adversarial_points(cells, rng, n):
    """points within a few metres of side planes, at corners, at exact layer / record boundaries"""
    lat = cells["lat"].astype(np.float64)
    lon = cells["lon"].astype(np.float64)
    ci = rng.integers(0, len(cells), n)
    k = rng.integers(0, 3, n)
    A = unit(lat[ci, k], lon[ci, k])
    B = unit(lat[ci, (k + 1) % 3], lon[ci, (k + 1) % 3])
    s = rng.uniform(-0.02, 1.02, n)[:, None]
    s[: n // 8] = rng.choice([0.0, 1.0], (n // 8, 1))          # exactly at corners
    Pd = A * (1 - s) + B * s
    Pd /= np.linalg.norm(Pd, axis=1)[:, None]
    nrm = np.cross(A, B)
    nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    d = rng.uniform(-6, 6, n)
    d[n // 8: n // 4] = 0.0
    # radii: half random, half exactly on a stored height (layer boundaries and record seams)
    h = cells["height"]
    nl = cells["numLayers"]
    j = (rng.uniform(0, 1, n) * (nl[ci] + 1)).astype(int)
    r = np.where(rng.uniform(0, 1, n) < 0.5, h[ci, j].astype(np.float64),
                 rng.uniform(h[ci, 0], h[ci, nl[ci]]))
    P = (Pd * r[:, None] + nrm * d[:, None]).astype(np.float32)
    # re-scale so that the fp32 length equals the stored height exactly where possible
    return P


def main():
    assert ol.Ref.available(), "build oracle/_ref first: make -C oracle ref"
    ref = ol.Ref()
    os.makedirs(GOLD, exist_ok=True)
    rng = np.random.default_rng(20261104)

    # ---- fixture grid: R2B1 x 40 levels (2 records per column, 640 records, 182 kB) ----------
    cells = icb.gen_grid(2, 1, 40)
    cells.tofile(os.path.join(GOLD, "r2b1_l40.ic"))
    b = ref.grid_bounds(cells)

    # ---- building blocks ------------------------------------------------------------------------
    sph = np.stack([rng.uniform(6.3e6, 7.2e6, 512), rng.uniform(-np.pi / 2, np.pi / 2, 512), rng.uniform(-np.pi, np.pi, 512)], -1).astype(np.float32)
    cart = ref.to_cartesian(sph)
    planes = ref.cell_planes(cells)
    lut = ref.resample_lut(ol.DEFAULT_LUT5, 300)
    v = np.concatenate([rng.uniform(-0.2, 1.2, 500), [0.0, 1.0, 0.5, -0.0, 1.0 / 300, 299.0 / 300]]).astype(np.float32)
    cls = ref.post_classify(lut, 0.0, 1.0, 1.0, v)
    cls2 = ref.post_classify(lut, 0.1, 0.9, 2.5, v)
    lcg = np.stack([ref.lcg(s0, s1, 16) for s0, s1 in [(0, 0), (1, 2), (511, 77), (262144 * 3 + 5, 400), (2**32 - 1, 12345)]])
    nl4 = int(cells["numLayers"][4])
    fh_r = np.concatenate([cells["height"][4, : nl4 + 1],
                           rng.uniform(cells["height"][4, 0] - 10, cells["height"][4, nl4] + 10, 200).astype(np.float32)])
    fh = ref.find_height(cells[4:5], fh_r)
    org = rng.normal(size=(256, 3)).astype(np.float32) * np.float32(2e7)
    dirs = -org / np.linalg.norm(org, axis=1)[:, None] + rng.normal(size=(256, 3)).astype(np.float32) * np.float32(0.25)
    dirs = (dirs / np.linalg.norm(dirs, axis=1)[:, None]).astype(np.float32)
    sh_hit, sh_tn, sh_tf = ref.intersect_sphere(org, dirs, float(b["sphHi"][0]))
    segs = np.full((256, 4), np.nan, np.float32)
    nseg = np.zeros(256, np.int32)
    for i in range(256):
        s = ref.shell_segments(org[i], dirs[i], 0.0, 1e10, float(b["sphLo"][0]), float(b["sphHi"][0]))
        nseg[i] = len(s)
        segs[i, : 2 * len(s)] = s.reshape(-1)
    rgba = rng.uniform(-0.1, 1.1, (512, 4)).astype(np.float32)
    packed = ref.srgb_pack(rgba)
    cam_va = ref.camera_view_all(b["volLo"], b["volHi"], 48 / 32, 48, 32)
    cam_or = ref.camera_orient([1.2e7, 3e6, 2.0e7], [0, 0, 0], [0, 1, 0], ol.FOVY_DEFAULT * 0.6, 1.5, 48, 32)
    np.savez_compressed(
        os.path.join(GOLD, "blocks.npz"), sph=sph, cart=cart, planes=planes, lut=lut, v=v, cls=cls, cls2=cls2, lcg=lcg,
        fh_r=fh_r, fh=fh, org=org, dirs=dirs, sh_hit=sh_hit, sh_tn=sh_tn, sh_tf=sh_tf, segs=segs, nseg=nseg,
        rgba=rgba, packed=packed, **{f"b_{k}": x for k, x in b.items()},
        **{f"camva_{k}": x for k, x in cam_va.items()}, **{f"camor_{k}": x for k, x in cam_or.items()})

    # ---- locate: random + adversarial positions -> record ids / values ---------------------------------
    n = 6000
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    r = rng.uniform(b["sphLo"][0] - 2e4, b["sphHi"][0] + 2e4, n)
    pos = np.concatenate([(d * r[:, None]).astype(np.float32), adversarial_points(cells, rng, 6000)])
    ids, vals = ref.locate(cells, pos)
    np.savez_compressed(os.path.join(GOLD, "locate_r2b1_l40.npz"), pos=pos, ids=ids, vals=vals)
    print("locate: hit fraction", (ids >= 0).mean())

    # ---- frames: A.3 marcher (composed from reference functions) ------------------------------------------
    W, H = 48, 32
    frames = {}
    eye = np.array([0.9e7, 0.35e7, 1.05e7], np.float32)          # close enough for the globe to fill the frame
    cam = ref.camera_orient(eye, [0, 0, 0], [0, 1, 0], ol.FOVY_DEFAULT, W / H, W, H)
    inside_eye = np.array([0, 0, float(b["sphLo"][0]) + 3.0e5], np.float32)   # camera inside the shell
    cam_in = ref.camera_orient(inside_eye, [5e6, 1e6, float(b["sphLo"][0]) + 3.2e5], [0, 0, 1], ol.FOVY_DEFAULT, W / H, W, H)
    for name, l, jit, cm, acc in (("smooth", lut, 0, cam, 0), ("sharp_jit", icb.sharp_lut(), 1, cam, 0),
                                  ("smooth_inside", lut, 1, cam_in, 0), ("smooth_acc3", lut, 1, cam, 3)):
        sc = ol.Scene(cells, l, camera=cm, unit_distance=16000.0, dt=4000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0],
                      jitter=jit, accum_id=acc)
        prev = None
        if acc:
            prev = rng.uniform(0, 1, (W * H, 4)).astype(np.float32)
        out = ref.march_frame(sc, W, H, accum=None if prev is None else prev.copy())
        frames[name + "_fb"] = out["fb"]
        frames[name + "_accum"] = out["accum"]
        frames[name + "_pp"] = out["perPixel"]
        if prev is not None:
            frames[name + "_prev"] = prev
        for k, x in cm.items():
            frames[f"{name}_cam_{k}"] = x
        print(name, "steps", out["steps"], "hits", out["hits"], "sec", round(out["seconds"], 2))
    # full id traces of a few pixels
    sc = ol.Scene(cells, icb.sharp_lut(), camera=cam, unit_distance=16000.0, dt=4000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0], jitter=1)
    for (x, y) in [(24, 16), (5, 16), (40, 9), (23, 30)]:
        ids, nst, c = ref.march_trace(sc, W, H, x, y)
        frames[f"trace_{x}_{y}_ids"] = ids
        frames[f"trace_{x}_{y}_rgba"] = c
    np.savez_compressed(os.path.join(GOLD, "frames_r2b1_l40.npz"), W=W, H=H, dt=4000.0, **frames)

    # ---- the reference's own integrator (Woodcock) for the statistical cross-check -----------------------------
    nfr = 16
    acc = np.zeros((W * H, 4), np.float64)
    acc2 = np.zeros((W * H, 4), np.float64)
    sec = 0.0
    for f in range(nfr):
        sc = ol.Scene(cells, lut, camera=cam, unit_distance=40000.0, r_inner=b["sphLo"][0], r_outer=b["sphHi"][0], jitter=1, accum_id=f)
        wd = ref.woodcock_frame(sc, W, H, b["volLo"], b["volHi"])
        acc += wd["accum"]
        acc2 += wd["accum"].astype(np.float64) ** 2
        sec += wd["seconds"]
    np.savez_compressed(os.path.join(GOLD, "woodcock_r2b1_l40.npz"), mean=(acc / nfr).astype(np.float32),
                        meansq=(acc2 / nfr).astype(np.float32), n_frames=nfr, unit_distance=40000.0,
                        **{f"cam_{k}": x for k, x in cam.items()})
    print("woodcock", nfr, "frames:", round(sec, 2), "s; mean alpha", (acc / nfr)[:, 3].mean())
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
