#!/usr/bin/env python
"""tests/golden/prism_*.npz: vectors of the REFERENCE's wedge sampler -- intersectWedgeEXT (UElems.h:214-311, Newton inverse
map + shape functions) compiled in place (oracle/_ref) -- on the prisms our locator finds, with the six wedge vertices laid
out as buildCuBQLAccel does (hostCode.cu:562-597).  The prism sampler (icb_set_sampler(ICB_SAMPLER_PRISM)) and its C
restatement (oracle/icon_oracle.c orc_prism_*) compute the same quantity in closed form and are checked against these.
Run in the build container (needs oracle/_ref):
    python tools/gen_golden_prism.py"""
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
ref, port = ol.Ref(), ol.Port()


def positions(cells, m, rng):
    """points inside random prisms (uniform barycentric, uniform radius), a share of them pushed towards edges,
    corners and layer interfaces"""
    rid = rng.integers(len(cells), size=m)
    lat, lon = cells["lat"][rid].astype(np.float64), cells["lon"][rid].astype(np.float64)
    w = rng.dirichlet([1, 1, 1], size=m)
    edge = rng.random(m) < 0.25
    w[edge, 0] *= 1e-3
    corner = rng.random(m) < 0.1
    w[corner, 1] *= 1e-3
    w /= w.sum(1)[:, None]
    U = np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], -1)
    d = (w[:, :, None] * U).sum(1)
    d /= np.linalg.norm(d, axis=1)[:, None]
    nl = cells["numLayers"][rid]
    lay = np.minimum((rng.random(m) * nl).astype(int), nl - 1)
    h0, h1 = cells["height"][rid, lay].astype(np.float64), cells["height"][rid, lay + 1].astype(np.float64)
    f = rng.random(m)
    f[rng.random(m) < 0.15] *= 1e-3
    return (d * (h0 + f * (h1 - h0))[:, None]).astype(np.float32)


def make(name, cells, m, seed):
    rng = np.random.default_rng(seed)
    prism = port.prism_build(cells)
    pos = positions(cells, m, rng)
    ids, vals, layers, pc = port.locate_prism(cells, prism, pos)
    cid, _ = ref.locate(cells, pos)
    assert np.array_equal(ids, cid)
    V = port.prism_wedges(cells, prism, ids, layers)
    acc, rv = ref.wedge(pos, V)
    acc[ids < 0] = 0
    ok = acc == 1
    span = V[:, :, 3].max(1) - V[:, :, 3].min(1)
    err = np.abs(vals - rv)[ok]
    print(f"{name}: {m} points, located {np.mean(ids >= 0):.3f}, reference accepts {ok.mean():.3f} of them, "
          f"max |closed form - Newton| = {err.max():.3g} ({(err / np.maximum(span[ok], 1e-6)).max():.3g} of the wedge's scalar span)")
    np.savez_compressed(os.path.join(GOLD, f"prism_{name}.npz"), pos=pos, ids=ids, layers=layers, wedges=V,
                        ref_accept=acc.astype(np.int8), ref_value=rv, port_value=vals, port_pcoords=pc, seed=seed)


cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
make("r2b1_l40", cells, 3000, 20240611)
# a patch of a finer grid (where the chord/sphere sag is below the layer thickness and the Newton iteration
# accepts most located prisms): the first 1200 records of R2B5 x 20 levels, `--num-cells 1200` style truncation
fine = icb.gen_grid(2, 5, 20)[:1200]
fine.tofile(os.path.join(GOLD, "r2b5_l20_head1200.ic"))
make("r2b5_l20_head1200", fine, 3000, 20240612)
