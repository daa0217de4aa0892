"""icb_trace_pixel_static against icb_trace_pixel, through the C ABI.

icb_trace_pixel_static is the variant of the per-pixel trace for hosts that cannot let the library allocate device
memory while they run: a present loop with frames in flight (icb_trace_pixel frees its staging buffers on return, and a
cudaFree waits for every frame still on the device), or a process whose device memory is already fully committed to
the grid replica and to the application's own pool.  Its staging is an array built into the library, shared by every
call on a GPU and never cleared as a whole -- so the test makes several calls in a row, a long ray first and shorter
ones after it, and a second handle afterwards: anything left behind by an earlier call would show up as a wrong count,
a wrong id or a wrong colour.  Results must equal icb_trace_pixel's bit for bit (ids, step count and fp32 RGBA: it is
the same kernel writing to a different address), and the goldens of the reference-built oracle must hold as well.

The file sorts last on purpose (pytest -x stops at the first failure); it passes alone in a fresh process -- the first
call then also does the once-per-GPU address look-up -- and after the rest of the suite.
"""
import os

import numpy as np
import pytest

import icb

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PIXELS = [(24, 16), (5, 16), (40, 9), (23, 30), (24, 16)]     # the first one again at the end: a long ray after short ones


def _scene(ctx, g, cells, lut):
    ctx.set_grid(cells)
    ctx.set_transfunc(lut)
    ctx.set_camera({k: g[f"sharp_jit_cam_{k}"] for k in ("org", "d00", "du", "dv")})
    ctx.set_params(unit_distance=16000.0, dt=4000.0, accum_id=0, jitter=True)


@pytest.mark.gpu
def test_static_trace_equals_allocating_trace():
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    lut = np.load(os.path.join(GOLD, "blocks.npz"))["lut"].copy()
    lut[:, 3] = 0.0
    lut[150:, 3] = 1.0
    W, H = int(g["W"]), int(g["H"])
    ctx = icb.Context(0)
    try:
        _scene(ctx, g, cells, lut)
        for (x, y) in PIXELS:                                  # consecutive calls share the built-in array
            ids_s, n_s, rgba_s = ctx.trace_pixel_static(x, y, W, H)
            ids_a, n_a, rgba_a = ctx.trace_pixel(x, y, W, H)
            assert n_s == n_a and (n_s > 0) == ((x, y) == (24, 16))     # the other golden pixels miss the globe: a count of 0
            assert np.array_equal(ids_s, ids_a)
            assert np.array_equal(rgba_s.view(np.uint32), rgba_a.view(np.uint32))
            assert np.array_equal(ids_s, g[f"trace_{x}_{y}_ids"])          # and the oracle's id sequence
        # a smooth, nearly transparent transfer function: rays run to the end of the shell (longer sequences, other colours)
        ctx.set_transfunc(np.load(os.path.join(GOLD, "blocks.npz"))["lut"])
        seen = 0
        for (x, y) in [(20, 14), (30, 20), (24, 16), (0, 0), (27, 12)]:      # (0, 0) misses the shell: nothing after a long ray
            ids_s, n_s, rgba_s = ctx.trace_pixel_static(x, y, W, H)
            ids_a, n_a, rgba_a = ctx.trace_pixel(x, y, W, H)
            assert n_s == n_a and np.array_equal(ids_s, ids_a)
            assert np.array_equal(rgba_s.view(np.uint32), rgba_a.view(np.uint32))
            seen += int(n_s > 0 and rgba_s[3] > 0)
        assert seen >= 3                                       # the comparison above was not between empty results
        # a short ids buffer: the count still covers the whole ray, only the first max_ids ids come back
        ids_s, n_s, _ = ctx.trace_pixel_static(24, 16, W, H, max_ids=3)
        ids_a, n_a, _ = ctx.trace_pixel(24, 16, W, H, max_ids=3)
        assert n_s == n_a and np.array_equal(ids_s, ids_a) and len(ids_s) == min(3, n_s)
        # a second handle on the same GPU uses the same array (after the first one's calls have returned)
        ctx2 = icb.Context(0)
        try:
            _scene(ctx2, g, cells, lut)
            ids_2, n_2, rgba_2 = ctx2.trace_pixel_static(40, 9, W, H)
            ids_a, n_a, rgba_a = ctx2.trace_pixel(40, 9, W, H)
            assert n_2 == n_a and np.array_equal(ids_2, ids_a) and np.array_equal(rgba_2.view(np.uint32), rgba_a.view(np.uint32))
        finally:
            ctx2.close()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_static_trace_rejects_what_does_not_fit():
    g = np.load(os.path.join(GOLD, "frames_r2b1_l40.npz"))
    cells = np.fromfile(os.path.join(GOLD, "r2b1_l40.ic"), icb.CELL_DTYPE)
    W, H = int(g["W"]), int(g["H"])
    ctx = icb.Context(0)
    try:
        _scene(ctx, g, cells, np.load(os.path.join(GOLD, "blocks.npz"))["lut"])
        with pytest.raises(icb.IcbError, match="ICB_TRACE_STATIC_MAX_IDS"):
            ctx.trace_pixel_static(24, 16, W, H, max_ids=icb.TRACE_STATIC_MAX_IDS + 1)
        ids, n, _ = ctx.trace_pixel_static(24, 16, W, H, max_ids=icb.TRACE_STATIC_MAX_IDS)     # the limit itself is fine
        assert n > 0 and len(ids) == min(n, icb.TRACE_STATIC_MAX_IDS)
        with pytest.raises(icb.IcbError):
            ctx.trace_pixel_static(W, 0, W, H)                 # pixel outside the frame
    finally:
        ctx.close()


def test_static_trace_limit_matches_header():
    """no GPU needed: the Python constant is the header's, and the symbol is exported"""
    hdr = open(os.path.join(os.path.dirname(GOLD), "..", "include", "icon_b200.h")).read()
    assert f"#define ICB_TRACE_STATIC_MAX_IDS {icb.TRACE_STATIC_MAX_IDS}" in hdr
    assert hasattr(icb.lib(), "icb_trace_pixel_static")
