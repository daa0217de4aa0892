"""ctypes binding of the C ABI in include/icon_b200.h (libicon_b200.so) plus the Python mirror of
the reference's small host objects (Camera, Transfunc, Frame -- common/camera.h, transfunc.h, fb.h).

The library is CUDA-only: ``Context()`` raises when no device / no built library is present; there
is no CPU fallback.  The grid generator entry points (``gen_grid``) need no device.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ICB_LIB", os.path.join(HERE, "libicon_b200.so"))   # ICB_LIB: kernel-variant experiments

CELL_DTYPE = np.dtype(
    [("lat", "<f4", 3), ("lon", "<f4", 3), ("numLayers", "<i4"), ("height", "<f4", 32), ("value", "<f4", 32)]
)
assert CELL_DTYPE.itemsize == 284  # ICONGrid.h:59-76

FLAG_JITTER = 1
LOCATE_WALK = 1
LOCATE_EXACT = 2
TRACE_STATIC_MAX_IDS = 16384   # ICB_TRACE_STATIC_MAX_IDS
TILE = 16


class IcbError(RuntimeError):
    pass


class Bounds(C.Structure):
    _fields_ = [("volLo", C.c_float * 3), ("volHi", C.c_float * 3), ("sphLo", C.c_float * 3), ("sphHi", C.c_float * 3),
                ("dataRange", C.c_float * 2), ("unitDistance", C.c_float)]


class Stats(C.Structure):
    _fields_ = [("numRecords", C.c_int64), ("numColumns", C.c_int64), ("numVertices", C.c_int64),
                ("numBins", C.c_int64), ("numBinEntries", C.c_int64), ("numGenericColumns", C.c_int64), ("numGlobalRecords", C.c_int64),
                ("deviceBytes", C.c_int64), ("setupSeconds", C.c_double),
                ("steps", C.c_uint64), ("hits", C.c_uint64), ("slowSteps", C.c_uint64), ("exactQueries", C.c_uint64),
                ("raysMarched", C.c_uint64), ("totalSteps", C.c_uint64), ("totalHits", C.c_uint64), ("kernelLaunches", C.c_int32), ("lastKernelMs", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def lib():
    """Load libicon_b200.so (fails loudly if it has not been built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise IcbError(f"{LIB_PATH} is missing: build it with `make -C {HERE}` (no fallback exists)")
        L = C.CDLL(LIB_PATH)
        L.icb_last_error.restype = C.c_char_p
        L.icb_last_error.argtypes = [C.c_void_p]
        L.icb_version.restype = C.c_char_p
        L.icb_gen_num_columns.restype = C.c_int64
        L.icb_gen_num_records.restype = C.c_int64
        _lib = L
    return _lib


_host = None
HOST_LIB_PATH = os.path.join(HERE, "libicon_hostgen.so")


def host_lib():
    """the host-only helpers (grid generator, Camera / resampleLUT mirror): libicon_hostgen.so -- the same objects that are
    linked into libicon_b200.so, without any CUDA code, so that a process that only needs them (bench.py --impl reference)
    never maps the product library"""
    global _host
    if _host is None:
        if _lib is not None or not os.path.exists(HOST_LIB_PATH):
            _host = lib()
        else:
            _host = C.CDLL(HOST_LIB_PATH)
            _host.icb_gen_num_columns.restype = C.c_int64
            _host.icb_gen_num_records.restype = C.c_int64
    return _host


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def gen_grid(n, k, levels, exaggeration=10.0, timestep=0):
    """Synthetic RnBk grid as a numpy array of 284-byte ICONCell records."""
    L = host_lib()
    nrec = L.icb_gen_num_records(n, k, levels)
    if nrec < 0:
        raise IcbError("gen_grid: bad arguments")
    cells = np.zeros(nrec, CELL_DTYPE)
    rc = L.icb_gen_grid(n, k, levels, C.c_float(exaggeration), int(timestep), C.c_void_p(cells.ctypes.data), C.c_int64(nrec))
    if rc != 0:
        raise IcbError(f"icb_gen_grid failed ({rc})")
    return cells


def gen_values(n, k, levels, timestep0, ntimesteps, first_record, num_records, out_ptr):
    """The scalars alone (32 floats per record) of a record range for consecutive timesteps, written to
    out_ptr[ntimesteps][num_records * 32] -- a rank's slab of a time series (icb_gen_values)."""
    rc = host_lib().icb_gen_values(int(n), int(k), int(levels), int(timestep0), int(ntimesteps), C.c_int64(first_record),
                                   C.c_int64(num_records), C.c_void_p(out_ptr))
    if rc != 0:
        raise IcbError(f"icb_gen_values failed ({rc})")


INGEST_REFERENCE, INGEST_FIXED, INGEST_SURFACE_LEVEL = 0, 1, 2


def ingest_normalize(var):
    """a data level shifted and scaled to [0,1] by its own extrema in double, narrowed to float (convert_icon.cpp:317-333)"""
    v = np.ascontiguousarray(var, np.float64).reshape(-1)
    out = np.empty(v.size, np.float32)
    rc = host_lib().icb_ingest_normalize(C.c_void_p(v.ctypes.data), C.c_int64(v.size), C.c_void_p(out.ctypes.data))
    if rc != 0:
        raise IcbError(f"icb_ingest_normalize failed ({rc})")
    return out


def ingest_columns(clon_vertices, clat_vertices, hsurf, hhl_heights, hhl_values, data_heights, data_values, max_layers=-1,
                   mode=INGEST_FIXED):
    """convert_icon's array -> `.ic` half (icb_ingest_columns): level fields in FILE order with their `height` indices;
    returns the 284-byte records.  hhl_values / data_values: sequences of float32 arrays, one per level, one value per cell."""
    L = host_lib()
    L.icb_ingest_columns.restype = C.c_int64
    clon = np.ascontiguousarray(clon_vertices, np.float64).reshape(-1)
    clat = np.ascontiguousarray(clat_vertices, np.float64).reshape(-1)
    hs = np.ascontiguousarray(hsurf, np.float64).reshape(-1)
    cells = hs.size
    hv = [np.ascontiguousarray(a, np.float32).reshape(-1) for a in hhl_values]
    dv = [np.ascontiguousarray(a, np.float32).reshape(-1) for a in data_values]
    hh = np.ascontiguousarray(hhl_heights, np.int32)
    dh = np.ascontiguousarray(data_heights, np.int32)
    hp = (C.c_void_p * max(1, len(hv)))(*[a.ctypes.data for a in hv])
    dp = (C.c_void_p * max(1, len(dv)))(*[a.ctypes.data for a in dv])
    args = [C.c_int64(cells), C.c_void_p(clon.ctypes.data), C.c_void_p(clat.ctypes.data), C.c_void_p(hs.ctypes.data),
            len(hv), C.c_void_p(hh.ctypes.data), hp, len(dv), C.c_void_p(dh.ctypes.data), dp, int(max_layers), int(mode)]
    n = L.icb_ingest_columns(*args, None, C.c_int64(0))
    if n < 0:
        raise IcbError(f"icb_ingest_columns failed ({n})")
    cells_out = np.zeros(n, CELL_DTYPE)
    n2 = L.icb_ingest_columns(*args, C.c_void_p(cells_out.ctypes.data), C.c_int64(n))
    if n2 != n:
        raise IcbError(f"icb_ingest_columns failed ({n2})")
    return cells_out


def read_ic(path, max_cells=-1, lat_range=(-90.0, 90.0), lon_range=(-180.0, 180.0)):
    """Load a `.ic` file exactly like icon_rt/hostCode.cu:717-758: truncate to --num-cells, then drop every
    record with any corner outside the lat/lon window (degrees)."""
    cells = np.fromfile(path, CELL_DTYPE)
    if max_cells >= 0:
        cells = cells[:max_cells]
    d2r = lambda d: np.float32(np.float32(d) * np.float32(np.pi) / np.float32(180.0))  # ICONGrid.h:26-29
    la0, la1, lo0, lo1 = d2r(lat_range[0]), d2r(lat_range[1]), d2r(lon_range[0]), d2r(lon_range[1])
    lat, lon = cells["lat"], cells["lon"]
    drop = (lat < la0).any(1) | (lat > la1).any(1) | (lon < lo0).any(1) | (lon > lo1).any(1)
    return np.ascontiguousarray(cells[~drop])


class Context:
    """One handle of the C ABI (one per GPU)."""

    def __init__(self, device=0):
        self.L = lib()
        h = C.c_void_p()
        rc = self.L.icb_create(int(device), C.byref(h))
        if rc != 0:
            raise IcbError(f"icb_create failed ({rc}): {self.L.icb_last_error(None).decode()}")
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.L.icb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc < 0:
            raise IcbError(f"{what} failed ({rc}): {self.L.icb_last_error(self.h).decode()}")
        return rc

    # ---- scene -------------------------------------------------------------------------------
    def set_locator(self, margin_m=-1.0, max_walk=-1):
        self._ck(self.L.icb_set_locator(self.h, C.c_float(margin_m), int(max_walk)), "icb_set_locator")

    def set_skip_ahead(self, enabled):
        self._ck(self.L.icb_set_skip_ahead(self.h, int(bool(enabled))), "icb_set_skip_ahead")

    def set_ray_split(self, chunk_log2=-1, max_heavy_tiles=-1, min_tile_cost_log2=-1, mode=-1):
        """icb_set_ray_split: compositing chunk = 2^chunk_log2 samples; up to max_heavy_tiles tiles per frame (cost >=
        2^min_tile_cost_log2 lane iterations) have the chunks of their rays marched by different lanes; mode 0 never, 1 always,
        2 auto (sort-first shares only); -1 keeps a value"""
        self._ck(self.L.icb_set_ray_split(self.h, int(chunk_log2), int(max_heavy_tiles), int(min_tile_cost_log2), int(mode)), "icb_set_ray_split")

    def set_integrator(self, name):
        """'march' (default), 'woodcock' (the reference's woodcockTrackingAE ray-gen), 'woodcock_grid' (its default
        ray-gen woodcockTrackingWithAccel over the macrocell majorant grid) or 'woodcock_sphere' (the same ray-gen over
        the shell accel, the accel mode the reference starts in)"""
        self._ck(self.L.icb_set_integrator(self.h, {"march": 0, "woodcock": 1, "woodcock_grid": 2, "woodcock_sphere": 3}[name]), "icb_set_integrator")

    def set_accel_shell(self, n_rad, n_lat, n_lon):
        """(rad, lat, lon) shell accel of the 'woodcock_sphere' integrator (reference: 1 x 1024 x 1024, hostCode.cu:654)"""
        self._ck(self.L.icb_set_accel_shell(self.h, int(n_rad), int(n_lat), int(n_lon)), "icb_set_accel_shell")
        self._shell_dims = (int(n_rad), int(n_lat), int(n_lon))

    def shell_majorants(self):
        """(maxOpacities[n], valueRanges[n,2]) of the shell accel as the ray-gen reads it"""
        n = int(np.prod(getattr(self, "_shell_dims", (1, 1024, 1024))))
        mo = np.empty(n, np.float32)
        vr = np.empty((n, 2), np.float32)
        self._ck(self.L.icb_get_shell_majorants(self.h, _fp(mo), _fp(vr), C.c_size_t(n)), "icb_get_shell_majorants")
        return mo, vr

    def set_accel_grid(self, nx, ny, nz):
        """macrocell grid of the 'woodcock_grid' integrator (reference: 256^3, hostCode.cu:670)"""
        self._ck(self.L.icb_set_accel_grid(self.h, int(nx), int(ny), int(nz)), "icb_set_accel_grid")
        self._accel_dims = (int(nx), int(ny), int(nz))

    def majorants(self):
        """(maxOpacities[n], valueRanges[n,2]) of the macrocell grid as the ray-gen reads it"""
        n = int(np.prod(getattr(self, "_accel_dims", (256, 256, 256))))
        mo = np.empty(n, np.float32)
        vr = np.empty((n, 2), np.float32)
        self._ck(self.L.icb_get_majorants(self.h, _fp(mo), _fp(vr), C.c_size_t(n)), "icb_get_majorants")
        return mo, vr

    def selftest_logf(self):
        """(mismatches of the FMA form, of the SSE2 form) of the device's logf(1 - k/2^24) against the host libm, all 2^24 k"""
        bad = (C.c_int64 * 2)()
        self._ck(self.L.icb_selftest_logf(self.h, bad), "icb_selftest_logf")
        return int(bad[0]), int(bad[1])

    def set_sampler(self, name):
        """'cell' (default: value[layer], ICONGrid.h:146-153) or 'prism' (wedge shape functions, UElems.h:176-187)"""
        self._ck(self.L.icb_set_sampler(self.h, {"cell": 0, "prism": 1}[name]), "icb_set_sampler")

    def set_grid(self, cells):
        cells = np.ascontiguousarray(cells)
        assert cells.dtype.itemsize == 284
        self._ck(self.L.icb_set_grid(self.h, C.c_void_p(cells.ctypes.data), C.c_size_t(len(cells))), "icb_set_grid")

    def set_scalars(self, values):
        v = np.ascontiguousarray(values, np.float32).reshape(-1, 32)
        self._ck(self.L.icb_set_scalars(self.h, _fp(v), C.c_size_t(len(v))), "icb_set_scalars")

    def set_scalars_dev(self, ptr, n, stream=0):
        self._ck(self.L.icb_set_scalars_dev(self.h, C.c_void_p(ptr), C.c_size_t(n), C.c_void_p(stream)), "icb_set_scalars_dev")

    def bounds(self):
        b = Bounds()
        self._ck(self.L.icb_get_bounds(self.h, C.byref(b)), "icb_get_bounds")
        return dict(volLo=np.array(b.volLo[:], np.float32), volHi=np.array(b.volHi[:], np.float32),
                    sphLo=np.array(b.sphLo[:], np.float32), sphHi=np.array(b.sphHi[:], np.float32),
                    dataRange=np.array(b.dataRange[:], np.float32), unitDistance=float(b.unitDistance))

    def set_transfunc(self, lut, value_range=(0.0, 1.0), opacity_scale=1.0):
        lut = np.ascontiguousarray(lut, np.float32).reshape(-1, 4)
        self._ck(self.L.icb_set_transfunc(self.h, _fp(lut), len(lut), C.c_float(value_range[0]), C.c_float(value_range[1]),
                                          C.c_float(opacity_scale)), "icb_set_transfunc")

    def set_camera(self, cam):
        self._ck(self.L.icb_set_camera(self.h, _f3(cam["org"]), _f3(cam["d00"]), _f3(cam["du"]), _f3(cam["dv"])), "icb_set_camera")

    def set_params(self, unit_distance=0.0, dt=0.0, ambient_color=(1.0, 1.0, 1.0), ambient_radiance=1.0, accum_id=0,
                   jitter=False, term_alpha=0.99):
        self._ck(self.L.icb_set_params(self.h, C.c_float(unit_distance), C.c_float(dt), _f3(ambient_color), C.c_float(ambient_radiance),
                                       int(accum_id), C.c_uint(FLAG_JITTER if jitter else 0), C.c_float(term_alpha)), "icb_set_params")

    def set_tiles(self, tile_ids):
        t = np.ascontiguousarray(tile_ids, np.int32)
        self._ck(self.L.icb_set_tiles(self.h, t.ctypes.data_as(C.POINTER(C.c_int32)), len(t)), "icb_set_tiles")

    # ---- hot path, device pointers (ints) ----------------------------------------------------
    def render(self, fb_ptr, accum_ptr, width, height, depth_ptr=0, stream=0):
        self._ck(self.L.icb_render(self.h, C.c_void_p(fb_ptr), C.c_void_p(accum_ptr), C.c_void_p(depth_ptr), int(width), int(height),
                                   C.c_void_p(stream)), "icb_render")

    def render_trace(self, fb_ptr, accum_ptr, per_pixel_ptr, width, height, stream=0):
        self._ck(self.L.icb_render_trace(self.h, C.c_void_p(fb_ptr), C.c_void_p(accum_ptr), C.c_void_p(per_pixel_ptr), int(width),
                                         int(height), C.c_void_p(stream)), "icb_render_trace")

    def render_tiles(self, rgba8_ptr, accum_ptr, width, height, stream=0):
        self._ck(self.L.icb_render_tiles(self.h, C.c_void_p(rgba8_ptr), C.c_void_p(accum_ptr), int(width), int(height),
                                         C.c_void_p(stream)), "icb_render_tiles")

    def scatter_tiles(self, rgba8_ptr, accum_ptr, tile_ids_ptr, n, fb_ptr, accum_out_ptr, width, height, stream=0):
        self._ck(self.L.icb_scatter_tiles(self.h, C.c_void_p(rgba8_ptr), C.c_void_p(accum_ptr), C.c_void_p(tile_ids_ptr), int(n),
                                          C.c_void_p(fb_ptr), C.c_void_p(accum_out_ptr), int(width), int(height), C.c_void_p(stream)),
                 "icb_scatter_tiles")

    def set_frame_slot(self, slot):
        """per-frame device state slot (0/1) used by set_tiles*/render*/tile_costs/stats: two frames in flight"""
        self._ck(self.L.icb_set_frame_slot(self.h, int(slot)), "icb_set_frame_slot")

    def render_host_async(self, width, height, fb_pinned):
        """fb_pinned: uint32 array from host_alloc; valid after frame_wait(slot)"""
        self._ck(self.L.icb_render_host_async(self.h, C.c_void_p(fb_pinned.ctypes.data), int(width), int(height)), "icb_render_host_async")

    def frame_wait(self, slot):
        self._ck(self.L.icb_frame_wait(self.h, int(slot)), "icb_frame_wait")

    def set_frames_in_flight(self, n):
        self._ck(self.L.icb_set_frames_in_flight(self.h, int(n)), "icb_set_frames_in_flight")

    def set_tile_costs(self, costs_ptr, num_tiles, stream=0):
        self._ck(self.L.icb_set_tile_costs(self.h, C.c_void_p(costs_ptr), int(num_tiles), C.c_void_p(stream)), "icb_set_tile_costs")

    def set_tiles_async(self, pinned_ptr, n, stream=0):
        self._ck(self.L.icb_set_tiles_async(self.h, C.c_void_p(pinned_ptr), int(n), C.c_void_p(stream)), "icb_set_tiles_async")

    def tile_costs(self, costs_ptr, num_tiles, stream=0):
        self._ck(self.L.icb_tile_costs(self.h, C.c_void_p(costs_ptr), int(num_tiles), C.c_void_p(stream)), "icb_tile_costs")

    def locate(self, pos_ptr, n, ids_ptr, values_ptr, flags=0, stream=0):
        self._ck(self.L.icb_locate(self.h, C.c_void_p(pos_ptr), C.c_size_t(n), C.c_void_p(ids_ptr), C.c_void_p(values_ptr),
                                   C.c_uint(flags), C.c_void_p(stream)), "icb_locate")

    # ---- hot path, host buffers ------------------------------------------------------------------
    def render_host(self, width, height, fb=None, accum=None, want_accum=True):
        """icb_render_host: frame down into host buffers.  want_accum=False passes NULL for the fp32 accumulation
        buffer (it then stays on the device between frames, like the reference's Frame::accumBuffer)."""
        fb = np.zeros(width * height, np.uint32) if fb is None else fb
        if want_accum and accum is None:
            accum = np.zeros((width * height, 4), np.float32)
        self._ck(self.L.icb_render_host(self.h, fb.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(accum) if want_accum else None,
                                        int(width), int(height)), "icb_render_host")
        return fb, accum

    def locate_host(self, pos, flags=0):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        ids = np.empty(len(pos), np.int32)
        vals = np.empty(len(pos), np.float32)
        self._ck(self.L.icb_locate_host(self.h, _fp(pos), C.c_size_t(len(pos)), ids.ctypes.data_as(C.POINTER(C.c_int32)), _fp(vals),
                                        C.c_uint(flags)), "icb_locate_host")
        return ids, vals

    def trace_pixel(self, x, y, width, height, max_ids=65536):
        ids = np.full(max_ids, -2, np.int32)
        rgba = np.zeros(4, np.float32)
        n = self._ck(self.L.icb_trace_pixel(self.h, int(x), int(y), int(width), int(height), ids.ctypes.data_as(C.POINTER(C.c_int32)),
                                            int(max_ids), _fp(rgba)), "icb_trace_pixel")
        return ids[: min(n, max_ids)].copy(), n, rgba

    def trace_pixel_static(self, x, y, width, height, max_ids=TRACE_STATIC_MAX_IDS):
        """icb_trace_pixel_static: trace_pixel for callers that must not let the library allocate (a present loop with
        frames in flight, a process whose device memory is fully committed).  Same results; max_ids is limited to
        TRACE_STATIC_MAX_IDS, the size of the staging array built into the library; calls on one GPU must not overlap."""
        ids = np.full(max(max_ids, 1), -2, np.int32)
        rgba = np.zeros(4, np.float32)
        n = self._ck(self.L.icb_trace_pixel_static(self.h, int(x), int(y), int(width), int(height),
                                                   ids.ctypes.data_as(C.POINTER(C.c_int32)), int(max_ids), _fp(rgba)),
                     "icb_trace_pixel_static")
        return ids[: min(n, max_ids)].copy(), n, rgba

    def reset_totals(self):
        self._ck(self.L.icb_reset_totals(self.h), "icb_reset_totals")

    def host_alloc(self, nbytes):
        """page-locked host memory as a numpy uint8 array (freed with the context)"""
        p = C.c_void_p()
        self._ck(self.L.icb_host_alloc(self.h, C.c_size_t(nbytes), C.byref(p)), "icb_host_alloc")
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,))

    def dev_download(self, dst_host_array, src_ptr, nbytes):
        self._ck(self.L.icb_dev_download(self.h, C.c_void_p(dst_host_array.ctypes.data), C.c_void_p(src_ptr), C.c_size_t(nbytes)),
                 "icb_dev_download")

    def debug_counters(self):
        out = (C.c_uint64 * 8)()
        self._ck(self.L.icb_debug_counters(self.h, out), "icb_debug_counters")
        names = ("warp_iters", "lane_iters", "fast", "layer", "fetch", "retry", "no_skip", "_")
        return dict(zip(names, [int(v) for v in out]))

    def stats(self):
        s = Stats()
        self._ck(self.L.icb_stats(self.h, C.byref(s)), "icb_stats")
        return s.as_dict()


class GroupStats(C.Structure):
    _fields_ = [("world", C.c_int32), ("localRanks", C.c_int32), ("framesInFlight", C.c_int32), ("pad", C.c_int32),
                ("frames", C.c_int64), ("hostMsPerFrame", C.c_double), ("waitMsPerFrame", C.c_double)]


GROUP_ID_BYTES = 128


def group_unique_id():
    """icb_group_unique_id (ncclGetUniqueId): made on rank 0, handed to every rank by the launcher's own means"""
    buf = (C.c_uint8 * GROUP_ID_BYTES)()
    rc = lib().icb_group_unique_id(buf)
    if rc != 0:
        lib().icb_group_last_error.restype = C.c_char_p
        raise IcbError(f"icb_group_unique_id failed ({rc}): {lib().icb_group_last_error(None).decode()}")
    return bytes(buf)


def partition_tiles(tiles_x, tiles_y, costs, n_ranks):
    """icb_partition_tiles (host only): the tile sets icb_group_render deals to the ranks; list of int32 arrays"""
    n = tiles_x * tiles_y
    ids = np.empty(n, np.int32)
    off = np.empty(n_ranks + 1, np.int32)
    c = None if costs is None else np.ascontiguousarray(costs, np.int32)
    rc = lib().icb_partition_tiles(int(tiles_x), int(tiles_y), None if c is None else c.ctypes.data_as(C.POINTER(C.c_int32)), int(n_ranks),
                                   ids.ctypes.data_as(C.POINTER(C.c_int32)), off.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc != 0:
        raise IcbError(f"icb_partition_tiles failed ({rc})")
    return [ids[off[r]:off[r + 1]].copy() for r in range(n_ranks)]


class Group:
    """icb_group_*: this process's share of a sort-first group (one rank per process: Group(ctx, rank, world, id);
    all ranks in one process: Group.local([ctx0, ctx1, ...])).  The handles stay owned by the caller."""

    def __init__(self, ctx=None, rank=0, world=1, uid=None, _local=None):
        self.L = lib()
        self.L.icb_group_last_error.restype = C.c_char_p
        self.L.icb_group_last_error.argtypes = [C.c_void_p]
        g = C.c_void_p()
        if _local is not None:
            arr = (C.c_void_p * len(_local))(*[c.h for c in _local])
            rc = self.L.icb_group_create_local(arr, len(_local), C.byref(g))
            self.ctxs = list(_local)
        else:
            idbuf = (C.c_uint8 * GROUP_ID_BYTES)(*(uid if uid is not None else bytes(GROUP_ID_BYTES)))
            rc = self.L.icb_group_create(ctx.h, int(rank), int(world), idbuf, C.byref(g))
            self.ctxs = [ctx]
        if rc != 0:
            raise IcbError(f"icb_group_create failed ({rc}): {self.L.icb_group_last_error(None).decode()}")
        self.g, self.rank, self.world = g, rank, world if _local is None else len(_local)

    @classmethod
    def local(cls, ctxs):
        return cls(_local=ctxs)

    def close(self):
        if getattr(self, "g", None):
            self.L.icb_group_destroy(self.g)
            self.g = None

    def _ck(self, rc, what):
        if rc < 0:
            raise IcbError(f"{what} failed ({rc}): {self.L.icb_group_last_error(self.g).decode()}")
        return rc

    def set_grid(self, cells):
        """cells: the records on rank 0 (built once there, device arrays broadcast over NVLink), None elsewhere"""
        if cells is None:
            return self._ck(self.L.icb_group_set_grid(self.g, None, C.c_size_t(0)), "icb_group_set_grid")
        cells = np.ascontiguousarray(cells)
        assert cells.dtype.itemsize == 284
        self._ck(self.L.icb_group_set_grid(self.g, C.c_void_p(cells.ctypes.data), C.c_size_t(len(cells))), "icb_group_set_grid")

    def set_transfunc(self, lut, value_range=(0.0, 1.0), opacity_scale=1.0):
        lut = np.ascontiguousarray(lut, np.float32).reshape(-1, 4)
        self._ck(self.L.icb_group_set_transfunc(self.g, _fp(lut), len(lut), C.c_float(value_range[0]), C.c_float(value_range[1]),
                                                C.c_float(opacity_scale)), "icb_group_set_transfunc")

    def set_camera(self, cam):
        self._ck(self.L.icb_group_set_camera(self.g, _f3(cam["org"]), _f3(cam["d00"]), _f3(cam["du"]), _f3(cam["dv"])), "icb_group_set_camera")

    def set_params(self, unit_distance=0.0, dt=0.0, ambient_color=(1.0, 1.0, 1.0), ambient_radiance=1.0, accum_id=0, jitter=False, term_alpha=0.99):
        self._ck(self.L.icb_group_set_params(self.g, C.c_float(unit_distance), C.c_float(dt), _f3(ambient_color), C.c_float(ambient_radiance),
                                             int(accum_id), C.c_uint(FLAG_JITTER if jitter else 0), C.c_float(term_alpha)), "icb_group_set_params")

    def begin(self, width, height, frames_in_flight=1):
        self._ck(self.L.icb_group_begin(self.g, int(width), int(height), int(frames_in_flight)), "icb_group_begin")
        self.W, self.H = width, height

    def render(self, fb_pinned=None):
        """one sort-first frame, enqueued without waiting for the device; returns the slot it occupies"""
        p = C.c_void_p(fb_pinned.ctypes.data) if fb_pinned is not None else None
        return self._ck(self.L.icb_group_render(self.g, p), "icb_group_render")

    def wait(self, slot=-1):
        self._ck(self.L.icb_group_wait(self.g, int(slot)), "icb_group_wait")

    def frame_ptr(self, slot):
        """device pointer of rank 0's assembled RGBA8 frame of `slot` (0 where rank 0 is not local)"""
        p = C.c_void_p()
        self._ck(self.L.icb_group_frame(self.g, int(slot), C.byref(p)), "icb_group_frame")
        return p.value or 0

    def stats(self):
        s = GroupStats()
        self._ck(self.L.icb_group_stats(self.g, C.byref(s)), "icb_group_stats")
        return {k: getattr(s, k) for k, _ in s._fields_ if k != "pad"}

    def reset_stats(self):
        self._ck(self.L.icb_group_reset_stats(self.g), "icb_group_reset_stats")

    def stream_scalars(self, values_pinned_ptr, n_records):
        self._ck(self.L.icb_group_stream_scalars(self.g, C.c_void_p(values_pinned_ptr), C.c_size_t(n_records)), "icb_group_stream_scalars")

    def scalar_slab(self, n_records):
        """(first_float, num_floats) of a timestep's n*32 floats that stream_scalars reads in this process"""
        a, c = C.c_size_t(0), C.c_size_t(0)
        self._ck(self.L.icb_group_scalar_slab(self.g, C.c_size_t(n_records), C.byref(a), C.byref(c)), "icb_group_scalar_slab")
        return int(a.value), int(c.value)

    def swap_scalars(self):
        self._ck(self.L.icb_group_swap_scalars(self.g), "icb_group_swap_scalars")


# ------------------------------------------------------------------------------------------------
# host objects mirroring the reference's; the arithmetic lives in the C++ host mirror
# (host/icon_host.hpp, same libm and operand order as common/camera.h) and is reached through the
# library's C wrappers, so Python callers get the same bits as C++ callers.
# ------------------------------------------------------------------------------------------------
f32 = np.float32
FOVY_DEFAULT = f32(90.0 * np.pi / 180.0)  # camera.h:108


def _cam_out():
    return [(C.c_float * 3)() for _ in range(4)]


def _cam_dict(o):
    return dict(org=np.array(o[0][:], np.float32), d00=np.array(o[1][:], np.float32),
                du=np.array(o[2][:], np.float32), dv=np.array(o[3][:], np.float32))


def camera_view_all(vol_lo, vol_hi, width, height, aspect=None, fovy=FOVY_DEFAULT):
    """Camera::viewAll + getScreen -> launch-param camera (camera.h:86-104, hostCode.cu:936-945)."""
    o = _cam_out()
    aspect = width / height if aspect is None else aspect
    host_lib().icb_camera_view_all(_f3(vol_lo), _f3(vol_hi), C.c_float(fovy), C.c_float(aspect), int(width), int(height), *o)
    return _cam_dict(o)


def camera_orient(eye, poi, up, width, height, aspect=None, fovy=FOVY_DEFAULT):
    """Camera::setOrientation + getScreen (camera.h:34-96)."""
    o = _cam_out()
    aspect = width / height if aspect is None else aspect
    host_lib().icb_camera_orient(_f3(eye), _f3(poi), _f3(up), C.c_float(fovy), C.c_float(aspect), int(width), int(height), *o)
    return _cam_dict(o)


def camera_orbit(vol_lo, vol_hi, width, height, theta_deg, aspect=None, fovy=FOVY_DEFAULT, pose_fovy=None):
    """SURVEY.md §8d cameras: the viewAll eye rotated about +y by theta around the box centre, up=(0,1,0).
    pose_fovy: field of view used for the viewAll eye distance only (default: fovy); a narrower `fovy` from the
    same eye zooms in on the globe."""
    base = camera_view_all(vol_lo, vol_hi, width, height, aspect, fovy if pose_fovy is None else pose_fovy)
    c = (np.asarray(vol_lo, np.float64) + np.asarray(vol_hi, np.float64)) / 2
    e = base["org"].astype(np.float64) - c
    th = np.deg2rad(theta_deg)
    eye = c + np.array([e[0] * np.cos(th) + e[2] * np.sin(th), e[1], -e[0] * np.sin(th) + e[2] * np.cos(th)])
    return camera_orient(eye.astype(np.float32), c.astype(np.float32), (0, 1, 0), width, height, aspect, fovy)


DEFAULT_LUT5 = np.array(  # icon_rt/hostCode.cu:828-834
    [[0.149, 0.015, 0.705, 1.0], [0.486, 0.603, 0.956, 0.75], [0.866, 0.866, 0.866, 0.5],
     [0.996, 0.690, 0.552, 0.25], [0.752, 0.298, 0.231, 0.0]], np.float32)


def load_ic(path, max_num_cells=-1, lat_range=(-90.0, 90.0), lon_range=(-180.0, 180.0)):
    """`.ic` loader with the reference's filtering (hostCode.cu:717-758; ranges in degrees): the surviving records"""
    L = host_lib()
    L.icb_load_ic.restype = C.c_int64
    args = (str(path).encode(), C.c_int64(max_num_cells), C.c_float(lat_range[0]), C.c_float(lat_range[1]), C.c_float(lon_range[0]),
            C.c_float(lon_range[1]))
    n = L.icb_load_ic(*args, None, C.c_int64(0))
    if n < 0:
        raise RuntimeError(f"icb_load_ic({path}) failed: {n}")
    cells = np.empty(n, CELL_DTYPE)
    got = L.icb_load_ic(*args, C.c_void_p(cells.ctypes.data), C.c_int64(n))
    if got != n:
        raise RuntimeError(f"icb_load_ic({path}): {got} records on the second pass, {n} on the first")
    return cells


def resample_lut(src, ndst):
    """common/dvr_course-common.h:44-70 through the C++ host mirror."""
    src = np.ascontiguousarray(src, np.float32).reshape(-1, 4)
    dst = np.empty((ndst, 4), np.float32)
    host_lib().icb_resample_lut(_fp(src), len(src), _fp(dst), int(ndst))
    return dst


def smooth_lut(n=300):
    """The reference's default 5-point LUT resampled to 300 entries (pipeline.cu:469-473)."""
    return resample_lut(DEFAULT_LUT5, n)


def sharp_lut(n=300):
    """SURVEY.md §8d 'sharp' TF: rgb ramp of the default LUT, alpha 0 below the middle entry, 1 above."""
    lut = resample_lut(DEFAULT_LUT5, n)
    lut[:, 3] = 0.0
    lut[n // 2:, 3] = 1.0
    return lut


class Transfunc:
    """common/transfunc.h:29-49"""

    def __init__(self, lut=None, value_range=(0.0, 1.0), opacity=1.0):
        self.opacity = float(opacity)
        self.value_range = (float(value_range[0]), float(value_range[1]))
        self.rel_range = (0.0, 1.0)
        self.lut = smooth_lut() if lut is None else np.asarray(lut, np.float32).reshape(-1, 4)

    def set_lut(self, lut):
        self.lut = np.asarray(lut, np.float32).reshape(-1, 4)

    @property
    def size(self):
        return len(self.lut)
