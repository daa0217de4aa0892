"""ctypes bindings of the two CPU checkers under oracle/ (TEST INFRASTRUCTURE).

``Port``  -> oracle/libicon_oracle.so  (our plain-C restatement, always buildable)
``Ref``   -> oracle/_ref/libicon_ref.so (the reference's own sources compiled in place;
             present only where /root/reference was available at build time)

Both expose the same Python surface so tests can run the same checks against either.
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "libicon_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libicon_ref.so")

CELL_DTYPE = np.dtype(
    [("lat", "<f4", 3), ("lon", "<f4", 3), ("numLayers", "<i4"), ("height", "<f4", 32), ("value", "<f4", 32)]
)
assert CELL_DTYPE.itemsize == 284

FOVY_DEFAULT = np.float32(90.0 * np.pi / 180.0)  # camera.h:108


def build(target="all"):
    subprocess.run(["make", "-C", ORACLE_DIR, target], check=True, capture_output=True)


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t))


class _Scene(C.Structure):
    _fields_ = [
        ("cells", C.c_void_p), ("numCells", C.c_int),
        ("lut", C.POINTER(C.c_float)), ("lutSize", C.c_int),
        ("tfLo", C.c_float), ("tfHi", C.c_float), ("opacityScale", C.c_float),
        ("org", C.c_float * 3), ("dir00", C.c_float * 3), ("dirDu", C.c_float * 3), ("dirDv", C.c_float * 3),
        ("ambientColor", C.c_float * 3), ("ambientRadiance", C.c_float), ("unitDistance", C.c_float),
        ("dt", C.c_float), ("termAlpha", C.c_float), ("rInner", C.c_float), ("rOuter", C.c_float),
        ("jitter", C.c_int), ("accumID", C.c_int),
        ("prismVtx", C.POINTER(C.c_float)), ("prismGeom", C.POINTER(C.c_float)),
        ("chunk", C.c_int),
    ]


class Scene:
    """Everything a frame needs, in reference terms (LaunchParams, Params.h:92-119)."""

    def __init__(self, cells, lut, tf_range=(0.0, 1.0), opacity_scale=1.0, camera=None, unit_distance=1000.0,
                 dt=None, term_alpha=0.99, r_inner=None, r_outer=None, jitter=0, accum_id=0,
                 ambient_color=(1.0, 1.0, 1.0), ambient_radiance=1.0, prism=None, chunk=0):
        self.cells = np.ascontiguousarray(cells)
        self.chunk = int(chunk)  # > 0: chunked compositing (what the CUDA marcher computes); 0: the plain A.3 statement
        self.prism = prism  # (vtx, geom) from Port.prism_build: the prism sampler instead of getValue
        self.lut = _f(lut).reshape(-1, 4)
        self.tf_range = (float(tf_range[0]), float(tf_range[1]))
        self.opacity_scale = float(opacity_scale)
        self.camera = camera  # dict(org, d00, du, dv)
        self.unit_distance = float(unit_distance)
        self.dt = float(dt) if dt is not None else self.unit_distance / 2
        self.term_alpha = float(term_alpha)
        self.r_inner = float(r_inner)
        self.r_outer = float(r_outer)
        self.jitter = int(jitter)
        self.accum_id = int(accum_id)
        self.ambient_color = tuple(float(x) for x in ambient_color)
        self.ambient_radiance = float(ambient_radiance)


class _Base:
    prefix = ""

    def __init__(self, path):
        self.lib = C.CDLL(path)
        self.path = path

    def fn(self, name, restype=None):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        return f

    # ---- building blocks ---------------------------------------------------
    def sizeof_cell(self):
        return self.fn("sizeof_cell", C.c_int)()

    def to_cartesian(self, sph):
        sph = _f(sph).reshape(-1, 3)
        out = np.empty_like(sph)
        self.fn("to_cartesian")(_p(sph), C.c_int(len(sph)), _p(out))
        return out

    def cell_planes(self, cells):
        cells = np.ascontiguousarray(cells)
        out = np.empty((len(cells), 3, 4), np.float32)
        self.fn("cell_planes")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(out))
        return out

    def find_height(self, cell, r):
        cell = np.ascontiguousarray(cell)
        r = _f(r)
        out = np.empty(len(r), np.int32)
        name = "find_height_n" if self.prefix == "orc_" else "find_height"
        self.fn(name)(C.c_void_p(cell.ctypes.data), _p(r), C.c_int(len(r)), _p(out, C.c_int32))
        return out

    def post_classify(self, lut, lo, hi, opacity_scale, v):
        lut = _f(lut).reshape(-1, 4)
        v = _f(v)
        out = np.empty((len(v), 4), np.float32)
        self.fn("post_classify")(_p(lut), C.c_int(len(lut)), C.c_float(lo), C.c_float(hi), C.c_float(opacity_scale),
                                 _p(v), C.c_int(len(v)), _p(out))
        return out

    def intersect_sphere(self, org, dirs, radius):
        org = _f(org).reshape(-1, 3)
        dirs = _f(dirs).reshape(-1, 3)
        n = len(org)
        hit = np.empty(n, np.int32)
        tn = np.empty(n, np.float32)
        tf = np.empty(n, np.float32)
        self.fn("intersect_sphere")(_p(org), _p(dirs), C.c_float(radius), C.c_int(n), _p(hit, C.c_int32), _p(tn), _p(tf))
        return hit, tn, tf

    def lcg(self, v0, v1, n):
        out = np.empty(n, np.float32)
        self.fn("lcg")(C.c_uint(v0), C.c_uint(v1), C.c_int(n), _p(out))
        return out

    def shell_segments(self, org, d, tmin, tmax, r_inner, r_outer):
        org = _f(org)
        d = _f(d)
        seg = np.zeros(4, np.float32)
        n = self.fn("shell_segments", C.c_int)(_p(org), _p(d), C.c_float(tmin), C.c_float(tmax),
                                                 C.c_float(r_inner), C.c_float(r_outer), _p(seg))
        return seg.reshape(2, 2)[:n].copy()

    def srgb_pack(self, rgba):
        rgba = _f(rgba).reshape(-1, 4)
        out = np.empty(len(rgba), np.uint32)
        self.fn("srgb_pack")(_p(rgba), C.c_int(len(rgba)), _p(out, C.c_uint32))
        return out

    def grid_bounds(self, cells):
        cells = np.ascontiguousarray(cells)
        a = [np.zeros(3, np.float32) for _ in range(4)]
        dr = np.zeros(2, np.float32)
        self.fn("grid_bounds")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(dr))
        return dict(volLo=a[0], volHi=a[1], sphLo=a[2], sphHi=a[3], dataRange=dr)

    def camera_view_all(self, vol_lo, vol_hi, aspect, W, H, fovy=FOVY_DEFAULT):
        o = [np.zeros(3, np.float32) for _ in range(4)]
        self.fn("camera_view_all")(_p(_f(vol_lo)), _p(_f(vol_hi)), C.c_float(fovy), C.c_float(aspect), C.c_int(W), C.c_int(H),
                                   _p(o[0]), _p(o[1]), _p(o[2]), _p(o[3]))
        return dict(org=o[0], d00=o[1], du=o[2], dv=o[3])

    def camera_orient(self, eye, poi, up, fovy, aspect, W, H):
        o = [np.zeros(3, np.float32) for _ in range(4)]
        self.fn("camera_orient")(_p(_f(eye)), _p(_f(poi)), _p(_f(up)), C.c_float(fovy), C.c_float(aspect), C.c_int(W), C.c_int(H),
                                 _p(o[0]), _p(o[1]), _p(o[2]), _p(o[3]))
        return dict(org=o[0], d00=o[1], du=o[2], dv=o[3])

    def resample_lut(self, src, ndst):
        src = _f(src).reshape(-1, 4)
        dst = np.empty((ndst, 4), np.float32)
        self.fn("resample_lut")(_p(src), C.c_int(len(src)), _p(dst), C.c_int(ndst))
        return dst


class Port(_Base):
    prefix = "orc_"

    def __init__(self):
        if not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(ORACLE_DIR, "icon_oracle.c")):
            build("port")
        super().__init__(PORT_SO)

    def num_threads(self):
        return self.fn("num_threads", C.c_int)()

    def locate(self, cells, pos):
        cells = np.ascontiguousarray(cells)
        pos = _f(pos).reshape(-1, 3)
        ids = np.empty(len(pos), np.int32)
        vals = np.empty(len(pos), np.float32)
        self.fn("locate")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(pos), C.c_int(len(pos)), _p(ids, C.c_int32), _p(vals))
        return ids, vals

    def _scene(self, sc):
        s = _Scene()
        s.cells = sc.cells.ctypes.data
        s.numCells = len(sc.cells)
        s.lut = _p(sc.lut)
        s.lutSize = len(sc.lut)
        s.tfLo, s.tfHi = sc.tf_range
        s.opacityScale = sc.opacity_scale
        for k, name in (("org", "org"), ("d00", "dir00"), ("du", "dirDu"), ("dv", "dirDv")):
            setattr(s, name, (C.c_float * 3)(*[float(x) for x in sc.camera[k]]))
        s.ambientColor = (C.c_float * 3)(*sc.ambient_color)
        s.ambientRadiance = sc.ambient_radiance
        s.unitDistance = sc.unit_distance
        s.dt, s.termAlpha, s.rInner, s.rOuter = sc.dt, sc.term_alpha, sc.r_inner, sc.r_outer
        s.jitter, s.accumID = sc.jitter, sc.accum_id
        s.chunk = getattr(sc, "chunk", 0)
        if getattr(sc, "prism", None) is not None:
            s.prismVtx, s.prismGeom = _p(sc.prism[0]), _p(sc.prism[1])
        return s

    # ---- prism sampler ---------------------------------
    def prism_build(self, cells):
        cells = np.ascontiguousarray(cells)
        vtx = np.zeros((len(cells), 32, 3), np.float32)
        geom = np.zeros((len(cells), 6), np.float32)
        self.fn("prism_build")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(vtx), _p(geom))
        return vtx, geom

    def locate_prism(self, cells, prism, pos):
        cells = np.ascontiguousarray(cells)
        pos = _f(pos).reshape(-1, 3)
        n = len(pos)
        ids, layers = np.empty(n, np.int32), np.empty(n, np.int32)
        vals, pc = np.empty(n, np.float32), np.empty((n, 3), np.float32)
        self.fn("locate_prism")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(prism[0]), _p(prism[1]), _p(pos), C.c_int(n),
                                _p(ids, C.c_int32), _p(vals), _p(layers, C.c_int32), _p(pc))
        return ids, vals, layers, pc

    def prism_wedges(self, cells, prism, ids, layers):
        """six float4 vertices (xyz + scalar) of the located prisms, in buildCuBQLAccel's order"""
        cells = np.ascontiguousarray(cells)
        out = np.zeros((len(ids), 6, 4), np.float32)
        f = self.fn("prism_wedge")
        for i, (rid, l) in enumerate(zip(ids, layers)):
            if rid < 0 or l < 0 or l >= cells[rid]["numLayers"]:
                continue
            f(C.c_void_p(cells[rid:rid + 1].ctypes.data), _p(prism[0][rid]), C.c_int(int(l)), _p(out[i]))
        return out

    def generate_rays(self, sc, xy, W, H):
        xy = np.ascontiguousarray(xy, np.int32).reshape(-1, 2)
        org = np.empty((len(xy), 3), np.float32)
        d = np.empty((len(xy), 3), np.float32)
        s = self._scene(sc)
        self.fn("generate_rays")(C.byref(s), _p(xy, C.c_int32), C.c_int(len(xy)), C.c_int(W), C.c_int(H), _p(org), _p(d))
        return org, d

    def march_frame(self, sc, W, H, rect=None, fb=None, accum=None, deadline_s=0.0, nthreads=0):
        x0, y0, x1, y1 = rect if rect else (0, 0, W, H)
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        pp = np.zeros((W * H, 4), np.int32)
        tot = np.zeros(3, np.float64)
        s = self._scene(sc)
        n = self.fn("march_frame", C.c_long)(C.byref(s), C.c_int(W), C.c_int(H), C.c_int(x0), C.c_int(y0), C.c_int(x1), C.c_int(y1),
                                              _p(fb, C.c_uint32), _p(accum), _p(pp, C.c_int32), C.c_double(deadline_s),
                                              _p(tot, C.c_double))
        return dict(fb=fb, accum=accum, perPixel=pp, pixels=n, steps=int(tot[0]), hits=int(tot[1]), seconds=float(tot[2]))

    def march_trace(self, sc, W, H, x, y, max_ids=65536):
        ids = np.full(max_ids, -2, np.int32)
        rgba = np.zeros(4, np.float32)
        s = self._scene(sc)
        n = self.fn("march_trace", C.c_int)(C.byref(s), C.c_int(W), C.c_int(H), C.c_int(x), C.c_int(y),
                                             _p(ids, C.c_int32), C.c_int(max_ids), _p(rgba))
        return ids[: min(n, max_ids)].copy(), n, rgba

    def woodcock_frame(self, sc, W, H, vol_lo, vol_hi, fb=None, accum=None):
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        s = self._scene(sc)
        self.fn("woodcock_frame")(C.byref(s), C.c_int(W), C.c_int(H), _p(_f(vol_lo)), _p(_f(vol_hi)), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum)


    # ---- macrocell majorants + the reference's default ray-gen ----
    def grid_ranges(self, cells, dims, lo, hi):
        cells = np.ascontiguousarray(cells)
        dims = np.asarray(dims, np.int32)
        out = np.empty((int(np.prod(dims)), 2), np.float32)
        self.fn("grid_ranges")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(dims, C.c_int32), _p(_f(lo)), _p(_f(hi)), _p(out))
        return out

    def max_opacities(self, ranges, lut, tf_range=(0.0, 1.0)):
        lut = _f(lut).reshape(-1, 4)
        ranges = _f(ranges)
        out = np.empty(len(ranges), np.float32)
        self.fn("max_opacities")(_p(ranges), C.c_long(len(ranges)), _p(lut), C.c_int(len(lut)), C.c_float(tf_range[0]),
                                 C.c_float(tf_range[1]), _p(out))
        return out

    def woodcock_grid_frame(self, sc, W, H, vol_lo, vol_hi, dims, max_op, fb=None, accum=None):
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        s = self._scene(sc)
        dims = np.asarray(dims, np.int32)
        self.fn("woodcock_grid_frame")(C.byref(s), C.c_int(W), C.c_int(H), _p(_f(vol_lo)), _p(_f(vol_hi)), _p(dims, C.c_int32),
                                       _p(_f(vol_lo)), _p(_f(vol_hi)), _p(_f(max_op)), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum)

    def shell_ranges(self, cells, dims, sph_lo, sph_hi):
        cells = np.ascontiguousarray(cells)
        dims = np.asarray(dims, np.int32)
        out = np.empty((int(np.prod(dims)), 2), np.float32)
        self.fn("shell_ranges")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(dims, C.c_int32), _p(_f(sph_lo)), _p(_f(sph_hi)), _p(out))
        return out

    def woodcock_sphere_frame(self, sc, W, H, vol_lo, vol_hi, dims, sph_lo, sph_hi, max_op, fb=None, accum=None):
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        s = self._scene(sc)
        dims = np.asarray(dims, np.int32)
        self.fn("woodcock_sphere_frame")(C.byref(s), C.c_int(W), C.c_int(H), _p(_f(vol_lo)), _p(_f(vol_hi)), _p(dims, C.c_int32),
                                         _p(_f(sph_lo)), _p(_f(sph_hi)), _p(_f(max_op)), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum)


class Ref(_Base):
    prefix = "ref_"

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        super().__init__(REF_SO)

    def num_threads(self):
        return self.fn("hardware_concurrency", C.c_int)()

    def locate(self, cells, pos, nthreads=0):
        cells = np.ascontiguousarray(cells)
        pos = _f(pos).reshape(-1, 3)
        ids = np.empty(len(pos), np.int32)
        vals = np.empty(len(pos), np.float32)
        self.fn("locate")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(pos), C.c_int(len(pos)),
                          _p(ids, C.c_int32), _p(vals), C.c_int(nthreads))
        return ids, vals

    def _set(self, sc, vol_lo=(0, 0, 0), vol_hi=(0, 0, 0)):
        self._keep = (sc.cells, sc.lut)
        self.fn("set_scene")(C.c_void_p(sc.cells.ctypes.data), C.c_int(len(sc.cells)), _p(_f(vol_lo)), _p(_f(vol_hi)),
                             _p(sc.lut), C.c_int(len(sc.lut)), C.c_float(sc.tf_range[0]), C.c_float(sc.tf_range[1]),
                             C.c_float(sc.opacity_scale), C.c_float(sc.unit_distance),
                             _p(_f(sc.ambient_color)), C.c_float(sc.ambient_radiance))
        self.fn("set_chunk")(C.c_int(getattr(sc, "chunk", 0)))
        cam = sc.camera
        self.fn("set_camera")(_p(_f(cam["org"])), _p(_f(cam["d00"])), _p(_f(cam["du"])), _p(_f(cam["dv"])))

    def generate_rays(self, sc, xy, W, H):
        xy = np.ascontiguousarray(xy, np.int32).reshape(-1, 2)
        org = np.empty((len(xy), 3), np.float32)
        d = np.empty((len(xy), 3), np.float32)
        self._set(sc)
        self.fn("generate_rays")(_p(xy, C.c_int32), C.c_int(len(xy)), C.c_int(W), C.c_int(H), C.c_int(sc.accum_id),
                                 C.c_int(sc.jitter), _p(org), _p(d))
        return org, d

    def sdda_callbacks(self, org, d, tmin, tmax, sph_lo, sph_hi, dims=(1, 1024, 1024), max_cb=64):
        t = np.zeros((max_cb, 2), np.float32)
        leaf = np.zeros(max_cb, np.int32)
        n = self.fn("sdda_callbacks", C.c_int)(_p(_f(org)), _p(_f(d)), C.c_float(tmin), C.c_float(tmax), _p(_f(sph_lo)), _p(_f(sph_hi)),
                                                _p(np.asarray(dims, np.int32), C.c_int32), C.c_int(max_cb), _p(t), _p(leaf, C.c_int32))
        return t[: min(n, max_cb)], leaf[: min(n, max_cb)], n

    def march_frame(self, sc, W, H, rect=None, fb=None, accum=None, deadline_s=0.0, nthreads=0):
        x0, y0, x1, y1 = rect if rect else (0, 0, W, H)
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        pp = np.zeros((W * H, 4), np.int32)
        tot = np.zeros(3, np.float64)
        self._set(sc)
        n = self.fn("march_frame", C.c_long)(C.c_int(W), C.c_int(H), C.c_int(x0), C.c_int(y0), C.c_int(x1), C.c_int(y1),
                                              C.c_int(sc.accum_id), _p(fb, C.c_uint32), _p(accum), C.c_int(sc.jitter),
                                              C.c_float(sc.dt), C.c_float(sc.term_alpha), C.c_float(sc.r_inner), C.c_float(sc.r_outer),
                                              _p(pp, C.c_int32), C.c_int(nthreads), C.c_double(deadline_s), _p(tot, C.c_double))
        return dict(fb=fb, accum=accum, perPixel=pp, pixels=n, steps=int(tot[0]), hits=int(tot[1]), seconds=float(tot[2]))

    def march_trace(self, sc, W, H, x, y, max_ids=65536):
        ids = np.full(max_ids, -2, np.int32)
        rgba = np.zeros(4, np.float32)
        self._set(sc)
        n = self.fn("march_trace", C.c_int)(C.c_int(W), C.c_int(H), C.c_int(x), C.c_int(y), C.c_int(sc.accum_id), C.c_int(sc.jitter),
                                             C.c_float(sc.dt), C.c_float(sc.term_alpha), C.c_float(sc.r_inner), C.c_float(sc.r_outer),
                                             _p(ids, C.c_int32), C.c_int(max_ids), _p(rgba))
        return ids[: min(n, max_ids)].copy(), n, rgba

    def woodcock_grid_frame(self, sc, W, H, vol_lo, vol_hi, dims, max_op):
        """one frame of the reference's default ray-gen woodcockTrackingWithAccel (GRID_ACCEL_MODE, grid over the volume
        bounds as hostCode.cu:668-670); returns accumBuffer * (accumID+1) like woodcock_frame"""
        fb = np.zeros(W * H, np.uint32)
        accum = np.zeros((W * H, 4), np.float32)
        self._set(sc, vol_lo, vol_hi)
        dims = np.asarray(dims, np.int32)
        mo = _f(max_op)
        sec = self.fn("woodcock_grid_frame", C.c_double)(C.c_int(W), C.c_int(H), C.c_int(sc.accum_id), _p(dims, C.c_int32),
                                                         _p(_f(vol_lo)), _p(_f(vol_hi)), _p(mo), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum * np.float32(sc.accum_id + 1), seconds=sec)

    def woodcock_sphere_frame(self, sc, W, H, vol_lo, vol_hi, dims, sph_lo, sph_hi, max_op):
        """one frame of woodcockTrackingWithAccel in SPHERE_ACCEL_MODE (sdda over volume.accel, the mode the reference starts
        in); returns accumBuffer * (accumID+1) like woodcock_frame"""
        fb = np.zeros(W * H, np.uint32)
        accum = np.zeros((W * H, 4), np.float32)
        self._set(sc, vol_lo, vol_hi)
        dims = np.asarray(dims, np.int32)
        sec = self.fn("woodcock_sphere_frame", C.c_double)(C.c_int(W), C.c_int(H), C.c_int(sc.accum_id), _p(dims, C.c_int32),
                                                           _p(_f(sph_lo)), _p(_f(sph_hi)), _p(_f(max_op)), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum * np.float32(sc.accum_id + 1), seconds=sec)

    # ---- the reference's own majorant builders (hostCode.cu kernels on host shims, oracle/ref_kernel_shims.h) ----
    def grid_ranges(self, cells, dims, lo, hi):
        """initGrid + buildGrid_ICON (hostCode.cu:205-297) as buildICONGrid launches them (:668-682)"""
        cells = np.ascontiguousarray(cells)
        dims = np.asarray(dims, np.int32)
        out = np.empty((int(np.prod(dims)), 2), np.float32)
        self.fn("grid_ranges")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(dims, C.c_int32), _p(_f(lo)), _p(_f(hi)), _p(out))
        return out

    def shell_ranges(self, cells, dims, sph_lo, sph_hi):
        """initGrid + buildShell_ICON (hostCode.cu:216-225, 299-336) as buildShellAccel launches them (:652-666)"""
        cells = np.ascontiguousarray(cells)
        dims = np.asarray(dims, np.int32)
        out = np.empty((int(np.prod(dims)), 2), np.float32)
        self.fn("shell_ranges")(C.c_void_p(cells.ctypes.data), C.c_int(len(cells)), _p(dims, C.c_int32), _p(_f(sph_lo)), _p(_f(sph_hi)), _p(out))
        return out

    def max_opacities(self, ranges, lut, tf_range=(0.0, 1.0), shell=False):
        """computeMaxOpacities (hostCode.cu:362-434), Grid or ShellAccel overload"""
        lut = _f(lut).reshape(-1, 4)
        ranges = _f(ranges)
        out = np.empty(len(ranges), np.float32)
        self.fn("max_opacities")(_p(ranges), C.c_long(len(ranges)), _p(lut), C.c_int(len(lut)), C.c_float(tf_range[0]),
                                 C.c_float(tf_range[1]), C.c_int(1 if shell else 0), _p(out))
        return out

    def wedge(self, pos, V):
        """the reference's intersectWedgeEXT (UElems.h:214-311) on (point, wedge) pairs"""
        pos = _f(pos).reshape(-1, 3)
        V = _f(V).reshape(-1, 24)
        acc = np.zeros(len(pos), np.int32)
        val = np.zeros(len(pos), np.float32)
        self.fn("wedge")(_p(pos), _p(V), C.c_int(len(pos)), _p(acc, C.c_int32), _p(val))
        return acc, val

    def woodcock_frame(self, sc, W, H, vol_lo, vol_hi):
        """one frame of the reference's woodcockTrackingAE; returns the un-normalised estimate
        (accumBuffer * (accumID+1)), i.e. colour/alpha of this frame's single collision per pixel"""
        fb = np.zeros(W * H, np.uint32)
        accum = np.zeros((W * H, 4), np.float32)
        self._set(sc, vol_lo, vol_hi)
        sec = self.fn("woodcock_frame", C.c_double)(C.c_int(W), C.c_int(H), C.c_int(sc.accum_id), _p(fb, C.c_uint32), _p(accum))
        return dict(fb=fb, accum=accum * np.float32(sc.accum_id + 1), seconds=sec)


DEFAULT_LUT5 = np.array(  # hostCode.cu:828-834
    [[0.149, 0.015, 0.705, 1.0], [0.486, 0.603, 0.956, 0.75], [0.866, 0.866, 0.866, 0.5],
     [0.996, 0.690, 0.552, 0.25], [0.752, 0.298, 0.231, 0.0]], np.float32)
