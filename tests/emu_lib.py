"""ctypes binding of tests/emu/libicon_emu.so: the device locator text compiled for the host
(TEST INFRASTRUCTURE; lets the walk / exact-query logic and the host grid builder be checked
against the oracle without a GPU).  Never used by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "emu")
EMU_SO = os.path.join(EMU_DIR, "libicon_emu.so")


class RenderParams(C.Structure):  # icb_types.h
    _fields_ = [("org", C.c_float * 3), ("d00", C.c_float * 3), ("du", C.c_float * 3), ("dv", C.c_float * 3),
                ("ambient", C.c_float * 3), ("ambientRadiance", C.c_float),
                ("unitDistance", C.c_float), ("dt", C.c_float), ("termAlpha", C.c_float),
                ("rInner", C.c_float), ("rOuter", C.c_float),
                ("tfLo", C.c_float), ("tfHi", C.c_float), ("opacityScale", C.c_float),
                ("tfSize", C.c_int32), ("width", C.c_int32), ("height", C.c_int32), ("accumID", C.c_int32),
                ("flags", C.c_uint32), ("skipPlaneScale", C.c_float), ("skipRadSlack", C.c_float), ("maxSkip", C.c_int32), ("skipBack", C.c_int32), ("tilesX", C.c_int32), ("tilesY", C.c_int32), ("numTiles", C.c_int32), ("volLo", C.c_float * 3), ("volHi", C.c_float * 3), ("chunkLog2", C.c_int32)]


class Emu:
    def __init__(self):
        subprocess.run(["make", "-C", EMU_DIR], check=True, capture_output=True)
        self.lib = C.CDLL(EMU_SO)
        self.lib.emu_error.restype = C.c_char_p
        assert self.lib.emu_sizeof_params() == C.sizeof(RenderParams)

    def build(self, cells, margin_m=16.0, max_walk=8):
        cells = np.ascontiguousarray(cells)
        rc = self.lib.emu_build(C.c_void_p(cells.ctypes.data), C.c_size_t(len(cells)), C.c_float(margin_m), int(max_walk))
        if rc != 0:
            raise RuntimeError(f"emu_build {rc}: {self.lib.emu_error().decode()}")
        info = np.zeros(8, np.int64)
        self.lib.emu_info(info.ctypes.data_as(C.POINTER(C.c_int64)))
        return dict(zip(["nRec", "nCol", "nVert", "binN", "binEntries", "nGeneric", "boundaryEdges", "nGlobal"], info.tolist()))

    def set_log_table(self, on):
        """True: the Woodcock ray-gens read logf(1-u) from the host-built table (round 1's form); False (default): they compute it"""
        self.lib.emu_set_log_table(1 if on else 0)

    def set_sampler(self, name):
        self.lib.emu_set_sampler({"cell": 0, "prism": 1}[name])

    def locate(self, pos, flags=0, chunk=0):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        ids = np.empty(len(pos), np.int32)
        vals = np.empty(len(pos), np.float32)
        ctr = np.zeros(4, np.uint64)
        self.lib.emu_locate(pos.ctypes.data_as(C.POINTER(C.c_float)), C.c_size_t(len(pos)), ids.ctypes.data_as(C.POINTER(C.c_int32)),
                            vals.ctypes.data_as(C.POINTER(C.c_float)), C.c_uint(flags), int(chunk),
                            ctr.ctypes.data_as(C.POINTER(C.c_uint64)))
        return ids, vals, dict(steps=int(ctr[0]), hits=int(ctr[1]), slow=int(ctr[2]), exact=int(ctr[3]))

    def _params(self, sc, W, H):
        p = RenderParams()
        for k, name in (("org", "org"), ("d00", "d00"), ("du", "du"), ("dv", "dv")):
            setattr(p, name, (C.c_float * 3)(*[float(v) for v in sc.camera[k]]))
        p.ambient = (C.c_float * 3)(*sc.ambient_color)
        p.ambientRadiance = sc.ambient_radiance
        p.unitDistance, p.dt, p.termAlpha = sc.unit_distance, sc.dt, sc.term_alpha
        p.rInner, p.rOuter = sc.r_inner, sc.r_outer
        p.tfLo, p.tfHi = sc.tf_range
        p.opacityScale = sc.opacity_scale
        p.tfSize = len(sc.lut)
        p.width, p.height, p.accumID, p.flags = W, H, sc.accum_id, (1 if sc.jitter else 0)
        return p

    def woodcock_frame(self, sc, W, H, vol_lo, vol_hi, fb=None, accum=None):
        """one frame of the Woodcock integrator (device text on the host); buffers as the reference's Frame"""
        p = self._params(sc, W, H)
        p.volLo = (C.c_float * 3)(*[float(v) for v in vol_lo])
        p.volHi = (C.c_float * 3)(*[float(v) for v in vol_hi])
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        lut = np.ascontiguousarray(sc.lut, np.float32)
        ctr = np.zeros(3, np.uint64)
        self.lib.emu_woodcock_frame(C.byref(p), lut.ctypes.data_as(C.POINTER(C.c_float)), fb.ctypes.data_as(C.POINTER(C.c_uint32)),
                                    accum.ctypes.data_as(C.POINTER(C.c_float)), ctr.ctypes.data_as(C.POINTER(C.c_uint64)))
        return dict(fb=fb, accum=accum, steps=int(ctr[0]), hits=int(ctr[1]), rays=int(ctr[2]))

    def woodcock_grid_frame(self, sc, W, H, vol_lo, vol_hi, dims, max_op, fb=None, accum=None):
        """one frame of the reference's default ray-gen (device text on the host) over the given majorant grid"""
        p = self._params(sc, W, H)
        p.volLo = (C.c_float * 3)(*[float(v) for v in vol_lo])
        p.volHi = (C.c_float * 3)(*[float(v) for v in vol_hi])
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        lut = np.ascontiguousarray(sc.lut, np.float32)
        mo = np.ascontiguousarray(max_op, np.float32)
        d = np.asarray(dims, np.int32)
        ctr = np.zeros(3, np.uint64)
        self.lib.emu_woodcock_grid_frame(C.byref(p), lut.ctypes.data_as(C.POINTER(C.c_float)), d.ctypes.data_as(C.POINTER(C.c_int32)),
                                         mo.ctypes.data_as(C.POINTER(C.c_float)), fb.ctypes.data_as(C.POINTER(C.c_uint32)),
                                         accum.ctypes.data_as(C.POINTER(C.c_float)), ctr.ctypes.data_as(C.POINTER(C.c_uint64)))
        return dict(fb=fb, accum=accum, steps=int(ctr[0]), hits=int(ctr[1]), rays=int(ctr[2]))

    def woodcock_sphere_frame(self, sc, W, H, vol_lo, vol_hi, dims, sph_lo, sph_hi, max_op, fb=None, accum=None):
        """one frame of the reference's ray-gen in SPHERE_ACCEL_MODE (device text on the host) over the given shell accel"""
        p = self._params(sc, W, H)
        p.volLo = (C.c_float * 3)(*[float(v) for v in vol_lo])
        p.volHi = (C.c_float * 3)(*[float(v) for v in vol_hi])
        fb = np.zeros(W * H, np.uint32) if fb is None else fb
        accum = np.zeros((W * H, 4), np.float32) if accum is None else accum
        lut = np.ascontiguousarray(sc.lut, np.float32)
        mo = np.ascontiguousarray(max_op, np.float32)
        d = np.asarray(dims, np.int32)
        lo, hi = np.ascontiguousarray(sph_lo, np.float32), np.ascontiguousarray(sph_hi, np.float32)
        ctr = np.zeros(3, np.uint64)
        fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
        self.lib.emu_woodcock_sphere_frame(C.byref(p), fp(lut), d.ctypes.data_as(C.POINTER(C.c_int32)), fp(lo), fp(hi), fp(mo),
                                           fb.ctypes.data_as(C.POINTER(C.c_uint32)), fp(accum), ctr.ctypes.data_as(C.POINTER(C.c_uint64)))
        return dict(fb=fb, accum=accum, steps=int(ctr[0]), hits=int(ctr[1]), rays=int(ctr[2]))

    def march_pixel(self, sc, W, H, x, y, max_ids=65536, skip=False, margin_m=16.0, max_skip=1024, chunk_log2=10):
        """sc: oracle_lib.Scene; skip=True turns on the skip-ahead with the slack icb_api.cu derives"""
        p = RenderParams()
        p.chunkLog2 = chunk_log2
        if skip:
            M = max(float(np.abs(np.asarray(sc.camera["org"], np.float64)).max()), sc.r_outer)
            slack = 16.0 * M * 2.0 ** -23
            p.skipRadSlack = slack
            p.skipPlaneScale = 1.0 + slack / margin_m
            p.maxSkip = max_skip
        for k, name in (("org", "org"), ("d00", "d00"), ("du", "du"), ("dv", "dv")):
            setattr(p, name, (C.c_float * 3)(*[float(v) for v in sc.camera[k]]))
        p.ambient = (C.c_float * 3)(*sc.ambient_color)
        p.ambientRadiance = sc.ambient_radiance
        p.unitDistance, p.dt, p.termAlpha = sc.unit_distance, sc.dt, sc.term_alpha
        p.rInner, p.rOuter = sc.r_inner, sc.r_outer
        p.tfLo, p.tfHi = sc.tf_range
        p.opacityScale = sc.opacity_scale
        p.tfSize = len(sc.lut)
        p.width, p.height, p.accumID, p.flags = W, H, sc.accum_id, (1 if sc.jitter else 0)
        ids = np.full(max_ids, -2, np.int32)
        rgba = np.zeros(4, np.float32)
        lut = np.ascontiguousarray(sc.lut, np.float32)
        n = self.lib.emu_march_pixel(C.byref(p), lut.ctypes.data_as(C.POINTER(C.c_float)), int(x), int(y),
                                     ids.ctypes.data_as(C.POINTER(C.c_int32)), int(max_ids), rgba.ctypes.data_as(C.POINTER(C.c_float)),
                                     1 if skip else 0)
        return ids[: min(n, max_ids)].copy(), n, rgba
