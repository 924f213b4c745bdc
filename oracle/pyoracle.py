"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/libignis_oracle.so (see oracle/oracle.h).

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libignis_oracle.so")


class Counters(C.Structure):
    _fields_ = [("scene_nodes", C.c_uint64), ("prim_nodes", C.c_uint64), ("tri_packets", C.c_uint64),
                ("tri_tests", C.c_uint64), ("entities", C.c_uint64)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        lib.igo_scene_create.restype = vp
        lib.igo_scene_create.argtypes = [vp]
        lib.igo_scene_destroy.argtypes = [vp]
        lib.igo_trace.argtypes = [vp, vp, C.c_uint64, C.c_uint32, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
        lib.igo_brute.argtypes = [vp, vp, C.c_uint64, C.c_uint32, C.c_int, vp, vp, vp, vp, vp, vp, vp]
        lib.igo_normalize_rays.argtypes = [vp, C.c_uint64]
        lib.igo_kat_ray_triangle.argtypes = [vp, vp, vp, C.c_float, C.c_float, vp]
        lib.igo_kat_ray_box.argtypes = [vp, vp, vp, C.c_float, C.c_float, vp]
        lib.igo_scene_info.argtypes = [vp, vp, vp, vp]
        lib.igo_render.argtypes = [vp, vp, vp, vp, vp, C.c_int]
        lib.igo_render.restype = C.c_int
        lib.igo_render_aov.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, C.c_int]
        lib.igo_render_aov.restype = C.c_int
        lib.igo_kat_rng.argtypes = [C.c_int32] * 6 + [C.c_uint32, vp, vp, vp]
        _lib = lib
    return _lib


class OracleScene:
    def __init__(self, tables):
        self._lib = load()
        self._tables = tables  # keep the blobs alive
        sc = tables.c_scene()
        self._h = self._lib.igo_scene_create(C.addressof(sc))
        if not self._h:
            raise ValueError("oracle rejected the scene tables")

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.igo_scene_destroy(self._h)
            self._h = None

    @staticmethod
    def normalized(rays):
        r = np.array(rays, copy=True)
        load().igo_normalize_rays(r.ctypes.data, r.size)
        return r

    def trace(self, rays, ray_flags=0, any_hit=False, threads=0, normalize=True):
        r = self.normalized(rays) if normalize else np.ascontiguousarray(rays)
        n = r.size
        ent, prim = np.empty(n, np.int32), np.empty(n, np.int32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        cnt = Counters()
        self._lib.igo_trace(self._h, r.ctypes.data, n, ray_flags, int(any_hit), threads, ent.ctypes.data,
                            prim.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data, C.addressof(cnt))
        self.counters = {k: getattr(cnt, k) for k, _ in Counters._fields_}
        return ent, prim, t, u, v

    def brute(self, rays, ray_flags=0, threads=0, normalize=True):
        r = self.normalized(rays) if normalize else np.ascontiguousarray(rays)
        n = r.size
        ent, prim = np.empty(n, np.int32), np.empty(n, np.int32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        tie, second = np.empty(n, np.uint8), np.empty(n, np.float32)
        self._lib.igo_brute(self._h, r.ctypes.data, n, ray_flags, threads, ent.ctypes.data, prim.ctypes.data,
                            t.ctypes.data, u.ctypes.data, v.ctypes.data, tie.ctypes.data, second.ctypes.data)
        return ent, prim, t, u, v, tie.astype(bool), second

    def render(self, shading, width, height, spi, iterations=1, seed=0, frame=0, rays=None, threads=0, film=None,
               first_iteration=0, aovs=None):
        """Restated Device::render: adds `iterations` iterations into a (h, w, 3) f32 film of RGB SUMS.
        `shading` is an ignis-masterthesis_b200.shading.Shading; rays != None selects igtrace (list) mode.
        aovs: list of up to three (h, w, 3) f32 arrays the technique's AOVs are added into (igo_render_aov)."""
        from importlib import import_module
        S = import_module("ignis-masterthesis_b200").shading
        if rays is not None:
            rays = np.ascontiguousarray(rays)
            width, height = rays.size, 1
        if film is None:
            film = np.zeros((height, width, 3), np.float32)
        sh = shading.c_shading()
        stats = np.zeros(3, np.uint64)
        self.ray_stats = np.zeros(3, np.uint64)
        for it in range(first_iteration, first_iteration + iterations):
            rs = S.RenderSettings(rays=rays.ctypes.data if rays is not None else None,
                                  ray_count=rays.size if rays is not None else 0, spi=spi, width=width, height=height,
                                  iteration=it, frame=frame, user_seed=seed)
            a = [x.ctypes.data if x is not None else None for x in (list(aovs or []) + [None] * 3)[:3]]
            rc = self._lib.igo_render_aov(self._h, C.addressof(sh), C.addressof(rs), film.ctypes.data, a[0], a[1], a[2],
                                          stats.ctypes.data, threads)
            if rc != 0:
                raise ValueError("oracle rejected the shading tables / settings")
            self.ray_stats += stats
        return film

    def info(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._lib.igo_scene_info(self._h, C.addressof(a), C.addressof(b), C.addressof(c))
        return {"node4": a.value, "tri4": b.value, "tris": c.value}


def tonemap(image_sum, iterations, method=0, use_gamma=True, scale=1.0, exposure_factor=1.0, exposure_offset=0.0):
    """Restated Device::tonemap of an (h, w, 3) f32 image of sums -> (h, w) u32 packed ARGB."""
    img = np.ascontiguousarray(image_sum, np.float32)
    out = np.empty(img.shape[:2], np.uint32)
    lib = load()
    lib.igo_tonemap.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_float, C.c_float, C.c_float, C.c_void_p]
    lib.igo_tonemap.restype = None
    inv = np.float32(1.0) / np.float32(iterations) if iterations > 0 else np.float32(0)
    lib.igo_tonemap(img.ctypes.data, img.shape[0] * img.shape[1], method, 1 if use_gamma else 0, float(np.float32(scale) * inv),
                    exposure_factor, exposure_offset, out.ctypes.data)
    return out


def texture_eval(shading, tex_id, uv):
    """rgb of texture `tex_id` of an ignis-masterthesis_b200 Shading at the (n, 2) coordinates `uv`."""
    uv = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    out = np.zeros((uv.shape[0], 3), np.float32)
    sh = shading.c_shading()
    lib = load()
    lib.igo_texture_eval.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
    if lib.igo_texture_eval(C.addressof(sh), tex_id, uv.ctypes.data, uv.shape[0], out.ctypes.data) != 0:
        raise ValueError("texture id out of range")
    return out


def kat_rng(sample, it, frame, x, y, user, n=8):
    seed = C.c_uint32()
    raw, f32 = np.zeros(n, np.uint32), np.zeros(n, np.float32)
    load().igo_kat_rng(sample, it, frame, x, y, user, n, C.addressof(seed), raw.ctypes.data, f32.ctypes.data)
    return seed.value, raw, f32


def kat_ray_triangle(tri, org, direction, tmin=0.0, tmax=3.4028234e38):
    tri = np.ascontiguousarray(tri, np.float32).reshape(9)
    org = np.ascontiguousarray(org, np.float32)
    d = np.ascontiguousarray(direction, np.float32)
    out = np.zeros(3, np.float32)
    hit = load().igo_kat_ray_triangle(tri.ctypes.data, org.ctypes.data, d.ctypes.data, tmin, tmax, out.ctypes.data)
    return (bool(hit), out)


def kat_ray_box(box, org, direction, tmin=0.0, tmax=3.4028234e38):
    box = np.ascontiguousarray(box, np.float32).reshape(6)
    org = np.ascontiguousarray(org, np.float32)
    d = np.ascontiguousarray(direction, np.float32)
    out = np.zeros(1, np.float32)
    hit = load().igo_kat_ray_box(box.ctypes.data, org.ctypes.data, d.ctypes.data, tmin, tmax, out.ctypes.data)
    return (bool(hit), float(out[0]))
