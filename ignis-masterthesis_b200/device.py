"""Host-side mirror of the reference's `IG::Device` (src/runtime/device/Device.h:14-74) over the C ABI.

Same lifetime rules as the reference: one scene per device, `assign_scene` before `trace`
(Runtime.cpp:478-487), ray directions are normalised on upload (Device.cpp:585-607), outputs are the
five hit columns of the primary stream (src/artic/driver/streams.art:80-100).
"""
import ctypes as C

import numpy as np

from . import _lib
from .tables import RAY_DTYPE, SceneTables


class Device:
    def __init__(self, device=0, acquire_stats=False):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        setup = _lib.Setup(device=device, acquire_stats=1 if acquire_stats else 0)
        _lib.check(self._lib.igb200_create(C.byref(setup), C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.igb200_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # Device::assignScene
    def assign_scene(self, tables: SceneTables):
        sc = tables.c_scene()
        _lib.check(self._lib.igb200_assign_scene(self._h, C.byref(sc)))

    # Runtime::trace(rays, out) for hit records; host numpy buffers in, host numpy buffers out.
    def trace(self, rays, ray_flags=0, any_hit=False):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        n = rays.size
        ent = np.empty(n, np.int32)
        prim = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        u = np.empty(n, np.float32)
        v = np.empty(n, np.float32)
        _lib.check(self._lib.igb200_trace_rays(self._h, rays.ctypes.data, n, ray_flags, 1 if any_hit else 0,
                                               ent.ctypes.data, prim.ctypes.data, t.ctypes.data, u.ctypes.data,
                                               v.ctypes.data))
        return ent, prim, t, u, v

    # Same call writing into caller-provided host arrays (e.g. pinned memory); no allocation per call.
    def trace_into(self, rays, ent, prim, t, u, v, ray_flags=0, any_hit=False):
        n = rays.size
        assert rays.dtype == RAY_DTYPE and rays.flags.c_contiguous
        for a in (ent, prim, t, u, v):
            assert a is None or (a.size == n and a.itemsize == 4 and a.flags.c_contiguous)
        p = lambda a: None if a is None else a.ctypes.data
        _lib.check(self._lib.igb200_trace_rays(self._h, rays.ctypes.data, n, ray_flags, 1 if any_hit else 0,
                                               p(ent), p(prim), p(t), p(u), p(v)))

    # Same call on device-resident buffers (raw device pointers, e.g. torch tensors' data_ptr()).
    def trace_device(self, d_rays, n, d_ent, d_prim, d_t, d_u, d_v, ray_flags=0, any_hit=False, stream=0):
        _lib.check(self._lib.igb200_trace_rays_device(self._h, d_rays, n, ray_flags, 1 if any_hit else 0,
                                                      d_ent, d_prim, d_t, d_u, d_v, stream))

    # Device::getStatistics
    def trace_stats(self):
        st = _lib.TraceStats()
        _lib.check(self._lib.igb200_get_trace_stats(self._h, C.byref(st)))
        return {"rays": st.rays, "nodes_visited": st.nodes_visited, "tris_tested": st.tris_tested,
                "entities_entered": st.entities_entered, "kernel_ms": st.kernel_ms}

    # ---- wavefront path tracing: Device::render / getFramebufferForHost (Device.h:53,63-66) ----
    def assign_shading(self, shading):
        sh = shading.c_shading()
        _lib.check(self._lib.igb200_assign_shading(self._h, C.addressof(sh)))

    def render(self, width=0, height=0, spi=1, iteration=0, frame=0, seed=0, rays=None, iterations=1):
        """`iterations` iterations (spi samples per pixel each) accumulated into the film, in one wavefront.
        rays != None: igtrace (list) mode."""
        from .shading import RenderSettings
        if rays is not None:
            rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        rs = RenderSettings(rays=rays.ctypes.data if rays is not None else None, ray_count=rays.size if rays is not None else 0,
                            spi=spi, width=width, height=height, iteration=iteration, frame=frame, user_seed=seed)
        if iterations == 1:
            _lib.check(self._lib.igb200_render(self._h, C.addressof(rs)))
        else:
            _lib.check(self._lib.igb200_render_iterations(self._h, C.addressof(rs), iterations))
        self._film = (rays.size, 1) if rays is not None else (width, height)

    def framebuffer(self, copy=True, name=None):
        """(film or AOV `name` (h, w, 3) f32 of RGB SUMS over iterations, iteration count of that image)."""
        ptr, it = C.c_void_p(), C.c_uint64()
        _lib.check(self._lib.igb200_framebuffer(self._h, name.encode() if name else None, 0, C.byref(ptr), C.byref(it)))
        st = self.render_stats()
        n = st["film"][0] * st["film"][1] * 3
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(n,)).reshape(st["film"][1], st["film"][0], 3)
        return (arr.copy() if copy else arr), it.value

    def framebuffer_device(self):
        ptr, it = C.c_void_p(), C.c_uint64()
        _lib.check(self._lib.igb200_framebuffer(self._h, None, 1, C.byref(ptr), C.byref(it)))
        return ptr.value, it.value

    def set_framebuffer(self, d_src, count, iterations):
        _lib.check(self._lib.igb200_set_framebuffer(self._h, d_src, count, iterations))

    def clear_framebuffer(self, name=None):
        """Device::clearAllFramebuffer, or Device::clearFramebuffer(name) when a name is given ("Color" = the film)."""
        if name is None:
            _lib.check(self._lib.igb200_clear_framebuffer(self._h))
        else:
            _lib.check(self._lib.igb200_clear_aov(self._h, name.encode()))

    def tonemap(self, aov=None, method=0, use_gamma=True, scale=1.0, exposure_factor=1.0, exposure_offset=0.0):
        """Device::tonemap: (h, w) u32 packed ARGB of the film or AOV divided by its iteration count."""
        from .shading import TonemapSettings
        ts = TonemapSettings(aov=aov.encode() if aov else None, method=method, use_gamma=1 if use_gamma else 0, scale=scale,
                             exposure_factor=exposure_factor, exposure_offset=exposure_offset)
        w, h = self._film
        out = np.empty((h, w), np.uint32)
        _lib.check(self._lib.igb200_tonemap(self._h, C.addressof(ts), out.ctypes.data))
        return out

    def resize(self, width, height):
        _lib.check(self._lib.igb200_resize(self._h, width, height))
        self._film = (width, height)

    def render_stats(self):
        from .shading import RenderStats
        st = RenderStats()
        _lib.check(self._lib.igb200_get_render_stats(self._h, C.addressof(st)))
        return {"camera_rays": st.camera_rays, "bounce_rays": st.bounce_rays, "shadow_rays": st.shadow_rays,
                "rounds": st.rounds, "kernel_launches": st.kernel_launches, "device_ms": st.device_ms,
                "film": getattr(self, "_film", (0, 0))}

    def bvh_info(self):
        nb, tb, lb, d = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32()
        _lib.check(self._lib.igb200_get_bvh_info(self._h, C.byref(nb), C.byref(tb), C.byref(lb), C.byref(d)))
        return {"node_bytes": nb.value, "tri_bytes": tb.value, "leaf_bytes": lb.value, "max_depth": d.value}
