"""Host-side mirror of the reference's `IG::Device` (src/runtime/device/Device.h:14-74) over the C ABI.

Same lifetime rules as the reference: one scene per device, `assign_scene` before `trace`
(Runtime.cpp:478-487), ray directions are normalised on upload (Device.cpp:585-607), outputs are the
five hit columns of the primary stream (src/artic/driver/streams.art:80-100).
"""
import ctypes as C

import numpy as np

from . import _lib
from .tables import RAY_DTYPE, SceneTables


PARTITION_AUTO, PARTITION_ITERATIONS, PARTITION_TILE_ROWS = 0, 1, 2


def partition_share(total_iterations, part, parts, mode=PARTITION_AUTO):
    """igb200_partition_share: (first iteration offset, iteration count, tile_rank, tile_count) of partition `part`."""
    lib = _lib.load()
    out = [C.c_uint32() for _ in range(4)]
    _lib.check(lib.igb200_partition_share(total_iterations, part, parts, mode, *[C.byref(o) for o in out]))
    return tuple(o.value for o in out)


class Device:
    def __init__(self, device=0, acquire_stats=False, _borrowed=None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self._owned = _borrowed is None
        if _borrowed is not None:  # a member of a DeviceGroup: the group owns the handle
            self._h = C.c_void_p(_borrowed)
            self.device = self._lib.igb200_device_ordinal(self._h)
            return
        setup = _lib.Setup(device=device, acquire_stats=1 if acquire_stats else 0)
        _lib.check(self._lib.igb200_create(C.byref(setup), C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            if self._owned:
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

    NOALLOC_MAX_RAYS = 65536  # IGB200_NOALLOC_MAX_RAYS

    # igb200_trace_rays_noalloc: same results, scratch in a static device array (no cudaMalloc on the call)
    def trace_noalloc_into(self, rays, ent, prim, t, u, v, ray_flags=0, any_hit=False):
        n = rays.size
        assert rays.dtype == RAY_DTYPE and rays.flags.c_contiguous
        p = lambda a: None if a is None else a.ctypes.data
        _lib.check(self._lib.igb200_trace_rays_noalloc(self._h, rays.ctypes.data if n else None, n, ray_flags, 1 if any_hit else 0,
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

    def render(self, width=0, height=0, spi=1, iteration=0, frame=0, seed=0, rays=None, iterations=1, pipelined=False):
        """`iterations` iterations (spi samples per pixel each) accumulated into the film, in one wavefront.
        rays != None: igtrace (list) mode.  pipelined: igb200_render_pipelined (one iteration; its tail continues in the
        next call, the film is complete after flush() or any call that reads it)."""
        rs, rays = self._settings(width, height, spi, iteration, frame, seed, rays)
        if pipelined:
            assert iterations == 1
            _lib.check(self._lib.igb200_render_pipelined(self._h, C.addressof(rs)))
        elif iterations == 1:
            _lib.check(self._lib.igb200_render(self._h, C.addressof(rs)))
        else:
            _lib.check(self._lib.igb200_render_iterations(self._h, C.addressof(rs), iterations))

    def flush(self):
        _lib.check(self._lib.igb200_flush(self._h))

    def _settings(self, width, height, spi, iteration, frame, seed, rays):
        from .shading import RenderSettings
        if rays is not None:
            rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        self._film = (rays.size, 1) if rays is not None else (width, height)
        return RenderSettings(rays=rays.ctypes.data if rays is not None else None, ray_count=rays.size if rays is not None else 0,
                              spi=spi, width=width, height=height, iteration=iteration, frame=frame, user_seed=seed), rays

    def render_partition(self, width, height, spi, total_iterations, part, parts, mode=PARTITION_AUTO, iteration=0, frame=0, seed=0):
        """igb200_render_partition: this device's share of a `total_iterations` request (partial film)."""
        rs, _ = self._settings(width, height, spi, iteration, frame, seed, None)
        _lib.check(self._lib.igb200_render_partition(self._h, C.addressof(rs), total_iterations, part, parts, mode))

    def set_iteration_count(self, iterations, name=None):
        _lib.check(self._lib.igb200_set_iteration_count(self._h, name.encode() if name else None, iterations))

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

    SECTION_NAMES = ("begin_round", "generate", "trace_primary", "scatter", "shade", "trace_shadow")

    def section_stats(self, reset=False):
        """igb200_get_section_stats: per-kernel-kind device ms and launch counts (devices created with acquire_stats)."""
        st = _lib.SectionStats()
        _lib.check(self._lib.igb200_get_section_stats(self._h, C.addressof(st), 1 if reset else 0))
        out = {n: {"ms": st.kernel_ms[i], "launches": st.launches[i]} for i, n in enumerate(self.SECTION_NAMES)}
        out["framebuffer_host_update"] = {"ms": st.framebuffer_host_update_ms, "launches": st.framebuffer_host_updates}
        return out

    def bvh_info(self):
        nb, tb, lb, d = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32()
        _lib.check(self._lib.igb200_get_bvh_info(self._h, C.byref(nb), C.byref(tb), C.byref(lb), C.byref(d)))
        return {"node_bytes": nb.value, "tri_bytes": tb.value, "leaf_bytes": lb.value, "max_depth": d.value}


class Comm:
    """One process per GPU (torchrun / MPI): igb200_comm_* -- an NCCL communicator next to a Device.
    `unique_id()` on rank 0, hand the 128 bytes to the other ranks, then Comm(dev, id, nranks, rank) everywhere."""

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        _lib.check(_lib.load().igb200_comm_unique_id(buf))
        return bytes(buf)

    def __init__(self, dev, uid, nranks, rank):
        self._lib, self.dev, self.nranks, self.rank = _lib.load(), dev, nranks, rank
        self._h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        _lib.check(self._lib.igb200_comm_create(dev._h, buf, nranks, rank, C.byref(self._h)))

    def render(self, width, height, spi, total_iterations, mode=PARTITION_AUTO, iteration=0, frame=0, seed=0):
        rs, _ = self.dev._settings(width, height, spi, iteration, frame, seed, None)
        _lib.check(self._lib.igb200_comm_render(self._h, C.addressof(rs), total_iterations, mode))

    def reduce_film(self, total_iterations, root=0, mode=PARTITION_AUTO, warm=True):
        """One ncclReduce(sum) per image onto `root`; returns the reduce time in ms (CUDA events)."""
        ms = C.c_float()
        _lib.check(self._lib.igb200_comm_reduce_film(self._h, root, total_iterations, mode, 1 if warm else 0, C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.igb200_comm_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close


class DeviceGroup:
    """One process, N devices (igb200_group_*): scene replicated, requests partitioned, films reduced onto member 0."""

    def __init__(self, devices, acquire_stats=False):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        devices = list(range(devices)) if isinstance(devices, int) else list(devices)
        arr = (C.c_int32 * len(devices))(*devices)
        _lib.check(self._lib.igb200_group_create(arr, len(devices), 1 if acquire_stats else 0, C.byref(self._h)))
        self.members = [Device(_borrowed=self._lib.igb200_group_device(self._h, i)) for i in range(len(devices))]
        self.root = self.members[0]

    def __len__(self):
        return len(self.members)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            for m in self.members:
                m.close()
            self._lib.igb200_group_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def assign_scene(self, tables):
        sc = tables.c_scene()
        _lib.check(self._lib.igb200_group_assign_scene(self._h, C.byref(sc)))

    def assign_shading(self, shading):
        sh = shading.c_shading()
        _lib.check(self._lib.igb200_group_assign_shading(self._h, C.addressof(sh)))

    def resize(self, width, height):
        _lib.check(self._lib.igb200_group_resize(self._h, width, height))
        for m in self.members:
            m._film = (width, height)

    def clear_framebuffer(self):
        _lib.check(self._lib.igb200_group_clear_framebuffer(self._h))

    def render(self, width, height, spi, iterations, mode=PARTITION_AUTO, iteration=0, frame=0, seed=0):
        rs, _ = self.root._settings(width, height, spi, iteration, frame, seed, None)
        for m in self.members:
            m._film = (width, height)
        _lib.check(self._lib.igb200_group_render(self._h, C.addressof(rs), iterations, mode))

    def framebuffer(self, copy=True, name=None):
        return self.root.framebuffer(copy=copy, name=name)

    def trace(self, rays, ray_flags=0, any_hit=False):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        n = rays.size
        ent, prim = np.empty(n, np.int32), np.empty(n, np.int32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        _lib.check(self._lib.igb200_group_trace_rays(self._h, rays.ctypes.data, n, ray_flags, 1 if any_hit else 0, ent.ctypes.data,
                                                     prim.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data))
        return ent, prim, t, u, v

    def render_stats(self):
        from .shading import RenderStats
        st, ms = RenderStats(), C.c_float()
        _lib.check(self._lib.igb200_group_get_render_stats(self._h, C.addressof(st), C.byref(ms)))
        return {"camera_rays": st.camera_rays, "bounce_rays": st.bounce_rays, "shadow_rays": st.shadow_rays, "rounds": st.rounds,
                "kernel_launches": st.kernel_launches, "device_ms": st.device_ms, "film_reduce_ms": ms.value}
