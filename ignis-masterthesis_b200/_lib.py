"""ctypes binding of the C ABI declared in include/ignis_b200.h.

The shared library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no CPU
fallback: if the library is missing, loading raises, and ``igb200_create`` fails without a B200.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("IGB200_LIB") or os.path.join(_HERE, "lib", "libignis_b200.so")  # IGB200_LIB: experiment builds

# Every symbol include/ignis_b200.h declares (checked by tests/test_abi.py).
ABI_SYMBOLS = [
    "igb200_abi_version",
    "igb200_last_error",
    "igb200_create",
    "igb200_destroy",
    "igb200_assign_scene",
    "igb200_trace_rays",
    "igb200_trace_rays_device",
    "igb200_get_trace_stats",
    "igb200_get_bvh_info",
    "igb200_debug_check_bvh8",
    "igb200_debug_bvh8_quality",
    "igb200_assign_shading",
    "igb200_render",
    "igb200_render_iterations",
    "igb200_render_pipelined",
    "igb200_flush",
    "igb200_framebuffer",
    "igb200_clear_framebuffer",
    "igb200_clear_aov",
    "igb200_tonemap",
    "igb200_resize",
    "igb200_set_framebuffer",
    "igb200_get_render_stats",
    "igb200_get_section_stats",
    # ABI v6: multi-GPU
    "igb200_partition_share",
    "igb200_render_partition",
    "igb200_set_iteration_count",
    "igb200_device_ordinal",
    "igb200_comm_unique_id",
    "igb200_comm_create",
    "igb200_comm_destroy",
    "igb200_comm_render",
    "igb200_comm_reduce_film",
    "igb200_group_create",
    "igb200_group_destroy",
    "igb200_group_size",
    "igb200_group_device",
    "igb200_group_assign_scene",
    "igb200_group_assign_shading",
    "igb200_group_resize",
    "igb200_group_clear_framebuffer",
    "igb200_group_render",
    "igb200_group_trace_rays",
    "igb200_group_get_render_stats",
    "igb200_trace_rays_noalloc",
]


class Ray(C.Structure):
    _fields_ = [("org", C.c_float * 3), ("dir", C.c_float * 3), ("tmin", C.c_float), ("tmax", C.c_float)]


class LookupEntry(C.Structure):
    _fields_ = [("type_id", C.c_uint32), ("flags", C.c_uint32), ("offset", C.c_uint64)]


class EntityLeaf(C.Structure):
    _fields_ = [
        ("bbox_min", C.c_float * 3),
        ("entity_id", C.c_int32),
        ("bbox_max", C.c_float * 3),
        ("shape_id", C.c_int32),
        ("local", C.c_float * 12),
        ("flags", C.c_uint32),
        ("mat_id", C.c_int32),
        ("user1", C.c_int32),
        ("user2", C.c_int32),
    ]


class Scene(C.Structure):
    _fields_ = [
        ("shape_lookups", C.c_void_p),
        ("shape_count", C.c_uint64),
        ("shape_data", C.c_void_p),
        ("shape_bytes", C.c_uint64),
        ("entities", C.c_void_p),
        ("entity_count", C.c_uint64),
        ("leaves", C.c_void_p * 2),
        ("leaf_count", C.c_uint64 * 2),
        ("scene_bbox", C.c_float * 6),
        ("material_count", C.c_uint32),
        ("_pad", C.c_uint32),
    ]


class Setup(C.Structure):
    _fields_ = [("device", C.c_int32), ("acquire_stats", C.c_int32), ("debug_trace", C.c_int32), ("_pad", C.c_int32)]


class SectionStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_float * 6), ("launches", C.c_uint64 * 6), ("framebuffer_host_update_ms", C.c_float),
                ("framebuffer_host_updates", C.c_uint32)]


class TraceStats(C.Structure):
    _fields_ = [
        ("rays", C.c_uint64),
        ("nodes_visited", C.c_uint64),
        ("tris_tested", C.c_uint64),
        ("entities_entered", C.c_uint64),
        ("kernel_ms", C.c_float),
        ("_pad", C.c_float),
    ]


assert C.sizeof(Ray) == 32 and C.sizeof(LookupEntry) == 16 and C.sizeof(EntityLeaf) == 96

_lib = None


def load():
    """Load libignis_b200.so (once). Raises if it has not been built -- never falls back to anything."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the backend has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
    lib.igb200_abi_version.restype = i32
    lib.igb200_last_error.restype = C.c_char_p
    lib.igb200_create.argtypes = [C.POINTER(Setup), C.POINTER(vp)]
    lib.igb200_destroy.argtypes = [vp]
    lib.igb200_destroy.restype = None
    lib.igb200_assign_scene.argtypes = [vp, C.POINTER(Scene)]
    lib.igb200_trace_rays.argtypes = [vp, vp, u64, u32, i32, vp, vp, vp, vp, vp]
    lib.igb200_trace_rays_noalloc.argtypes = [vp, vp, u64, u32, i32, vp, vp, vp, vp, vp]
    lib.igb200_trace_rays_device.argtypes = [vp, vp, u64, u32, i32, vp, vp, vp, vp, vp, vp]
    lib.igb200_get_trace_stats.argtypes = [vp, C.POINTER(TraceStats)]
    lib.igb200_get_bvh_info.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u32)]
    lib.igb200_debug_check_bvh8.argtypes = [vp, u32, u32, vp]
    lib.igb200_debug_bvh8_quality.argtypes = [vp, u32, u32, vp]
    lib.igb200_assign_shading.argtypes = [vp, vp]
    lib.igb200_render.argtypes = [vp, vp]
    lib.igb200_render_iterations.argtypes = [vp, vp, u32]
    lib.igb200_render_pipelined.argtypes = [vp, vp]
    lib.igb200_flush.argtypes = [vp]
    lib.igb200_framebuffer.argtypes = [vp, C.c_char_p, i32, C.POINTER(vp), C.POINTER(u64)]
    lib.igb200_clear_framebuffer.argtypes = [vp]
    lib.igb200_clear_aov.argtypes = [vp, C.c_char_p]
    lib.igb200_tonemap.argtypes = [vp, vp, vp]
    lib.igb200_resize.argtypes = [vp, u32, u32]
    lib.igb200_set_framebuffer.argtypes = [vp, vp, u64, u64]
    lib.igb200_get_render_stats.argtypes = [vp, vp]
    lib.igb200_get_section_stats.argtypes = [vp, vp, i32]
    pu32 = C.POINTER(u32)
    lib.igb200_partition_share.argtypes = [u32, u32, u32, u32, pu32, pu32, pu32, pu32]
    lib.igb200_render_partition.argtypes = [vp, vp, u32, u32, u32, u32]
    lib.igb200_set_iteration_count.argtypes = [vp, C.c_char_p, u64]
    lib.igb200_device_ordinal.argtypes = [vp]
    lib.igb200_comm_unique_id.argtypes = [vp]
    lib.igb200_comm_create.argtypes = [vp, vp, u32, u32, C.POINTER(vp)]
    lib.igb200_comm_destroy.argtypes = [vp]
    lib.igb200_comm_destroy.restype = None
    lib.igb200_comm_render.argtypes = [vp, vp, u32, u32]
    lib.igb200_comm_reduce_film.argtypes = [vp, u32, u32, u32, i32, C.POINTER(C.c_float)]
    lib.igb200_group_create.argtypes = [C.POINTER(C.c_int32), u32, i32, C.POINTER(vp)]
    lib.igb200_group_destroy.argtypes = [vp]
    lib.igb200_group_destroy.restype = None
    lib.igb200_group_size.argtypes = [vp]
    lib.igb200_group_size.restype = u32
    lib.igb200_group_device.argtypes = [vp, u32]
    lib.igb200_group_device.restype = vp
    lib.igb200_group_assign_scene.argtypes = [vp, C.POINTER(Scene)]
    lib.igb200_group_assign_shading.argtypes = [vp, vp]
    lib.igb200_group_resize.argtypes = [vp, u32, u32]
    lib.igb200_group_clear_framebuffer.argtypes = [vp]
    lib.igb200_group_render.argtypes = [vp, vp, u32, u32]
    lib.igb200_group_trace_rays.argtypes = [vp, vp, u64, u32, i32, vp, vp, vp, vp, vp]
    lib.igb200_group_get_render_stats.argtypes = [vp, vp, C.POINTER(C.c_float)]
    _lib = lib
    return lib


HOST_LIB_PATH = os.path.join(_HERE, "lib", "libig_runtime_b200.so")
_host = None


def load_host():
    """Load libig_runtime_b200.so, the C++ host side above the C ABI (loader pieces: sky bake, scene reader)."""
    global _host
    if _host is None:
        load()  # dependency, resolved through the rpath as well
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError(f"{HOST_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        h = C.CDLL(HOST_LIB_PATH)
        fp = C.POINTER(C.c_float)
        h.ig_b200_bake_sky.argtypes = [fp, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, fp]
        h.ig_b200_bake_sky.restype = C.c_int
        h.ig_b200_hosek_radiance.argtypes = [C.c_int] + [C.c_double] * 5
        h.ig_b200_hosek_radiance.restype = C.c_double
        _host = h
    return _host


class BackendError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"ignis_b200 error {code}: {message}")
        self.code = code


def check(rc):
    if rc != 0:
        raise BackendError(rc, load().igb200_last_error().decode("utf-8", "replace"))
