"""Record layouts shared by the host mirror, the ctypes bindings and the tests.

Every dtype here is byte-for-byte the C struct of include/rt3.h, which in turn mirrors the
reference's records (paths under /root/reference):
  Vertex        src/assets/mod.rs:127-133 == shaders/include/datatypes.slang:5-9
  GeometryInfo  shaders/include/datatypes.slang:11-19 (+ transform used by hit_logic.slang:23)
  GConst        src/renderer/mod.rs:47-63 == shaders/include/datatypes.slang:28-43
  RayDesc       (HLSL built-in) ; RayPayload shaders/include/datatypes.slang:45-50
"""
import ctypes as C

import numpy as np

VERTEX = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("t", "<f4", 2)])
GEOMETRY = np.dtype(
    [
        ("base_color", "<f4", 4),
        ("base_color_tex", "<i4"),
        ("metallic", "<f4"),
        ("index_offset", "<u4"),
        ("vertex_offset", "<u4"),
        ("emission", "<f4", 4),
        ("roughness", "<f4"),
        ("index_count", "<u4"),
        ("_pad", "<f4", 2),
        ("xform", "<f4", 12),
    ]
)
PARAMS = np.dtype(
    [
        ("proj", "<f4", 16),
        ("view", "<f4", 16),
        ("proj_inv", "<f4", 16),
        ("view_inv", "<f4", 16),
        ("window_size", "<f4", 2),
        ("frame", "<u4"),
        ("blendfactor", "<f4"),
        ("bounces", "<u4"),
        ("samples", "<u4"),
        ("proberng", "<u4"),
        ("cell_size", "<f4"),
        ("mouse", "<u4", 2),
        ("flags", "<u4"),
        ("pad", "<u4"),
    ]
)
RAY = np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4")])
HIT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")])

assert VERTEX.itemsize == 32 and GEOMETRY.itemsize == 112 and PARAMS.itemsize == 304
assert RAY.itemsize == 32 and HIT.itemsize == 16

MISS = 0xFFFFFFFF
BACKGROUND_DEPTH = np.float32(100000.0)  # shaders/include/datatypes.slang:3

FLAG_ENV = 1
FLAG_BLUENOISE = 2
FLAG_NEE = 4
FLAG_SPECULAR = 8
BUILD_FAST_TRACE, BUILD_FAST_BUILD, BUILD_TWO_LEVEL, BUILD_TREELET, BUILD_ROTATIONS = 1, 2, 4, 8, 16

IDENTITY_3X4 = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float32)


class Rect(C.Structure):
    _fields_ = [("x0", C.c_uint32), ("y0", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


class BuildStats(C.Structure):
    _fields_ = [
        ("ms_flatten", C.c_double),
        ("ms_morton", C.c_double),
        ("ms_sort", C.c_double),
        ("ms_hierarchy", C.c_double),
        ("ms_refit", C.c_double),
        ("ms_collapse", C.c_double),
        ("ms_total", C.c_double),
        ("node_bytes", C.c_uint64),
        ("tri_bytes", C.c_uint64),
        ("triangles", C.c_uint32),
        ("wide_nodes", C.c_uint32),
        ("bvh2_nodes", C.c_uint32),
        ("levels", C.c_uint32),
        ("sah_cost", C.c_float),
        ("scene_lo", C.c_float * 3),
        ("scene_hi", C.c_float * 3),
    ]

    def as_dict(self):
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = list(v) if hasattr(v, "__len__") else v
        return d


class Counters(C.Structure):
    _fields_ = [
        ("rays_primary", C.c_uint64),
        ("rays_secondary", C.c_uint64),
        ("rays_shadow", C.c_uint64),
        ("paths_started", C.c_uint64),
        ("paths_escaped", C.c_uint64),
    ]


class TraceStats(C.Structure):
    _fields_ = [
        ("nodes_per_ray", C.c_double),
        ("tris_per_ray", C.c_double),
        ("node_bytes_per_ray", C.c_double),
        ("tri_bytes_per_ray", C.c_double),
    ]


class TailStats(C.Structure):
    _fields_ = [("kernel_us", C.c_double), ("mean_warp_busy_us", C.c_double), ("mean_warp_tail_us", C.c_double), ("tail_fraction", C.c_double),
                ("warps", C.c_uint32), ("_pad", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "_pad"}


def ptr(a, ctype=C.c_void_p):
    """ctypes pointer to a contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctype)
