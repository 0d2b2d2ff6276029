"""ctypes binding of librt3.so (include/rt3.h).  Fails loudly when the CUDA library is missing:
there is no CPU fallback anywhere in this package."""
import ctypes as C
import os
import re

from .types import BuildStats, Counters, Rect, TailStats, TraceStats

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "librt3.so")
HEADER = os.path.join(_HERE, "..", "include", "rt3.h")

OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_STATE, ERR_OOM = 0, -1, -2, -3, -4, -5


class Rt3Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rt3 status {status}: {message}")
        self.status = status


def declared_symbols():
    """Every entry point include/rt3.h declares."""
    with open(HEADER) as f:
        return sorted(set(re.findall(r"\b(rt3_[a-z0-9_]+)\s*\(", f.read())))


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). raytracer3_b200 has no CPU fallback.")
    L = C.CDLL(SO_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    P = C.POINTER
    sig = {
        "rt3_create": (i32, [i32, P(vp)]),
        "rt3_destroy": (None, [vp]),
        "rt3_last_error": (C.c_char_p, [vp]),
        "rt3_abi_version": (u32, []),
        "rt3_sizeof": (C.c_size_t, [i32]),
        "rt3_scene_clear": (i32, [vp]),
        "rt3_scene_push_vertices": (i32, [vp, vp, u32, P(u32)]),
        "rt3_scene_push_indices": (i32, [vp, vp, u32, P(u32)]),
        "rt3_scene_push_geometry": (i32, [vp, vp, P(u32)]),
        "rt3_scene_push_instance": (i32, [vp, u32, u32, vp, P(u32)]),
        "rt3_scene_set_texture": (i32, [vp, u32, u32, u32, vp]),
        "rt3_scene_set_environment": (i32, [vp, u32, u32, vp]),
        "rt3_scene_set_bluenoise": (i32, [vp, u32, u32, vp]),
        "rt3_accel_build": (i32, [vp, u32, vp, P(BuildStats)]),
        "rt3_resolve_prim": (i32, [vp, u32, P(u32), P(u32), P(u32)]),
        "rt3_pass_gbuffer": (i32, [vp, vp, Rect, vp, vp, vp, vp, vp]),
        "rt3_pass_reference": (i32, [vp, vp, Rect, vp, vp, vp, vp, vp]),
        "rt3_pass_postprocess": (i32, [vp, vp, Rect, vp, vp, vp, vp]),
        "rt3_pass_present_bgra8": (i32, [vp, vp, Rect, vp, vp, vp, vp]),
        "rt3_probe_layout": (i32, [vp, P(u32), P(u32), P(u32)]),
        "rt3_pass_probe_trace": (i32, [vp, vp, vp, vp, vp, vp, vp, vp]),
        "rt3_pass_probe_trace_static": (i32, [vp, vp, vp, vp, vp, vp, vp, vp]),
        "rt3_pass_probe_sis": (i32, [vp, vp, vp, vp, vp]),
        "rt3_pass_probe_sh": (i32, [vp, vp, vp, vp, vp]),
        "rt3_pass_probe_interpolate": (i32, [vp, vp, vp, vp, vp, vp, vp]),
        "rt3_pass_gbuffer_tiles": (i32, [vp, vp, vp, u32, vp, vp, vp]),
        "rt3_pass_reference_tiles": (i32, [vp, vp, vp, u32, vp, vp, vp, vp, vp]),
        "rt3_untile_rgba32f": (i32, [vp, vp, u32, u32, u32, vp, vp, u32, u32, vp]),
        "rt3_scatter_tiles_rgba32f": (i32, [vp, vp, vp, u32, u32, u32, vp, u32, vp]),
        "rt3_trace_closest": (i32, [vp, vp, u64, vp, vp]),
        "rt3_trace_any": (i32, [vp, vp, u64, vp, vp]),
        "rt3_render_frame_host": (i32, [vp, vp, vp, vp, vp, vp]),
        "rt3_frame_submit_host": (i32, [vp, vp, vp, vp, P(u64)]),
        "rt3_frame_wait": (i32, [vp, u64]),
        "rt3_host_alloc": (i32, [vp, C.c_size_t, P(vp)]),
        "rt3_host_free": (i32, [vp, vp]),
        "rt3_trace_closest_host": (i32, [vp, vp, u64, vp]),
        "rt3_trace_any_host": (i32, [vp, vp, u64, vp]),
        "rt3_set_option": (i32, [vp, i32, i32]),
        "rt3_counters_read": (i32, [vp, P(Counters), i32]),
        "rt3_kernel_launches": (u64, [vp]),
        "rt3_measure_l2_bandwidth": (i32, [vp, C.c_size_t, i32, P(C.c_double)]),
        "rt3_trace_closest_stats": (i32, [vp, vp, u64, P(TraceStats)]),
        "rt3_trace_closest_tail_stats": (i32, [vp, vp, u64, P(TailStats)]),
        "rt3_profile_enable": (i32, [vp, i32]),
        "rt3_profile_read": (i32, [vp, P(C.c_double), P(u64)]),
        "rt3_accel_download": (i32, [vp, vp, C.c_size_t, vp, C.c_size_t]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    L._rt3_signatures = sig
    _lib = L
    return L
