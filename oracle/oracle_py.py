"""ctypes wrapper over oracle/_build/liboracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (raytracer3_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from raytracer3_b200.types import HIT, PARAMS, RAY, Counters, Rect, ptr

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("oracle.cpp", "oracle.h")] + [
        os.path.join(_HERE, "..", "include", f) for f in ("rt3.h", "rt3_shared_math.h")]
    if not force and os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_create.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_uint32]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_set_texture.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_scene_set_environment.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_scene_set_bluenoise.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_scene_build.restype = C.c_double
        L.orc_scene_build.argtypes = [C.c_void_p]
        L.orc_num_triangles.restype = C.c_uint32
        L.orc_num_triangles.argtypes = [C.c_void_p]
        L.orc_scene_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_resolve_prim.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_world_triangles.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_world_edges.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_scene_view.restype = C.c_uint64
        L.orc_scene_view.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int, C.c_int]
        L.orc_trace_any.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int, C.c_int]
        L.orc_pass_gbuffer.argtypes = [C.c_void_p, C.c_void_p, Rect, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_set_ray_capture.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]
        L.orc_pass_reference.argtypes = [C.c_void_p, C.c_void_p, Rect, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_pass_postprocess.argtypes = [C.c_void_p, C.c_void_p, Rect, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_pass_present_bgra8.argtypes = [C.c_void_p, C.c_void_p, Rect, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_pass_probe_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_pass_probe_sh.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_pass_probe_sis.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_pass_probe_interpolate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        # KAT hooks
        u32, f32, vp = C.c_uint32, C.c_float, C.c_void_p
        sig = {
            "orc_hash_jenkins": (u32, [u32]), "orc_zcurve": (u32, [u32, u32]), "orc_murmur3": (u32, [u32, u32]),
            "orc_uniform_float": (f32, [u32, u32]), "orc_rng_seed": (u32, [u32, u32, u32]), "orc_onb": (None, [vp, vp]),
            "orc_diffuse_sample": (None, [f32, f32, vp]), "orc_sincos_2pi": (None, [f32, vp]), "orc_atan2": (f32, [f32, f32]),
            "orc_asin": (f32, [f32]), "orc_log2": (f32, [f32]), "orc_exp2": (f32, [f32]), "orc_equirect_uv": (None, [vp, vp]),
            "orc_f32_to_f16": (u32, [f32]), "orc_f16_to_f32": (f32, [u32]), "orc_rgb9e5_pack": (u32, [vp]),
            "orc_rgb9e5_unpack": (None, [u32, vp]), "orc_pack_normal": (u32, [vp]), "orc_unpack_normal": (None, [u32, vp]),
            "orc_pack_color": (u32, [vp]), "orc_unpack_color": (None, [u32, vp]),
            "orc_gbuffer_pack": (None, [vp, vp, vp, f32, f32, vp]), "orc_gbuffer_unpack": (None, [vp, vp]),
            "orc_primary_ray": (None, [vp, u32, u32, vp]), "orc_ray_triangle": (C.c_int, [vp, vp, vp, vp, vp, vp]),
            "orc_r2": (None, [u32, vp]), "orc_vdc": (f32, [u32]), "orc_sample_2d": (None, [vp, u32, u32, u32, u32, u32, vp]),
            "orc_agx": (None, [vp, vp]), "orc_contract_probe": (f32, [f32, f32, f32]),
            "orc_specular_sample": (C.c_int, [f32, vp, vp, f32, f32, vp]), "orc_specular_evaluate": (None, [f32, vp, vp, vp, vp]),
            "orc_specular_dominant_direction": (None, [vp, vp, f32, vp]),
            "orc_layered_sample": (C.c_int, [vp, vp, f32, f32, vp, f32, f32, vp]), "orc_layered_evaluate": (None, [vp, vp, f32, f32, vp, vp, vp]),
            "orc_octa_decode": (None, [f32, f32, vp]), "orc_octa_encode": (None, [vp, vp]), "orc_sh3_evaluate": (None, [vp, vp]),
            "orc_sh3_unproject_cos_lobe": (None, [vp, vp, vp]), "orc_world_pos_from_depth": (None, [vp, u32, u32, f32, vp]),
            "orc_probe_sh_capacity": (u32, [vp]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def f3(a):
    return np.ascontiguousarray(a, np.float32)


class OracleScene:
    """CPU oracle scene: same inputs as rt3_scene_push_* / rt3_accel_build."""

    def __init__(self, scene, threads=None):
        L = lib()
        self.L = L
        self.threads = threads or os.cpu_count() or 1
        v = np.ascontiguousarray(scene["vertices"])
        i = np.ascontiguousarray(scene["indices"], np.uint32)
        g = np.ascontiguousarray(scene["geometries"])
        inst = scene.get("instances") or []
        first = np.array([a for a, _, _ in inst], np.uint32)
        count = np.array([b for _, b, _ in inst], np.uint32)
        xf = np.ascontiguousarray(np.array([c for _, _, c in inst], np.float32).reshape(-1))
        self.h = L.orc_scene_create(ptr(v), len(v), ptr(i), len(i), ptr(g), len(g), ptr(first) if inst else None,
                                    ptr(count) if inst else None, ptr(xf) if inst else None, len(inst))
        for slot, img in (scene.get("textures") or {}).items():
            img = np.ascontiguousarray(img, np.uint8)
            L.orc_scene_set_texture(self.h, slot, img.shape[1], img.shape[0], ptr(img))
        env = scene.get("environment")
        if env is not None:
            env = np.ascontiguousarray(env, np.float32)
            L.orc_scene_set_environment(self.h, env.shape[1], env.shape[0], ptr(env))
        bn = scene.get("bluenoise")
        if bn is not None:
            bn = np.ascontiguousarray(bn, np.uint8)
            L.orc_scene_set_bluenoise(self.h, bn.shape[1], bn.shape[0], ptr(bn))
        self.build_ms = L.orc_scene_build(self.h)
        self.num_triangles = L.orc_num_triangles(self.h)

    def __del__(self):
        try:
            self.L.orc_scene_destroy(self.h)
        except Exception:
            pass

    def bounds(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self.L.orc_scene_bounds(self.h, ptr(lo), ptr(hi))
        return lo, hi

    def resolve_prim(self, prim):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.L.orc_resolve_prim(self.h, int(prim), C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def world_triangles(self):
        out = np.zeros((self.num_triangles, 3, 3), np.float32)
        self.L.orc_world_triangles(self.h, ptr(out))
        return out

    def world_edges(self):
        """(v0, e1, e2) of every triangle in global id order: the records the traversal tests"""
        out = np.zeros((self.num_triangles, 3, 3), np.float32)
        self.L.orc_world_edges(self.h, ptr(out))
        return out

    def trace_closest(self, rays, brute=False, bvh2=False):
        """Default: the BVH4 + SSE structure (the timed CPU arm); bvh2=True: the scalar BVH2; brute=True: every triangle (the definition)."""
        rays = np.ascontiguousarray(rays)
        assert rays.dtype == RAY
        out = np.zeros(len(rays), HIT)
        self.L.orc_trace_closest(self.h, ptr(rays), len(rays), ptr(out), 0 if brute else (2 if bvh2 else 1), self.threads)
        return out

    def trace_any(self, rays, brute=False, bvh2=False):
        rays = np.ascontiguousarray(rays)
        out = np.zeros((len(rays) + 31) // 32, np.uint32)
        self.L.orc_trace_any(self.h, ptr(rays), len(rays), ptr(out), 0 if brute else (2 if bvh2 else 1), self.threads)
        return out

    def pass_gbuffer(self, params, rect=None, gbuffer=None, depth=None):
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        rect = rect or (0, 0, w, h)
        gb = np.zeros((h, w, 4), np.uint32) if gbuffer is None else gbuffer
        dp = np.zeros((h, w), np.float32) if depth is None else depth
        ids = np.zeros((h, w, 2), np.uint32)
        bary = np.zeros((h, w, 2), np.float32)
        self.L.orc_pass_gbuffer(self.h, ptr(params), Rect(*rect), ptr(gb), ptr(dp), ptr(ids), ptr(bary), self.threads)
        return gb, dp, ids, bary

    def pass_reference(self, params, gbuffer, depth, prev_light=None, rect=None, light=None):
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        rect = rect or (0, 0, w, h)
        lt = np.zeros((h, w, 4), np.float32) if light is None else light
        cnt = Counters()
        self.L.orc_pass_reference(self.h, ptr(params), Rect(*rect), ptr(gbuffer), ptr(depth), ptr(prev_light), ptr(lt), self.threads,
                                  C.byref(cnt))
        return lt, cnt

    def pass_probe_trace(self, params, gbuffer, depth, prev_atlas=None, directions=None):
        """trace_probes.slang -> probe atlas [(H/16)*8, (W/16)*8, 4]."""
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        atlas = np.zeros((h // 16 * 8, w // 16 * 8, 4), np.float32)
        self.L.orc_pass_probe_trace(self.h, ptr(params), ptr(gbuffer), ptr(depth), None if directions is None else ptr(directions),
                                    None if prev_atlas is None else ptr(prev_atlas), ptr(atlas), self.threads)
        return atlas

    def pass_probe_sis(self, params, gbuffer):
        """structured_importance_sampling.slang -> u16 direction words [(H/16)*8, (W/16)*8]."""
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        out = np.zeros((h // 16 * 8, w // 16 * 8), np.uint16)
        self.L.orc_pass_probe_sis(ptr(params), ptr(gbuffer), ptr(out), self.threads)
        return out

    def pass_probe_sh(self, params, atlas):
        """spherical_harmonic_conversion.slang -> [capacity, 9] SH-L2 records, Z-curve indexed by (probe.x * 3 + channel, probe.y)."""
        sh = np.zeros((self.L.orc_probe_sh_capacity(ptr(params)), 9), np.float32)
        self.L.orc_pass_probe_sh(ptr(params), ptr(atlas), ptr(sh), self.threads)
        return sh

    def pass_probe_interpolate(self, params, gbuffer, depth, sh, light=None):
        """interpolate_probes.slang -> Light (pixels that the shader leaves untouched keep the values of `light`)."""
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        out = np.zeros((h, w, 4), np.float32) if light is None else np.array(light, np.float32, copy=True)
        self.L.orc_pass_probe_interpolate(ptr(params), ptr(gbuffer), ptr(depth), ptr(sh), ptr(out), self.threads)
        return out

    def pass_present_bgra8(self, params, depth, inp, rect=None):
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        rect = rect or (0, 0, w, h)
        out = np.zeros((h, w), np.uint32)
        self.L.orc_pass_present_bgra8(self.h, ptr(params), Rect(*rect), ptr(depth), ptr(inp), ptr(out), self.threads)
        return out

    def pass_postprocess(self, params, depth, inp, rect=None):
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        rect = rect or (0, 0, w, h)
        out = np.zeros((h, w, 4), np.float32)
        self.L.orc_pass_postprocess(self.h, ptr(params), Rect(*rect), ptr(depth), ptr(inp), ptr(out), self.threads)
        return out
