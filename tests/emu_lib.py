"""ctypes wrapper over tests/emu/libemu.so (CPU emulation of the device BVH logic; test infrastructure)."""
import ctypes as C
import os
import subprocess

import numpy as np

from raytracer3_b200.types import HIT, ptr

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "emu", "libemu.so")
_lib = None
_wlibs = {}


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "emu", "emu.cpp")
        hdr = os.path.join(_HERE, "..", "raytracer3_b200", "csrc", "bvh_logic.h")
        hdr2 = os.path.join(_HERE, "emu", "bvh_experiments.h")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(src), os.path.getmtime(hdr), os.path.getmtime(hdr2)):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-march=x86-64-v3", "-Wno-unknown-pragmas"]
                                  + os.environ.get("RT3_EMU_DEFINES", "").split()          # experiments only (e.g. -DRT3_TREELET_LEAVES=9)
                                  + ["-I", os.path.join(_HERE, "..", "include"), "-shared", "-o", _SO, src])
        L = C.CDLL(_SO)
        L.emu_build.restype = C.c_void_p
        L.emu_build.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_build_dp.restype = C.c_void_p
        L.emu_build_dp.argtypes = [C.c_void_p, C.c_uint32]
        L.emu_build_split.restype = C.c_void_p
        L.emu_build_split.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_float]
        L.emu_build_ex.restype = C.c_void_p
        L.emu_build_ex.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_float, C.c_uint32, C.c_int]
        L.emu_build_treelet.restype = C.c_void_p
        L.emu_build_treelet.argtypes = [C.c_void_p, C.c_uint32, C.c_int]
        L.emu_build_ploc.restype = C.c_void_p
        L.emu_build_ploc.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_uint32]
        L.emu_ploc_rounds.restype = C.c_uint32
        L.emu_ploc_rounds.argtypes = [C.c_void_p]
        L.emu_wrap.restype = C.c_void_p
        L.emu_wrap.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        L.emu_set_input_edges.argtypes = [C.c_int]
        L.emu_destroy.argtypes = [C.c_void_p]
        L.emu_export.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.emu_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.emu_check_triangles.restype = C.c_uint32
        L.emu_check_triangles.argtypes = [C.c_void_p]
        L.emu_check_node_v2.restype = C.c_uint32
        L.emu_check_node_v2.argtypes = [C.c_void_p]
        L.emu_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p]
        L.emu_trace_any.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        _lib = L
    return _lib


def warp_lib(balance=False, defines=(), load=True):
    """tests/emu/warp_emu.cpp: the production kernel source (csrc/trace_kernel.cuh) over a lock-step warp emulation.
    balance=True builds it with -DRT3_TRI_BALANCE=1 (the warp-balanced record phase, an experiment not in the product build)."""
    defines = tuple(defines) + (("RT3_TRI_BALANCE=1",) if balance else ())
    key = "_".join(d.replace("=", "") for d in defines)
    if key not in _wlibs:
        so = os.path.join(_HERE, "emu", f"libwemu_{key}.so" if key else "libwemu.so")
        src = os.path.join(_HERE, "emu", "warp_emu.cpp")
        csrc = os.path.join(_HERE, "..", "raytracer3_b200", "csrc")
        deps = [src, os.path.join(_HERE, "emu", "warp_emu.h"), os.path.join(csrc, "bvh_logic.h"), os.path.join(csrc, "trace_kernel.cuh"),
                os.path.join(_HERE, "..", "include", "rt3_shared_math.h")]
        if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-march=x86-64-v3", "-Wno-unknown-pragmas"]
                                  + ["-D" + d for d in defines]
                                  + ["-I", os.path.join(_HERE, "..", "include"), "-shared", "-o", so, src])
        if not load:                                              # __graft_entry__.build(): compile the checker, do not open it
            return so
        L = C.CDLL(so)
        L.wemu_trace.restype = C.c_int
        L.wemu_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.wemu_last_error.restype = C.c_char_p
        L.wemu_tri_balance.restype = C.c_int
        L.wemu_phase_stats.argtypes = [C.c_void_p]
        _wlibs[key] = L
    return _wlibs[key]


def pass_lib(load=True):
    """tests/emu/pass_emu.cpp: the product's pass kernels (csrc/pass_kernels.cuh) + traversal kernel over the warp emulation."""
    if "pass" not in _wlibs:
        so = os.path.join(_HERE, "emu", "libpemu.so")
        src = os.path.join(_HERE, "emu", "pass_emu.cpp")
        csrc = os.path.join(_HERE, "..", "raytracer3_b200", "csrc")
        deps = [src, os.path.join(_HERE, "emu", "warp_emu.h"), os.path.join(_HERE, "..", "include", "rt3_shared_math.h")] + [
            os.path.join(csrc, f) for f in ("bvh_logic.h", "trace_kernel.cuh", "pass_kernels.cuh", "pass_types.h")]
        if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-march=x86-64-v3", "-Wno-unknown-pragmas",
                                   "-I", os.path.join(_HERE, "..", "include"), "-shared", "-o", so, src])
        if not load:
            return so
        L = C.CDLL(so)
        L.pemu_shade_records.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.pemu_frame.restype = C.c_int
        L.pemu_frame.argtypes = [C.c_void_p] * 13
        L.pemu_last_error.restype = C.c_char_p
        _wlibs["pass"] = L
    return _wlibs["pass"]


class PassEmu:
    """Whole frames of the product's kernels on the CPU: scene tables from the oracle's scene, acceleration structure from the CPU
    emulation of the device builder over the oracle's (v0, e1, e2) records."""

    def __init__(self, oracle_scene, **build_options):
        self.P = pass_lib()
        self.orc = oracle_scene
        self.view = np.zeros(int(oracle_scene.L.orc_scene_view(oracle_scene.h, None, 0)), np.uint8)
        oracle_scene.L.orc_scene_view(oracle_scene.h, ptr(self.view), len(self.view))
        edges = oracle_scene.world_edges()
        self.bvh = EmuBvh(tris=edges, edges=True, **build_options)
        self.nodes, self.recs = self.bvh.export()
        self.shade_rec = np.zeros((max(len(edges), 1), 4), np.uint32)
        self.P.pemu_shade_records(ptr(self.view), len(edges), ptr(self.shade_rec))

    def frame(self, params, prev_light=None, postprocess=False, strips=1, round_lengths=None, shade64=True):
        """strips: the number of concurrent wavefronts rt3_launch_reference would split the frame into (pass_types.h); round_lengths: a
        uint64[4096] array that receives the queue length at the start of every round, summed over the strips (regeneration statistics)"""
        w, h = int(params["window_size"][0][0]), int(params["window_size"][0][1])
        self.P.pemu_set_strips.argtypes = [C.c_int]
        self.P.pemu_set_round_lengths.argtypes = [C.c_void_p]
        self.P.pemu_set_shade64.argtypes = [C.c_int, C.c_uint32]
        self.P.pemu_set_shade64(1 if shade64 else 0, len(self.shade_rec) if self.orc.num_triangles else 0)
        self.P.pemu_set_strips(int(strips))
        self.P.pemu_set_round_lengths(None if round_lengths is None else ptr(round_lengths))
        out = dict(gbuffer=np.zeros((h, w, 4), np.uint32), depth=np.zeros((h, w), np.float32), ids=np.zeros((h, w, 2), np.uint32),
                   bary=np.zeros((h, w, 2), np.float32), light=np.zeros((h, w, 4), np.float32), counters=np.zeros(5, np.uint64))
        post = np.zeros((h, w, 4), np.float32) if postprocess else None
        rc = self.P.pemu_frame(ptr(self.view), ptr(self.shade_rec), ptr(self.nodes), ptr(self.recs), ptr(params), ptr(out["gbuffer"]), ptr(out["depth"]),
                               ptr(out["ids"]), ptr(out["bary"]), None if prev_light is None else ptr(prev_light), ptr(out["light"]),
                               None if post is None else ptr(post), ptr(out["counters"]))
        if rc != 0:
            raise RuntimeError("warp emulation: " + self.P.pemu_last_error().decode())
        out["post"] = post
        return out


class EmuBvh:
    def __init__(self, tris=None, nodes=None, tri_records=None, optimal_collapse=True, split_beta=0.0, ploc_radius=0, treelet_passes=0, edges=False):
        """optimal_collapse: children of the wide nodes by the cost dynamic programme (the device default) or greedily (RT3_BUILD_FAST_BUILD);
        split_beta > 0: triangle splitting with a budget of split_beta x triangles extra references (RT3_BUILD_SPLIT);
        edges: `tris` holds (v0, e1, e2) records (e.g. the oracle's, OracleScene.world_edges) instead of three vertices"""
        self.L = lib()
        self.L.emu_set_input_edges(1 if edges else 0)
        if tris is not None:
            t = np.ascontiguousarray(tris, np.float32)
            if (treelet_passes > 0) + (ploc_radius > 0) + (split_beta > 0) > 1:                  # combinations (experiments)
                self.h = self.L.emu_build_ex(ptr(t), len(t), 1 if optimal_collapse else 0, split_beta, ploc_radius, treelet_passes)
            elif treelet_passes > 0:
                self.h = self.L.emu_build_treelet(ptr(t), len(t), treelet_passes)
            elif ploc_radius > 0:
                self.h = self.L.emu_build_ploc(ptr(t), len(t), 1 if optimal_collapse else 0, ploc_radius)
            elif split_beta > 0:
                self.h = self.L.emu_build_split(ptr(t), len(t), 1 if optimal_collapse else 0, split_beta)
            else:
                self.h = (self.L.emu_build_dp if optimal_collapse else self.L.emu_build)(ptr(t), len(t))
        else:
            n, r = np.ascontiguousarray(nodes, np.uint32), np.ascontiguousarray(tri_records, np.uint32)
            self.h = self.L.emu_wrap(ptr(n), len(n), ptr(r), len(r))

    def __del__(self):
        try:
            self.L.emu_destroy(self.h)
        except Exception:
            pass

    def info(self):
        a, s = np.zeros(4, np.uint32), np.zeros(2, np.float32)
        self.L.emu_info(self.h, ptr(a), ptr(s))
        return dict(nodes=int(a[0]), tris=int(a[1]), levels=int(a[2]), input_tris=int(a[3]), sah_nodes=float(s[0]), sah_tris=float(s[1]))

    def ploc_rounds(self):
        return int(self.L.emu_ploc_rounds(self.h))

    def export(self):
        """(wide nodes [n, 20] u32, leaf records [m, 12] u32) as the device holds them"""
        i = self.info()
        nodes, recs = np.zeros((max(i["nodes"], 1), 20), np.uint32), np.zeros((max(i["tris"], 1), 12), np.uint32)
        self.L.emu_export(self.h, ptr(nodes), ptr(recs))
        return nodes, recs

    def kernel_trace(self, rays, any_hit=False, blocks=2, d_count=False, balance=False, defines=(), static=False):
        """the production kernel k_trace_dyn on the warp emulation: hits (closest) or result bits (any); also the number of warp collectives run"""
        W = warp_lib(balance, defines)
        nodes, recs = self.export()
        rays = np.ascontiguousarray(rays)
        hits, bits, coll = np.zeros(len(rays), HIT), np.zeros((len(rays) + 31) // 32 + 1, np.uint32), np.zeros(1, np.uint64)
        rc = W.wemu_trace(ptr(nodes), ptr(recs), None, ptr(rays), len(rays), 2 if static else (1 if any_hit else 0), blocks, 1 if d_count else 0, ptr(hits), ptr(bits), ptr(coll))
        if rc != 0:
            raise RuntimeError("warp emulation: " + W.wemu_last_error().decode())
        ph = np.zeros((8, 3), np.uint64)
        W.wemu_phase_stats(ptr(ph))
        self.phase_stats = ph          # [phase][warp executions with >= 1 lane, lanes taking part, times reached]; phases: see warp_emu.cpp
        return (bits[:(len(rays) + 31) // 32] if any_hit else hits), int(coll[0])

    def check_triangles(self):
        return self.L.emu_check_triangles(self.h)

    def check_node_format(self):
        """hit-byte-first node format (bvh_logic.h): permutation table, record_index / dense_records, valid24 of every node"""
        return self.L.emu_check_node_v2(self.h)

    def trace_closest(self, rays, brute=False):
        rays = np.ascontiguousarray(rays)
        out, st = np.zeros(len(rays), HIT), np.zeros(2)
        self.L.emu_trace_closest(self.h, ptr(rays), len(rays), ptr(out), 1 if brute else 0, ptr(st))
        return out, st

    def trace_any(self, rays, brute=False):
        rays = np.ascontiguousarray(rays)
        out = np.zeros((len(rays) + 31) // 32, np.uint32)
        self.L.emu_trace_any(self.h, ptr(rays), len(rays), ptr(out), 1 if brute else 0)
        return out
