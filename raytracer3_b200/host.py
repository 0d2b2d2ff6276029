"""Host-side mirror of the reference's renderer API for the ray-tracing path, over the C ABI.

The reference is Rust; no Rust toolchain exists in this image, so the host side above librt3.so is
this Python module (tests and bench are Python; INTEGRATION.md shows the Rust `extern "C"` binding).
Names, argument meaning and error behaviour follow the reference (paths under /root/reference):

  RenderContext       Context + RayTracingContext            src/renderer/vulkan/mod.rs:86-231, raytracing.rs:56-86
  RenderWorld         global append-only scene buffers + AS  src/renderer/world/mod.rs:103-180
  RenderGraph.image   named transient image resources        src/renderer/render_graph/mod.rs:440-483
  RayTracingPass      builder .shader .constants .read .write .read_write .launch(WorkSize2D)
                                                             src/renderer/render_graph/executions.rs:80-121, build.rs:66-193,371-384
  ComputePass         .dispatch(WorkSize2D) (postprocess)    src/renderer/render_graph/executions.rs:21-55
  WorkSize2D          FullScreen / FractionalFullScreen / X / XY   executions.rs:57-78

Errors: the reference unwraps every Result (src/renderer/mod.rs:30), i.e. panics; here every non-zero
rt3_status raises Rt3Error with rt3_last_error()'s message.  Device memory is torch CUDA tensors
(plumbing only); passes are recorded on the current torch CUDA stream, the analogue of the reference's
command buffer.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from .types import PARAMS, BuildStats, Counters, Rect, TailStats, TraceStats, ptr


def _torch():
    import torch

    return torch


class RenderContext:
    """One per CUDA device (rt3_context)."""

    def __init__(self, device=0):
        self.L = _lib.load()
        h = C.c_void_p()
        st = self.L.rt3_create(int(device), C.byref(h))
        if st != 0:
            raise _lib.Rt3Error(st, self.L.rt3_last_error(None).decode())
        self.h = h
        self.device = int(device)

    def check(self, status):
        if status != 0:
            raise _lib.Rt3Error(status, self.L.rt3_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.rt3_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self):
        torch = _torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def set_strips(self, n):
        """RT3_OPT_STRIPS: concurrent wavefronts per reference-pass call (0 = default)."""
        self.check(self.L.rt3_set_option(self.h, 1, int(n)))

    def kernel_launches(self):
        return int(self.L.rt3_kernel_launches(self.h))

    def counters(self, reset=False):
        c = Counters()
        self.check(self.L.rt3_counters_read(self.h, C.byref(c), 1 if reset else 0))
        return c

    def profile_enable(self, on=True):
        self.check(self.L.rt3_profile_enable(self.h, 1 if on else 0))

    def profile_read(self):
        ms, n = C.c_double(), C.c_uint64()
        self.check(self.L.rt3_profile_read(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def measure_l2_bandwidth(self, nbytes=32 << 20, iters=10):
        out = C.c_double()
        self.check(self.L.rt3_measure_l2_bandwidth(self.h, nbytes, iters, C.byref(out)))
        return out.value


class RenderWorld:
    """Global vertex / index / geometry / instance buffers and the acceleration structure."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.num_instances = 0
        self.build_stats = None

    def push_vertices(self, vertices):
        v = np.ascontiguousarray(vertices)
        first = C.c_uint32()
        self.ctx.check(self.ctx.L.rt3_scene_push_vertices(self.ctx.h, ptr(v), len(v), C.byref(first)))
        return first.value

    def push_indices(self, indices):
        i = np.ascontiguousarray(indices, np.uint32)
        first = C.c_uint32()
        self.ctx.check(self.ctx.L.rt3_scene_push_indices(self.ctx.h, ptr(i), len(i), C.byref(first)))
        return first.value

    def push_geometry(self, geometry):
        g = np.ascontiguousarray(geometry)
        out = C.c_uint32()
        self.ctx.check(self.ctx.L.rt3_scene_push_geometry(self.ctx.h, ptr(g), C.byref(out)))
        return out.value

    def push_instance(self, first_geometry, geometry_count, xform3x4):
        x = np.ascontiguousarray(xform3x4, np.float32).reshape(12)
        out = C.c_uint32()
        self.ctx.check(self.ctx.L.rt3_scene_push_instance(self.ctx.h, first_geometry, geometry_count, ptr(x), C.byref(out)))
        self.num_instances += 1
        return out.value

    def set_texture(self, slot, rgba8):
        img = np.ascontiguousarray(rgba8, np.uint8)
        self.ctx.check(self.ctx.L.rt3_scene_set_texture(self.ctx.h, slot, img.shape[1], img.shape[0], ptr(img)))

    def set_environment(self, rgb):
        img = np.ascontiguousarray(rgb, np.float32)
        self.ctx.check(self.ctx.L.rt3_scene_set_environment(self.ctx.h, img.shape[1], img.shape[0], ptr(img)))

    def set_bluenoise(self, rgba8):
        img = np.ascontiguousarray(rgba8, np.uint8)
        self.ctx.check(self.ctx.L.rt3_scene_set_bluenoise(self.ctx.h, img.shape[1], img.shape[0], ptr(img)))

    def clear(self):
        self.ctx.check(self.ctx.L.rt3_scene_clear(self.ctx.h))
        self.num_instances = 0

    def load_scene(self, scene):
        """Upload a scenes.py dict; geometry offsets in the dict are relative to the dict's own buffers."""
        v0 = self.push_vertices(scene["vertices"])
        i0 = self.push_indices(scene["indices"])
        g0 = None
        for g in scene["geometries"]:
            g = np.array(g, copy=True).reshape(1)
            g["vertex_offset"] += v0
            g["index_offset"] += i0
            k = self.push_geometry(g)
            g0 = k if g0 is None else g0
        for first, count, xf in scene.get("instances") or []:
            self.push_instance((g0 or 0) + first, count, xf)
        for slot, img in (scene.get("textures") or {}).items():
            self.set_texture(slot, img)
        if scene.get("environment") is not None:
            self.set_environment(scene["environment"])
        if scene.get("bluenoise") is not None:
            self.set_bluenoise(scene["bluenoise"])

    def create_acceleration_structure(self, flags=1):
        """RayTracingContext::create_acceleration_structure (raytracing.rs:88-148), PREFER_FAST_TRACE by default."""
        bs = BuildStats()
        self.ctx.check(self.ctx.L.rt3_accel_build(self.ctx.h, flags, self.ctx.stream(), C.byref(bs)))
        self.build_stats = bs
        return bs

    def download_accel(self):
        """(nodes [n,20] u32, tris [m,12] u32 view of the 48-byte records) of the built acceleration structure."""
        bs = self.build_stats
        nodes = np.zeros((bs.wide_nodes, 20), np.uint32)
        tris = np.zeros((bs.triangles, 12), np.uint32)
        self.ctx.check(self.ctx.L.rt3_accel_download(self.ctx.h, ptr(nodes), nodes.nbytes, ptr(tris), tris.nbytes))
        return nodes, tris

    def resolve_prim(self, prim):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.ctx.check(self.ctx.L.rt3_resolve_prim(self.ctx.h, int(prim), C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    # raw queries ---------------------------------------------------------------------------------
    def trace_closest(self, rays_dev, hits_dev=None):
        torch = _torch()
        n = rays_dev.shape[0]
        if hits_dev is None:
            hits_dev = torch.empty((n, 4), dtype=torch.float32, device=rays_dev.device)
        self.ctx.check(self.ctx.L.rt3_trace_closest(self.ctx.h, rays_dev.data_ptr(), n, hits_dev.data_ptr(), self.ctx.stream()))
        return hits_dev

    def trace_any(self, rays_dev, bits_dev=None):
        torch = _torch()
        n = rays_dev.shape[0]
        if bits_dev is None:
            bits_dev = torch.empty(((n + 31) // 32,), dtype=torch.int32, device=rays_dev.device)
        self.ctx.check(self.ctx.L.rt3_trace_any(self.ctx.h, rays_dev.data_ptr(), n, bits_dev.data_ptr(), self.ctx.stream()))
        return bits_dev

    def trace_closest_host(self, rays, out=None):
        from .types import HIT
        rays = np.ascontiguousarray(rays)
        out = np.empty(len(rays), HIT) if out is None else out
        self.ctx.check(self.ctx.L.rt3_trace_closest_host(self.ctx.h, ptr(rays), len(rays), ptr(out)))
        return out

    def trace_any_host(self, rays, out=None):
        rays = np.ascontiguousarray(rays)
        out = np.empty((len(rays) + 31) // 32, np.uint32) if out is None else out
        self.ctx.check(self.ctx.L.rt3_trace_any_host(self.ctx.h, ptr(rays), len(rays), ptr(out)))
        return out

    def trace_tail_stats(self, rays_dev):
        """Ramp-up / tail of one persistent traversal launch over `rays_dev` (instrumented kernel, rt3.h rt3_tail_stats)."""
        ts = TailStats()
        self.ctx.check(self.ctx.L.rt3_trace_closest_tail_stats(self.ctx.h, rays_dev.data_ptr(), rays_dev.shape[0], C.byref(ts)))
        return ts

    def trace_stats(self, rays_dev):
        ts = TraceStats()
        self.ctx.check(self.ctx.L.rt3_trace_closest_stats(self.ctx.h, rays_dev.data_ptr(), rays_dev.shape[0], C.byref(ts)))
        return ts


@dataclass(frozen=True)
class WorkSize2D:
    """executions.rs:57-78"""
    kind: str = "FullScreen"
    x: int = 0
    y: int = 0

    @staticmethod
    def FullScreen():
        return WorkSize2D("FullScreen")

    @staticmethod
    def FractionalFullScreen(x, y):
        return WorkSize2D("FractionalFullScreen", x, y)

    @staticmethod
    def X(x):
        return WorkSize2D("X", x, 1)

    @staticmethod
    def XY(x, y):
        return WorkSize2D("XY", x, y)

    def size(self, window):
        w, h = window
        if self.kind == "FullScreen":
            return w, h
        if self.kind == "FractionalFullScreen":
            return -(-w // self.x), -(-h // self.y)
        return self.x, self.y


_FORMATS = {"B8G8R8A8": ("int32", 1), "RGBA32UI": ("int32", 4), "R32F": ("float32", 1), "RGBA32F": ("float32", 4), "RG32UI": ("int32", 2), "RG32F": ("float32", 2), "R16UI": ("int16", 1)}


class RenderGraph:
    """Named full-screen images backed by torch CUDA tensors (render_graph/mod.rs:440-483); passes run in
    declaration order on the current stream, so no barrier synthesis is needed."""

    def __init__(self, ctx, world, window_size):
        self.ctx, self.world = ctx, world
        self.window = (int(window_size[0]), int(window_size[1]))
        self.resources = {}
        self.frame_number = 0

    def image(self, fmt, name):
        torch = _torch()
        if name not in self.resources:
            dt, ch = _FORMATS[fmt]
            w, h = self.window
            shape = (h, w, ch) if ch > 1 else (h, w)
            self.resources[name] = torch.zeros(shape, dtype=getattr(torch, dt), device=f"cuda:{self.ctx.device}")
        return name

    def buffer(self, name, shape, dtype="float32"):
        """A resource that is not window sized (ImageSize::Custom / buffers of render_graph/mod.rs:440-483): probe atlas, SH records."""
        torch = _torch()
        if name not in self.resources or tuple(self.resources[name].shape) != tuple(shape):
            self.resources[name] = torch.zeros(tuple(shape), dtype=getattr(torch, dtype), device=f"cuda:{self.ctx.device}")
        return name

    def tensor(self, name):
        return self.resources[name]


class _PassBuilder:
    def __init__(self, rg, name):
        self.rg, self.name = rg, name
        self._shader, self._entry, self._constants = "", "main", None
        self._bind = {}

    def shader(self, path):
        self._shader = path
        return self

    def entry(self, name):
        self._entry = name
        return self

    def constants(self, gconst):
        p = np.ascontiguousarray(gconst)
        assert p.dtype == PARAMS and p.size == 1
        self._constants = p
        return self

    def read(self, slot, res):
        self._bind[slot] = res
        return self

    write = read
    read_write = read

    def _ptr(self, slot, required=True):
        if slot not in self._bind:
            if required:
                raise KeyError(f"pass '{self.name}' ({self._shader}): resource slot '{slot}' not bound")
            return None
        return C.c_void_p(self.rg.tensor(self._bind[slot]).data_ptr())


class RayTracingPass(_PassBuilder):
    """RayTracingPass::new(rg, name) ... .launch(WorkSize2D) -> vkCmdTraceRaysKHR(x, y, 1)."""

    @staticmethod
    def new(rg, name):
        return RayTracingPass(rg, name)

    def launch(self, work=WorkSize2D.FullScreen()):
        ctx = self.rg.ctx
        if self._constants is None:
            raise ValueError("constants(&gconst) not set")
        x, y = work.size(self.rg.window)
        rect = Rect(0, 0, x, y)
        p = ptr(self._constants)
        if self._shader == "gbuffer":
            st = ctx.L.rt3_pass_gbuffer(ctx.h, p, rect, self._ptr("gbuffer"), self._ptr("gbuffer_depth"), self._ptr("ids", False),
                                        self._ptr("bary", False), ctx.stream())
        elif self._shader == "refrence_mode":
            st = ctx.L.rt3_pass_reference(ctx.h, p, rect, self._ptr("gbuffer"), self._ptr("gbuffer_depth"), self._ptr("PrevLight", False),
                                          self._ptr("Light"), ctx.stream())
        elif self._shader == "trace_probes":                      # shaders/old/trace_probes.slang, launched over the probe atlas
            st = ctx.L.rt3_pass_probe_trace(ctx.h, p, self._ptr("gbuffer"), self._ptr("gbuffer_depth"), self._ptr("directions", False),
                                            self._ptr("prev_probe_atlas", False), self._ptr("probe_atlas"), ctx.stream())
        else:
            raise ValueError(f"unknown ray-tracing shader '{self._shader}' (known: gbuffer, refrence_mode, trace_probes)")
        ctx.check(st)
        return self.name


class ComputePass(_PassBuilder):
    @staticmethod
    def new(rg, name):
        return ComputePass(rg, name)

    def dispatch(self, work=WorkSize2D.FullScreen()):
        ctx = self.rg.ctx
        x, y = work.size(self.rg.window)
        p = ptr(self._constants)
        if self._shader == "postprocess":
            ctx.check(ctx.L.rt3_pass_postprocess(ctx.h, p, Rect(0, 0, x, y), self._ptr("Depth"), self._ptr("In"), self._ptr("Out"), ctx.stream()))
        elif self._shader == "present":                            # postprocess + the swapchain blit's conversion to B8G8R8A8_UNORM (swapchain.rs:27-48)
            ctx.check(ctx.L.rt3_pass_present_bgra8(ctx.h, p, Rect(0, 0, x, y), self._ptr("Depth"), self._ptr("In"), self._ptr("Out"), ctx.stream()))
        elif self._shader == "structured_importance_sampling":     # shaders/old/structured_importance_sampling.slang, 8x8 threads per probe
            ctx.check(ctx.L.rt3_pass_probe_sis(ctx.h, p, self._ptr("gbuffer"), self._ptr("out"), ctx.stream()))
        elif self._shader == "spherical_harmonic_conversion":
            ctx.check(ctx.L.rt3_pass_probe_sh(ctx.h, p, self._ptr("probe_atlas"), self._ptr("out"), ctx.stream()))
        elif self._shader == "interpolate_probes":
            ctx.check(ctx.L.rt3_pass_probe_interpolate(ctx.h, p, self._ptr("gbuffer"), self._ptr("gbuffer_depth"), self._ptr("sh_coeficents"),
                                                       self._ptr("Light"), ctx.stream()))
        else:
            raise ValueError(f"unknown compute shader '{self._shader}' (known: postprocess, present, structured_importance_sampling, "
                             "spherical_harmonic_conversion, interpolate_probes)")
        return self.name


class Renderer:
    """The per-frame `commands` system (src/renderer/mod.rs:65-106) for the ray-tracing path:
    gbuffer pass -> reference-mode pass (ping-pong Light/PrevLight) -> optional postprocess."""

    def __init__(self, ctx, world, width, height):
        self.ctx, self.world = ctx, world
        self.rg = RenderGraph(ctx, world, (width, height))
        rg = self.rg
        rg.image("RGBA32UI", "gbuffer"); rg.image("R32F", "gbuffer_depth"); rg.image("RG32UI", "ids"); rg.image("RG32F", "bary")
        rg.image("RGBA32F", "light0"); rg.image("RGBA32F", "light1"); rg.image("RGBA32F", "out")
        self.cur = 0

    def frame(self, gconst, postprocess=False, debug_ids=False):
        rg = self.rg
        light, prev = f"light{self.cur}", f"light{1 - self.cur}"
        b = RayTracingPass.new(rg, "gbuffer").shader("gbuffer").constants(gconst).write("gbuffer", "gbuffer").write("gbuffer_depth", "gbuffer_depth")
        if debug_ids:
            b.write("ids", "ids").write("bary", "bary")
        b.launch(WorkSize2D.FullScreen())
        (RayTracingPass.new(rg, "reference").shader("refrence_mode").constants(gconst).read("gbuffer", "gbuffer").read("gbuffer_depth", "gbuffer_depth")
         .read("PrevLight", prev).write("Light", light).launch(WorkSize2D.FullScreen()))
        if postprocess:
            (ComputePass.new(rg, "postprocess").shader("postprocess").constants(gconst).read("Depth", "gbuffer_depth").read("In", light)
             .write("Out", "out").dispatch(WorkSize2D.FullScreen()))
        self.cur = 1 - self.cur
        rg.frame_number += 1
        return rg.tensor(light)

    def probe_frame(self, gconst, sis=False):
        """The screen-space probe path: gbuffer -> [structured_importance_sampling ->] trace_probes ->
        spherical_harmonic_conversion -> interpolate_probes, probe atlas ping-ponged for the temporal blend.  Returns the Light tensor.
        sis=True runs the direction-selection pass and binds its u16 image to trace_probes; otherwise the uniform assignment."""
        rg = self.rg
        p = np.ascontiguousarray(gconst)
        aw, ah, nsh = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        self.ctx.check(self.ctx.L.rt3_probe_layout(ptr(p), C.byref(aw), C.byref(ah), C.byref(nsh)))
        rg.buffer("probe_atlas0", (ah.value, aw.value, 4)); rg.buffer("probe_atlas1", (ah.value, aw.value, 4)); rg.buffer("probe_sh", (nsh.value, 9))
        self.probe_cur = getattr(self, "probe_cur", 0)
        atlas, prev = f"probe_atlas{self.probe_cur}", f"probe_atlas{1 - self.probe_cur}"
        (RayTracingPass.new(rg, "gbuffer").shader("gbuffer").constants(gconst).write("gbuffer", "gbuffer").write("gbuffer_depth", "gbuffer_depth")
         .launch(WorkSize2D.FullScreen()))
        tp = (RayTracingPass.new(rg, "trace_probes").shader("trace_probes").constants(gconst).read("gbuffer", "gbuffer").read("gbuffer_depth", "gbuffer_depth")
              .read("prev_probe_atlas", prev).write("probe_atlas", atlas))
        if sis:
            rg.buffer("probe_directions", (ah.value, aw.value), "int16")
            (ComputePass.new(rg, "sis").shader("structured_importance_sampling").constants(gconst).read("gbuffer", "gbuffer").write("out", "probe_directions")
             .dispatch(WorkSize2D.FractionalFullScreen(2, 2)))
            tp.read("directions", "probe_directions")
        tp.launch(WorkSize2D.FractionalFullScreen(2, 2))
        (ComputePass.new(rg, "sh").shader("spherical_harmonic_conversion").constants(gconst).read("probe_atlas", atlas).write("out", "probe_sh")
         .dispatch(WorkSize2D.FractionalFullScreen(2, 2)))
        (ComputePass.new(rg, "interpolate").shader("interpolate_probes").constants(gconst).read("gbuffer", "gbuffer").read("gbuffer_depth", "gbuffer_depth")
         .read("sh_coeficents", "probe_sh").write("Light", "light0").dispatch(WorkSize2D.FullScreen()))
        self.probe_cur = 1 - self.probe_cur
        rg.frame_number += 1
        return rg.tensor("light0")

    def render_frame_host(self, gconst, light_host, prev_light_host=None):
        """End-to-end path with HOST buffers (params in, framebuffer out)."""
        p = np.ascontiguousarray(gconst)
        self.ctx.check(self.ctx.L.rt3_render_frame_host(self.ctx.h, ptr(p), None if prev_light_host is None else C.c_void_p(prev_light_host),
                                                        C.c_void_p(light_host), None, None))


    def submit_frame_host(self, gconst, light_host, prev_light_host=None):
        """Frames in flight (FRAMES_IN_FLIGHT = 3, swapchain.rs:8): enqueue one frame and its read-back, return a ticket."""
        p = np.ascontiguousarray(gconst)
        t = C.c_uint64(0)
        self.ctx.check(self.ctx.L.rt3_frame_submit_host(self.ctx.h, ptr(p), None if prev_light_host is None else C.c_void_p(prev_light_host),
                                                        C.c_void_p(light_host), C.byref(t)))
        return t.value

    def wait_frame(self, ticket):
        self.ctx.check(self.ctx.L.rt3_frame_wait(self.ctx.h, ticket))


class TiledRenderer:
    """Rank-local renderer for screen-tile sharding (tiles.py): this rank's tiles of a W x H frame are rendered into
    tile-major slabs with no communication while samples accumulate; the finished image is DELIVERED in bands of rows — rank r ends up with
    rows tiles.band_rows()[r], the part it presents or reads back — by `exchange`:

      p2p   rt3_scatter_tiles_rgba32f: one kernel of ours stores every tile straight into its owner's band buffer over NVLink peer memory
            (torch symmetric memory provides the mapping and the barrier that follows); no collective library on the data path
      a2a   one NCCL all_to_all_single (every rank sends (N-1)/N of its slab, receives its band) + rt3_untile_rgba32f

    `gather` (all_gather of the slabs to EVERY rank + full-frame un-tile) is kept for callers that need the whole image on each GPU; it
    moves N times the bytes."""

    def __init__(self, ctx, world, width, height, rank, world_size, tile=64):
        torch = _torch()
        from . import tiles as T
        self.ctx, self.world = ctx, world
        self.W, self.H, self.rank, self.world_size, self.tile = width, height, rank, world_size, tile
        self.rects = np.ascontiguousarray(T.rank_tiles(width, height, rank, world_size, tile))
        self.all_rects = np.ascontiguousarray(T.all_tiles(width, height, world_size, tile))
        assert self.rects["w"].max() == tile and self.rects["h"].max() == tile, "every rank needs at least one full tile"
        n = len(self.rects)
        dev = f"cuda:{ctx.device}"
        self.dev = dev
        self.gbuffer = torch.zeros((n, tile, tile, 4), dtype=torch.int32, device=dev)
        self.depth = torch.zeros((n, tile, tile), dtype=torch.float32, device=dev)
        self.light = [torch.zeros((n, tile, tile, 4), dtype=torch.float32, device=dev) for _ in range(2)]
        self.cur = 0
        self.pixels = int((self.rects["w"].astype(np.int64) * self.rects["h"]).sum())
        # delivery in bands
        self.bands = T.band_rows(height, world_size, tile)
        self.row0, self.row1 = self.bands[rank]
        self.band_h_max = max(1, max(b - a for a, b in self.bands))
        self.counts = T.exchange_plan(width, height, world_size, tile)           # counts[s][d]: tiles of rank s inside the band of rank d
        self.n_real = int(self.counts[rank].sum())
        self.band_rects = np.ascontiguousarray(T.band_tiles(width, height, rank, world_size, tile))
        self.exchange_mode, self.exchange_note = None, ""
        self.side = torch.cuda.Stream(device=dev)                                 # delivery of frame i overlaps the rendering of frame i + 1
        self.ev_rendered_slab = [torch.cuda.Event(), torch.cuda.Event()]
        self.ev_delivered_slab = [None, None]
        self.ev_last_delivery = None
        self.gathered = None
        self.frame_buf = None

    # ---- rendering -------------------------------------------------------------------------------------------------------
    def frame(self, gconst):
        L, h, st = self.ctx.L, self.ctx.h, self.ctx.stream()
        p = np.ascontiguousarray(gconst)
        light, prev = self.light[self.cur], self.light[1 - self.cur]
        self.ctx.check(L.rt3_pass_gbuffer_tiles(h, ptr(p), ptr(self.rects), len(self.rects), self.gbuffer.data_ptr(), self.depth.data_ptr(), st))
        self.ctx.check(L.rt3_pass_reference_tiles(h, ptr(p), ptr(self.rects), len(self.rects), self.gbuffer.data_ptr(), self.depth.data_ptr(),
                                                   prev.data_ptr(), light.data_ptr(), st))
        self.cur = 1 - self.cur
        return light

    # ---- delivery --------------------------------------------------------------------------------------------------------
    def setup_exchange(self, mode="auto", slots=2):
        """Allocate `slots` band buffers ([band rows, W, 4] RGBA32F each) and choose the delivery path ('p2p', 'a2a' or 'auto' = p2p when
        torch symmetric memory can map the peers' buffers, agreed on by all ranks)."""
        torch = _torch()
        W, bh, N = self.W, self.band_h_max, self.world_size
        self.slots = slots
        per = bh * W * 4
        buf = None
        if N == 1:
            mode = "local"
        if mode in ("auto", "p2p"):
            import torch.distributed as dist
            ok, hdl = 1, None
            try:
                import torch.distributed._symmetric_memory as symm
                buf = symm.empty(slots * per, dtype=torch.float32, device=self.dev)
                hdl = symm.rendezvous(buf, dist.group.WORLD)
                ptrs = [int(x) for x in hdl.buffer_ptrs]
                assert len(ptrs) == N and all(ptrs)
            except Exception as e:  # noqa: BLE001
                ok, self.exchange_note = 0, f"symmetric memory unavailable: {type(e).__name__}: {e}"[:200]
            flag = torch.tensor([ok], dtype=torch.int32, device=self.dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)                           # every rank takes the same path
            if int(flag.item()) == 1:
                self.hdl, self.peer_ptrs = hdl, ptrs
                from . import tiles as T
                real = self.rects[: self.n_real]
                self.dst = []
                for k in range(slots):
                    d = np.zeros(self.n_real, np.uint64)
                    for i, r in enumerate(real):
                        owner = T.band_of_row(int(r["y0"]), self.H, N, self.tile)
                        d[i] = ptrs[owner] + 4 * (k * per + ((int(r["y0"]) - self.bands[owner][0]) * W + int(r["x0"])) * 4)
                    self.dst.append(d)
                mode = "p2p"
            elif mode == "p2p":
                raise RuntimeError("peer-memory delivery requested but " + (self.exchange_note or "another rank could not map it"))
            else:
                mode, buf = "a2a", None
        if buf is None:
            buf = torch.zeros(slots * per, dtype=torch.float32, device=self.dev)
        self.band_store = buf
        self.band = [buf[k * per:(k + 1) * per].view(bh, W, 4) for k in range(slots)]
        if mode == "a2a":
            T2 = self.tile * self.tile * 4
            self.in_splits = [int(c) * T2 for c in self.counts[self.rank]]
            self.out_splits = [int(self.counts[s][self.rank]) * T2 for s in range(N)]
            self.recv = torch.zeros(max(1, sum(self.out_splits)), dtype=torch.float32, device=self.dev)
        self.exchange_mode = mode
        return mode

    def exchange(self, light, slot=0):
        """Deliver the finished slab: afterwards (in stream order) band[slot][: row1 - row0] holds this rank's rows of the frame."""
        L, h = self.ctx.L, self.ctx.h
        T = self.tile
        if self.exchange_mode == "p2p":
            self.ctx.check(L.rt3_scatter_tiles_rgba32f(h, ptr(self.rects), ptr(self.dst[slot]), self.n_real, T, T, light.data_ptr(), self.W, self.ctx.stream()))
            self.hdl.barrier(channel=0)                                           # every rank's stores have landed before anyone reads its band
        elif self.exchange_mode == "a2a":
            import torch.distributed as dist
            send = light.view(-1)[: self.n_real * T * T * 4]
            dist.all_to_all_single(self.recv[: sum(self.out_splits)], send, self.out_splits, self.in_splits)
            self.ctx.check(L.rt3_untile_rgba32f(h, ptr(self.band_rects), len(self.band_rects), T, T, self.recv.data_ptr(), self.band[slot].data_ptr(),
                                                self.W, self.row1 - self.row0, self.ctx.stream()))
        elif self.exchange_mode == "local":
            self.ctx.check(L.rt3_untile_rgba32f(h, ptr(self.rects), len(self.rects), T, T, light.data_ptr(), self.band[slot].data_ptr(), self.W, self.H,
                                                self.ctx.stream()))
        else:
            raise RuntimeError("setup_exchange() has not been called")
        return self.band[slot][: self.row1 - self.row0]

    def frame_exchange_overlapped(self, gconst, slot):
        """frame() on the current stream, exchange() on the side stream: the delivery of frame i runs while frame i + 1 is traced.  The Light
        slab a frame writes is the one delivered two frames earlier, so the current stream first waits for that delivery.  Call drain()
        before reading a band or stopping a clock."""
        torch = _torch()
        cur = torch.cuda.current_stream()
        c = self.cur
        if self.ev_delivered_slab[c] is not None:
            cur.wait_event(self.ev_delivered_slab[c])
        light = self.frame(gconst)
        self.ev_rendered_slab[c].record(cur)
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.ev_rendered_slab[c])
            band = self.exchange(light, slot)
            ev = torch.cuda.Event()
            ev.record(self.side)
        self.ev_delivered_slab[c] = ev
        self.ev_last_delivery = ev
        return band

    def drain(self):
        if self.ev_last_delivery is not None:
            _torch().cuda.current_stream().wait_event(self.ev_last_delivery)

    def read_back(self, band, host_frame):
        """This rank's band -> rows [row0, row1) of a page-locked host frame [H, W, 4] (shared by the ranks of the node: every rank reads back
        its own rows over its own PCIe link).  Asynchronous on the current stream."""
        if self.row1 > self.row0:
            host_frame[self.row0:self.row1].copy_(band, non_blocking=True)

    # ---- the whole image on every rank (N times the traffic of exchange) --------------------------------------------------
    def gather(self, light):
        """One all_gather of the rank-local slabs + un-tile into the full frame (every rank gets the image)."""
        torch = _torch()
        n = len(self.rects)
        if self.frame_buf is None:
            self.frame_buf = torch.zeros((self.H, self.W, 4), dtype=torch.float32, device=self.dev)
        if self.world_size > 1:
            import torch.distributed as dist
            if self.gathered is None:
                self.gathered = torch.zeros((self.world_size * n, self.tile, self.tile, 4), dtype=torch.float32, device=self.dev)
            dist.all_gather_into_tensor(self.gathered, light)
            slab, rects = self.gathered, self.all_rects
        else:
            slab, rects = light, self.rects
        self.ctx.check(self.ctx.L.rt3_untile_rgba32f(self.ctx.h, ptr(rects), len(rects), self.tile, self.tile, slab.data_ptr(),
                                                     self.frame_buf.data_ptr(), self.W, self.H, self.ctx.stream()))
        return self.frame_buf
