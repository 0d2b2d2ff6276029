#!/usr/bin/env python
"""Runs the BASELINE.json configs other than the bench headline (configs[0], [2], [3], [4]) on one GPU and prints one JSON line each.
  python tools/run_configs.py [cfg1] [cfg3] [cfg4] [cfg5]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import __graft_entry__ as g  # noqa: E402

g.build()
from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.camera import camera_for, make_params  # noqa: E402
from raytracer3_b200.host import RenderContext, Renderer, RenderWorld  # noqa: E402
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV, MISS  # noqa: E402

FLAGS = FLAG_ENV | FLAG_BLUENOISE


def timed(fn, reps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def atrium():
    sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
    sc["bluenoise"], _ = scenes.reference_blue_noise()
    return sc


def cfg1(ctx):
    """configs[0]: 1280x720 primary visibility, hit ids + depth, checked against the CPU oracle (lavapipe: unavailable on this box)."""
    from oracle.oracle_py import OracleScene
    sc = atrium()
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc); bs = w.create_acceleration_structure()
    W, H = 1280, 720
    p = make_params(camera_for(sc, "cam0", W, H), W, H, bounces=1, samples=1)
    r = Renderer(ctx, w, W, H)
    from raytracer3_b200.host import RayTracingPass, WorkSize2D
    def gb(i):
        (RayTracingPass.new(r.rg, "gbuffer").shader("gbuffer").constants(p).write("gbuffer", "gbuffer").write("gbuffer_depth", "gbuffer_depth")
         .write("ids", "ids").write("bary", "bary").launch(WorkSize2D.FullScreen()))
    for i in range(3):
        gb(i)
    ms = timed(gb, 20)
    ids = r.rg.tensor("ids").cpu().numpy().view(np.uint32)
    depth = r.rg.tensor("gbuffer_depth").cpu().numpy()
    bary = r.rg.tensor("bary").cpu().numpy()
    orc = OracleScene(sc)
    t0 = time.perf_counter()
    ogb, odp, oids, obary = orc.pass_gbuffer(p)
    cpu_s = time.perf_counter() - t0
    hit = oids[..., 0] != MISS
    return {"config": "cfg1 1280x720 primary visibility (synthetic_atrium)", "gpu_ms": ms, "gpu_mrays_s": W * H / ms / 1e3, "cpu_oracle_ms": cpu_s * 1e3,
            "cpu_threads": orc.threads, "ids_bit_exact": bool((ids == oids).all()), "depth_bit_exact": bool((depth.view(np.uint32) == odp.view(np.uint32)).all()),
            "bary_bit_exact": bool((bary.view(np.uint32) == obary.view(np.uint32)).all()), "hit_fraction": float(hit.mean()),
            "max_rel_depth_err": float((np.abs(depth[hit] - odp[hit]) / odp[hit]).max()), "triangles": bs.triangles}


def cfg3(ctx):
    """configs[2] on ONE GPU: 3840x2160, 64 spp progressive (64 frames x 1 spp, blendfactor 1/(n+1), ping-pong Light/PrevLight)."""
    sc = atrium()
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    W, H = 3840, 2160
    r = Renderer(ctx, w, W, H)
    cam = camera_for(sc, "cam0", W, H)
    def frame(n):
        r.frame(make_params(cam, W, H, frame=n, bounces=4, samples=1, blendfactor=1.0 / (n + 1), flags=FLAGS))
    for n in range(2):
        frame(n)
    ctx.counters(reset=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for n in range(64):
        frame(n)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    c = ctx.counters(reset=True)
    rays = c.rays_primary + c.rays_secondary
    light = r.rg.tensor(f"light{1 - r.cur}")
    return {"config": "cfg3 3840x2160 64 spp progressive, 1 GPU (synthetic_atrium)", "total_ms": ms, "ms_per_spp": ms / 64, "rays": int(rays),
            "mrays_s": rays / ms / 1e3, "spp_s": c.paths_started / (ms * 1e-3), "mean_radiance": float(light[..., :3].mean())}


def cfg4(ctx, log2n=25):
    """configs[3]: 2^25 closest + 2^25 any incoherent rays (surface-sampled, unsorted), traversal only, median of 20."""
    sc = atrium()
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc); bs = w.create_acceleration_structure()
    n = 1 << log2n
    t0 = time.perf_counter()
    closest, anyr = scenes.incoherent_rays(sc, n)
    gen_s = time.perf_counter() - t0
    dc = torch.from_numpy(closest.view(np.float32).reshape(n, 8)).cuda()
    da = torch.from_numpy(anyr.view(np.float32).reshape(n, 8)).cuda()
    hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    bits = torch.empty(((n + 31) // 32,), dtype=torch.int32, device="cuda")
    out = {"config": f"cfg4 2^{log2n} closest + 2^{log2n} any incoherent rays (synthetic_atrium)", "generate_s": gen_s}
    for name, fn in (("closest", lambda i: w.trace_closest(dc, hits)), ("any", lambda i: w.trace_any(da, bits))):
        for i in range(3):
            fn(i)
        ts = sorted(timed(fn, 1) for _ in range(20))
        med = ts[len(ts) // 2]
        out[f"{name}_ms_median"] = med
        out[f"{name}_mrays_s"] = n / med / 1e3
        out[f"{name}_hbm_gbs"] = n * (48 if name == "closest" else 32.125) / med / 1e6
    hit_frac = float((hits[:, 3].view(torch.int32) != -1).float().mean())
    occ = float(np.unpackbits(bits.cpu().numpy().view(np.uint8)).mean())
    ts_ = w.trace_stats(dc[: 1 << 20])
    out.update({"hit_fraction": hit_frac, "occluded_fraction": occ, "nodes_per_ray": ts_.nodes_per_ray, "tris_per_ray": ts_.tris_per_ray,
                "l2_bytes_per_ray": ts_.node_bytes_per_ray + ts_.tri_bytes_per_ray})
    return out


def cfg3_regen(ctx):
    """The 64 samples of configs[2] inside ONE reference-pass call (gconst.samples = 64, refrence_mode.slang:28): the wavefront with ray
    regeneration keeps one long queue (terminated paths are replaced in place by the pixel's next sample) instead of 64 x 3 shrinking ones,
    and the G-buffer pass runs once.  3840x2160 and 1920x1080; RT3_OPT_STRIPS = 1 and the default."""
    sc = atrium()
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    out = {"config": "64 spp in one frame call (regeneration), 1 GPU (synthetic_atrium)"}
    for W, H in ((1920, 1080), (3840, 2160)):
        r = Renderer(ctx, w, W, H)
        cam = camera_for(sc, "cam0", W, H)
        for strips in (1, 0):
            ctx.set_strips(strips)
            p = make_params(cam, W, H, frame=0, bounces=4, samples=64, flags=FLAGS)
            r.frame(make_params(cam, W, H, frame=0, bounces=4, samples=2, flags=FLAGS))
            ctx.counters(reset=True)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r.frame(p)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            c = ctx.counters(reset=True)
            rays = c.rays_primary + c.rays_secondary
            out[f"{W}x{H}_strips{strips or 'default'}"] = {"total_ms": ms, "rays": int(rays), "mrays_s": rays / ms / 1e3, "spp_s": c.paths_started / (ms * 1e-3),
                                                        "secondary_mrays_s": c.rays_secondary / ms / 1e3}
        ctx.set_strips(0)
        del r
    return out


def cfg5(ctx, grid=8):
    """configs[4] on ONE GPU: grid x grid instanced atrium, flattened single-level BVH and two-level (shared BLAS + TLAS over the
    instances): build time + 1080p 4-bounce path tracing for both; the two images must be bit-identical."""
    from raytracer3_b200.types import BUILD_TWO_LEVEL
    base = atrium()
    sc = scenes.instanced_grid(base, grid)
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc)
    W, H = 1920, 1080
    cam = camera_for(sc, "cam_grid", W, H)
    out = {"config": f"cfg5 {grid}x{grid} instanced synthetic_atrium, 1 GPU"}
    images = []
    for tag, flags in (("flattened", 1), ("two_level", 1 | BUILD_TWO_LEVEL)):
        bs = w.create_acceleration_structure(flags)
        bs2 = w.create_acceleration_structure(flags)      # second build: allocations warm
        r = Renderer(ctx, w, W, H)
        def frame(n):
            return r.frame(make_params(cam, W, H, frame=n, bounces=4, samples=1, flags=FLAGS))
        for n in range(3):
            frame(n)
        ctx.counters(reset=True)
        ms = timed(frame, 20)
        c = ctx.counters(reset=True)
        rays = (c.rays_primary + c.rays_secondary) / 20
        images.append(frame(5).clone())
        out[tag] = {"triangles": bs.triangles, "bvh_mb": (bs2.node_bytes + bs2.tri_bytes) / 1e6, "build_first": bs.as_dict(), "build_warm": bs2.as_dict(),
                    "frame_ms": ms, "rays_per_frame": rays, "mrays_s": rays / ms / 1e3}
        del r
    out["images_bit_identical"] = bool(torch.equal(images[0].view(torch.int32), images[1].view(torch.int32)))
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or ["cfg1", "cfg3", "cfg4", "cfg5"]
    ctx = RenderContext(0)
    for name in which:
        fn, arg = name, None
        if ":" in name:
            fn, arg = name.split(":")
        res = globals()[fn](ctx, int(arg)) if arg else globals()[fn](ctx)
        print(json.dumps(res), flush=True)
