#!/usr/bin/env python
"""Measurement of the SURVEY 8(f) "next" rows on one GPU, bench scene (synthetic_atrium, 1920x1080 unless stated), one JSON line each:
  F1  post-process (AgX + sky background): ms per pass, algorithmic bytes (36 B per pixel) against the measured HBM peak
  F2  the path tracer with RT3_FLAG_SPECULAR, with RT3_FLAG_NEE, and both: ms per frame, Mrays/s (all ray kinds counted)
  F4  screen-space probe frame: gbuffer + trace_probes + spherical_harmonic_conversion + interpolate_probes, ms per pass
Timing: CUDA events on the launch stream, 5 warm-ups, 30 repetitions."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.camera import camera_for, make_params  # noqa: E402
from raytracer3_b200.host import ComputePass, RayTracingPass, RenderContext, Renderer, RenderWorld, WorkSize2D  # noqa: E402
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV, FLAG_NEE, FLAG_SPECULAR  # noqa: E402


def timed(fn, reps=30, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    sc = bench.make_scene()
    ctx = RenderContext(0)
    w = RenderWorld(ctx); w.load_scene(sc); w.create_acceleration_structure()
    peak, peak_src = bench.hbm_peak()
    base = FLAG_ENV | FLAG_BLUENOISE
    for (W, H) in ((1920, 1080), (3840, 2160)):
        r = Renderer(ctx, w, W, H)
        cam = camera_for(sc, "cam0", W, H)
        p = make_params(cam, W, H, frame=0, bounces=4, samples=1, flags=base)
        light = r.frame(p)
        name = f"light{1 - r.cur}"

        def post(i):
            (ComputePass.new(r.rg, "postprocess").shader("postprocess").constants(p).read("Depth", "gbuffer_depth").read("In", name)
             .write("Out", "out").dispatch(WorkSize2D.FullScreen()))
        ms = timed(post)
        gbs = W * H * 36 / ms / 1e6
        print(json.dumps({"row": "F1 postprocess (AgX + sky background)", "size": f"{W}x{H}", "ms": ms, "algorithmic_bytes_per_pixel": 36,
                          "achieved_gbs": gbs, "hbm_peak_gbs": peak, "frac": gbs / peak, "peak_source": peak_src}), flush=True)
        del r
    W, H = 1920, 1080
    cam = camera_for(sc, "cam0", W, H)
    r = Renderer(ctx, w, W, H)
    for tag, flags in (("diffuse (bench headline)", base), ("F2 RT3_FLAG_SPECULAR", base | FLAG_SPECULAR), ("RT3_FLAG_NEE", base | FLAG_NEE),
                       ("F2 RT3_FLAG_SPECULAR | RT3_FLAG_NEE", base | FLAG_SPECULAR | FLAG_NEE)):
        ps = [make_params(cam, W, H, frame=i, bounces=4, samples=1, flags=flags) for i in range(40)]
        timed(lambda i: r.frame(ps[i]), reps=5)
        ctx.counters(reset=True)
        ms = timed(lambda i: r.frame(ps[i]), reps=30, warm=0)
        c = ctx.counters(reset=True)
        rays = (c.rays_primary + c.rays_secondary + c.rays_shadow) / 30
        print(json.dumps({"row": "path tracer 1920x1080, bounces=4, 1 spp, ENV|BLUENOISE + " + tag, "ms_per_frame": ms, "rays_per_frame": rays,
                          "shadow_rays_per_frame": c.rays_shadow / 30, "mrays_s": rays / ms / 1e3}), flush=True)
    # F4: the probe frame, pass by pass
    p = make_params(cam, W, H, frame=0, bounces=4, samples=1, flags=base)
    r.probe_frame(p)
    rg = r.rg
    atlas, prev = "probe_atlas0", "probe_atlas1"
    passes = {
        "gbuffer": lambda i: RayTracingPass.new(rg, "gbuffer").shader("gbuffer").constants(p).write("gbuffer", "gbuffer").write("gbuffer_depth", "gbuffer_depth")
        .launch(WorkSize2D.FullScreen()),
        "trace_probes": lambda i: RayTracingPass.new(rg, "trace_probes").shader("trace_probes").constants(p).read("gbuffer", "gbuffer")
        .read("gbuffer_depth", "gbuffer_depth").read("prev_probe_atlas", prev).write("probe_atlas", atlas).launch(WorkSize2D.FractionalFullScreen(2, 2)),
        "spherical_harmonic_conversion": lambda i: ComputePass.new(rg, "sh").shader("spherical_harmonic_conversion").constants(p).read("probe_atlas", atlas)
        .write("out", "probe_sh").dispatch(WorkSize2D.FractionalFullScreen(2, 2)),
        "interpolate_probes": lambda i: ComputePass.new(rg, "interpolate").shader("interpolate_probes").constants(p).read("gbuffer", "gbuffer")
        .read("gbuffer_depth", "gbuffer_depth").read("sh_coeficents", "probe_sh").write("Light", "light0").dispatch(WorkSize2D.FullScreen()),
    }
    out = {"row": "F4 screen-space probe frame 1920x1080 (one probe per 16x16 pixels, 8x8 directions: 518,400 probe rays)"}
    for k, fn in passes.items():
        out[k + "_ms"] = timed(fn)
    out["frame_ms"] = timed(lambda i: r.probe_frame(p))
    r.probe_frame(p, sis=True)
    out["structured_importance_sampling_ms"] = timed(lambda i: ComputePass.new(rg, "sis").shader("structured_importance_sampling").constants(p).read("gbuffer", "gbuffer")
                                                     .write("out", "probe_directions").dispatch(WorkSize2D.FractionalFullScreen(2, 2)))
    out["frame_with_sis_ms"] = timed(lambda i: r.probe_frame(p, sis=True))
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
