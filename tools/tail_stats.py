#!/usr/bin/env python
"""The fixed cost of a persistent traversal launch and what it does to small per-rank frames.
  python tools/tail_stats.py sweep        # T(n) of rt3_trace_closest (CUDA events) + the instrumented kernel's ramp / tail times, n = 38 k ... 3.6 M rays
  python tools/tail_stats.py rankframe    # one rank's share (1/8 of the 64x64 tiles) of the 3840x2160 frame on one GPU: ms per frame, strips 1 / 2
Run the second under `ncu --metrics gpu__time_duration.sum` to list the kernels of such a frame (profiles/r2l_*)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.camera import camera_for, make_params  # noqa: E402
from raytracer3_b200.host import RenderContext, RenderWorld, TiledRenderer  # noqa: E402
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV  # noqa: E402


def setup():
    sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
    sc["bluenoise"], _ = scenes.reference_blue_noise()
    ctx = RenderContext(0)
    w = RenderWorld(ctx)
    w.load_scene(sc)
    w.create_acceleration_structure()
    return sc, ctx, w


def sweep():
    sc, ctx, w = setup()
    n_max = 3_600_000
    closest, _ = scenes.incoherent_rays(sc, n_max)
    rays = torch.from_numpy(closest.view(np.float32).reshape(n_max, 8)).cuda()
    hits = torch.empty((n_max, 4), dtype=torch.float32, device="cuda")
    rows = []
    for n in (37_888, 151_552, 450_000, 900_000, 1_800_000, 3_600_000):
        r = rays[:n]
        for _ in range(3):
            w.trace_closest(r, hits)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            w.trace_closest(r, hits)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        ts = [w.trace_tail_stats(r).as_dict() for _ in range(5)][2:]
        row = {"rays": n, "launch_us_events": 1e3 * ms, "mrays_s": n / ms / 1e3}
        for k in ("kernel_us", "mean_warp_busy_us", "mean_warp_tail_us", "tail_fraction"):
            row[k] = float(np.mean([t[k] for t in ts]))
        row["warps"] = ts[0]["warps"]
        rows.append(row)
        print(json.dumps(row), flush=True)
    a, b = rows[-3], rows[-1]                  # two-point fit T(n) = fixed + n / rate on the large launches
    rate = (b["rays"] - a["rays"]) / (b["launch_us_events"] - a["launch_us_events"])
    print(json.dumps({"fit": "T(n) = fixed + n / rate from the 0.9 M and 3.6 M launches", "rate_mrays_s": rate, "fixed_us": a["launch_us_events"] - a["rays"] / rate}))


def rankframe():
    sc, ctx, w = setup()
    W, H, N = 3840, 2160, 8
    R = TiledRenderer(ctx, w, W, H, 0, N, tile=64)
    cam = camera_for(sc, "cam0", W, H)
    ps = [make_params(cam, W, H, frame=n, bounces=4, samples=1, blendfactor=1.0 / (n + 1), flags=FLAG_ENV | FLAG_BLUENOISE) for n in range(16)]
    for strips in (1, 2):
        ctx.set_strips(strips)
        for p in ps[:4]:
            R.frame(p)
        torch.cuda.synchronize()
        l0 = ctx.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for p in ps:
            R.frame(p)
        e1.record()
        torch.cuda.synchronize()
        print(json.dumps({"config": f"rank 0 of 8 of a {W}x{H} frame ({R.pixels} pixels in {R.n_real} tiles), one GPU", "strips": strips,
                          "ms_per_frame": e0.elapsed_time(e1) / len(ps), "kernel_launches_per_frame": (ctx.kernel_launches() - l0) / len(ps)}), flush=True)


if __name__ == "__main__":
    {"sweep": sweep, "rankframe": rankframe}[sys.argv[1] if len(sys.argv) > 1 else "sweep"]()
