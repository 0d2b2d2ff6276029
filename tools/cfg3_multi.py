#!/usr/bin/env python
"""BASELINE.json configs[2] as written: 3840x2160, 64 spp progressive accumulation (64 frames x 1 spp, blendfactor 1/(n+1),
Light / PrevLight ping-pong per rank), 64x64 screen tiles sharded over the ranks, ONE NCCL gather of the framebuffer after frame 63.
Strong scaling: the frame is fixed, the ranks split it.

  python tools/cfg3_multi.py                                                     # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/cfg3_multi.py [--verify]

Prints one JSON line on rank 0: wall time of the 64 frames + gather (CUDA events, max over ranks), rays, Mrays/s, spp/s, per-rank ray
counts (load balance).  --verify: rank 0 also renders the same 64 frames alone, un-tiled, and compares the gathered image bit for bit."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--w", type=int, default=3840)
    ap.add_argument("--h", type=int, default=2160)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--verify", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from raytracer3_b200 import scenes
    from raytracer3_b200.camera import camera_for, make_params
    from raytracer3_b200.host import RenderContext, Renderer, RenderWorld, TiledRenderer
    from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV
    flags = FLAG_ENV | FLAG_BLUENOISE
    sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
    sc["bluenoise"] = scenes.blue_noise_tile()
    ctx = RenderContext(local)
    w = RenderWorld(ctx); w.load_scene(sc); bs = w.create_acceleration_structure()      # scene + BVH replicated: every rank builds its own
    W, H = a.w, a.h
    cam = camera_for(sc, "cam0", W, H)
    params = [make_params(cam, W, H, frame=n, bounces=4, samples=1, blendfactor=1.0 / (n + 1), flags=flags) for n in range(a.spp)]
    R = TiledRenderer(ctx, w, W, H, rank, world, tile=64)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run():
        R.cur = 0
        for n in range(a.spp):
            light = R.frame(params[n])                       # accumulation stays rank-local: no communication per sample
        return R.gather(light)                               # one all_gather + un-tile per output image

    for n in range(2):                                       # warm-up (allocations, NCCL channel set-up)
        R.frame(params[n])
    R.gather(R.light[0])
    barrier()
    ctx.counters(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    frame = run()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    c = ctx.counters(reset=True)
    stats = torch.tensor([ms, float(c.rays_primary + c.rays_secondary), float(c.paths_started)], dtype=torch.float64, device=f"cuda:{local}")
    per_rank = [stats.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, stats)
    ok = None
    if a.verify and rank == 0:
        r1 = Renderer(ctx, w, W, H)
        for n in range(a.spp):
            alone = r1.frame(params[n])
        ok = bool(torch.equal(alone.view(torch.int32), frame.view(torch.int32)))
    if rank == 0:
        ms_max = max(float(s[0]) for s in per_rank)
        rays = sum(float(s[1]) for s in per_rank)
        paths = sum(float(s[2]) for s in per_rank)
        rr = [float(s[1]) for s in per_rank]
        print(json.dumps({"config": f"cfg3 {W}x{H} {a.spp} spp progressive, 64x64 tiles over {world} GPU(s), one gather after the last frame (synthetic_atrium)",
                          "n_gpus": world, "total_ms": ms_max, "ms_per_spp": ms_max / a.spp, "rays": int(rays), "mrays_s": rays / ms_max / 1e3,
                          "spp_s": paths / (ms_max * 1e-3), "rank_ms": [round(float(s[0]), 2) for s in per_rank],
                          "rank_rays_max_over_mean": max(rr) / (sum(rr) / len(rr)), "tiles_per_rank": int(len(R.rects)),
                          "gathered_equals_single_gpu_image": ok, "mean_radiance": float(frame[..., :3].mean()), "triangles": int(bs.triangles)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
