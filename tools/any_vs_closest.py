"""Closest-hit against any-hit traversal time on a frame's own last-bounce rays (captured from the CPU oracle at full size, in queue order)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle.oracle_py import OracleScene
from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.host import RenderContext, RenderWorld
from raytracer3_b200.types import BUILD_FAST_TRACE, RAY, ptr
sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
sc["bluenoise"], _ = scenes.reference_blue_noise()
W, H = 1920, 1080
orc = OracleScene(sc)
p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=0, bounces=4, samples=1, flags=3)
gb, dp, _, _ = orc.pass_gbuffer(p)
rays, valid = np.zeros((3, W * H), RAY), np.zeros((3, W * H), np.uint8)
orc.L.orc_set_ray_capture(ptr(rays), ptr(valid), W * H, 3)
orc.pass_reference(p, gb, dp)
orc.L.orc_set_ray_capture(None, None, 0, 0)
ctx = RenderContext(0)
w = RenderWorld(ctx)
w.load_scene(sc)
w.create_acceleration_structure(BUILD_FAST_TRACE)
for b in range(3):
    r = np.ascontiguousarray(rays[b][valid[b] != 0])
    d = torch.from_numpy(r.view(np.float32).reshape(-1, 8)).cuda()
    res = {}
    for name, fn in (("closest", w.trace_closest), ("any", w.trace_any)):
        for _ in range(3):
            fn(d)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            out = fn(d)
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 10
    hits = w.trace_closest(d).cpu().numpy().view(np.uint32)[:, 3]
    print(f"bounce {b + 1}: {len(r)} rays, closest {res['closest']:.4f} ms, any {res['any']:.4f} ms, ratio {res['any'] / res['closest']:.3f}, hit fraction {(hits != 0xFFFFFFFF).mean():.3f}", flush=True)
