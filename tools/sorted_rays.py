"""Closest-hit traversal time on a frame's own bounce rays in queue order against the same rays grouped by direction octant (stable, so the
pixel order survives inside a group) and by octant + Morton code of the origin: what ray re-ordering between bounces could buy at most."""
import os, sys
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


def timed(r):
    d = torch.from_numpy(np.ascontiguousarray(r).view(np.float32).reshape(-1, 8)).cuda()
    for _ in range(3):
        w.trace_closest(d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        w.trace_closest(d)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 10


def morton(o, lo, hi):
    q = np.clip(((o - lo) / (hi - lo) * 1023).astype(np.uint64), 0, 1023)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F; v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])


for b in range(3):
    r = rays[b][valid[b] != 0]
    f = r.view(np.float32).reshape(-1, 8)
    octant = (f[:, 4] < 0).astype(np.uint64) | ((f[:, 5] < 0).astype(np.uint64) << 1) | ((f[:, 6] < 0).astype(np.uint64) << 2)
    t0 = timed(r)
    t1 = timed(r[np.argsort(octant, kind="stable")])
    o = f[:, 0:3].astype(np.float64)
    key = (octant << 30) | morton(o, o.min(0), o.max(0) + 1e-9)
    t2 = timed(r[np.argsort(key, kind="stable")])
    t3 = timed(r[np.random.default_rng(1).permutation(len(r))])
    print(f"bounce {b + 1}: {len(r)} rays; queue order {t0:.4f} ms, by octant {t1:.4f} ms ({t1 / t0:.3f}), by octant + origin Morton {t2:.4f} ms ({t2 / t0:.3f}), shuffled {t3:.4f} ms ({t3 / t0:.3f})", flush=True)
