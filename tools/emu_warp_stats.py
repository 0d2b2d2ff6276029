"""SIMD statistics of the production traversal kernel (csrc/trace_kernel.cuh) on the CPU warp emulation (tests/emu/warp_emu.cpp):
how often a warp runs the node phase / a record step, with how many lanes, for the default loop and for the warp-balanced
record phase (-DRT3_TRI_BALANCE=1).  No GPU needed:  python tools/emu_warp_stats.py [detail] [rays] [--frame]   (--frame: the bench frame's own bounce rays, from the CPU oracle)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from emu_lib import EmuBvh  # noqa: E402
from raytracer3_b200 import scenes  # noqa: E402

# warp-instruction estimates per execution of a phase (from the SASS of k_trace_dyn<0,28,0>: node step incl. pop, one Moeller-Trumbore record
# step; the balanced step adds a 5-step prefix sum, 9 ray shuffles, the dispatch and merge loops)
VARIANTS = ([(False, ()), (False, ("RT3_TRI_REPS=1",))] + [(False, (f"RT3_TRI_WAIT={t}",)) for t in (6, 8, 10)]
            + [(True, ())] + [(True, (f"RT3_TRI_WAIT={t}",)) for t in (12, 16)])
# warp instructions per execution of a phase, from the SASS of k_trace_dyn<0,28,0> (node step incl. pop 258, one Moeller-Trumbore record step 95,
# balanced step 226, refill 150).  "time" model: node : record = 4.1 : 1 reproduces the measured RT3_TRI_REPS = 1 / 2 ratio of the launch
# (0.507 / 0.497 ms on B200) — a node step also waits for five dependent 16-byte loads.
COST = {"node": 258.0, "record": 95.0, "balanced": 226.0, "refill": 150.0}
NODE_TIME = 390.0


def frame_rays(sc, W=480, H=270, n_max=16384):
    """The bench frame's own secondary rays (4 bounces, ENV | BLUENOISE) from the CPU oracle, per bounce in pixel order — the order in
    which the wavefront queues hold them, up to the warp-aggregated compaction — at a reduced resolution; a contiguous run of each."""
    from oracle.oracle_py import OracleScene
    from raytracer3_b200.camera import camera_for, make_params
    from raytracer3_b200.types import RAY, ptr
    sc = dict(sc)
    sc["bluenoise"] = scenes.blue_noise_tile()
    orc = OracleScene(sc)
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=0, bounces=4, samples=1, flags=3)
    gb, dp, _, _ = orc.pass_gbuffer(p)
    rays, valid = np.zeros((3, W * H), RAY), np.zeros((3, W * H), np.uint8)
    orc.L.orc_set_ray_capture(ptr(rays), ptr(valid), W * H, 3)
    orc.pass_reference(p, gb, dp)
    orc.L.orc_set_ray_capture(None, None, 0, 0)
    out = []
    for b in range(3):
        r = rays[b][valid[b] != 0]
        start = max(0, (len(r) - n_max) // 2)
        out.append(np.ascontiguousarray(r[start:start + n_max]))
    return out


def main():
    frame = "--frame" in sys.argv
    argv = [a for a in sys.argv[1:] if a != "--frame"]
    detail = float(argv[0]) if len(argv) > 0 else 0.8
    n_rays = int(argv[1]) if len(argv) > 1 else 16384
    sc = scenes.synthetic_atrium(detail, environment=frame)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    b = EmuBvh(tris=T)
    if frame:
        sets = [(f"bounce{k + 1}", r, False) for k, r in enumerate(frame_rays(sc, n_max=n_rays))]
    else:
        c, a = scenes.incoherent_rays(sc, n_rays, seed=0x5EED0004, tris=T.astype(np.float64))
        sets = [("closest", c, False), ("any", a, True)]
    for name, rays, any_hit in sets:
        n_rays = len(rays)
        want, st = b.trace_closest(rays)
        print(f"{name}: {len(T)} triangles, {n_rays} rays, nodes/ray {st[0]:.2f}, triangle tests/ray {st[1]:.2f}, hit fraction {(want['prim'] != 0xFFFFFFFF).mean():.3f}")
        for balance, defs in VARIANTS:
            got, _ = b.kernel_trace(rays, any_hit=any_hit, blocks=1, balance=balance, defines=defs)
            if not any_hit:
                assert (got == want).all()
            ph = b.phase_stats.astype(np.float64)
            node_x, node_l = ph[0][0], ph[0][1]
            if balance:
                rec_x, rec_l = ph[3][0], ph[3][1]
                rest = rec_x * COST["balanced"] + ph[2][2] * 25.0 + ph[4][0] * COST["refill"]          # + the prefix sum every iteration
            else:
                rec_x, rec_l = ph[1][0], ph[1][1]
                rest = rec_x * COST["record"] + ph[4][0] * COST["refill"]
            slots, tslots = node_x * COST["node"] + rest, node_x * NODE_TIME + rest
            print(f"{name:8s} balance={int(balance)} {' '.join(defs):16s}: node steps/ray {node_x * 32 / n_rays:6.2f} at {node_l / max(node_x, 1):5.2f} lanes | "
                  f"record steps/ray {rec_x * 32 / n_rays:6.2f} at {rec_l / max(rec_x, 1):5.2f} lanes | refills {int(ph[4][0])} | "
                  f"warp instructions per ray {slots / n_rays:6.1f} (time model {tslots / n_rays:6.1f})")


if __name__ == "__main__":
    main()
