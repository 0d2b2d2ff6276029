"""SIMD statistics of the production traversal kernel (csrc/trace_kernel.cuh) on the CPU warp emulation (tests/emu/warp_emu.cpp):
how often a warp runs the node phase / a record step, with how many lanes, for the default loop and for the warp-balanced
record phase (-DRT3_TRI_BALANCE=1).  No GPU needed:  python tools/emu_warp_stats.py [detail] [rays]"""
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
VARIANTS = ([(False, ()), (False, ("RT3_TRI_REPS=1",)), (False, ("RT3_TRI_REPS=3",))] + [(False, (f"RT3_TRI_WAIT={t}",)) for t in (6, 8, 10)]
            + [(True, ())] + [(True, (f"RT3_TRI_WAIT={t}",)) for t in (8, 12, 16, 24)])
# node : record = 4.1 : 1 reproduces the measured RT3_TRI_REPS = 1 / 2 ratio of the traversal launch (0.507 / 0.497 ms on B200)
COST = {"node": 390.0, "record": 95.0, "balanced": 165.0, "refill": 190.0}


def main():
    detail = float(sys.argv[1]) if len(sys.argv) > 1 else 0.8
    n_rays = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
    sc = scenes.synthetic_atrium(detail, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, a = scenes.incoherent_rays(sc, n_rays, seed=0x5EED0004, tris=T.astype(np.float64))
    b = EmuBvh(tris=T)
    want, st = b.trace_closest(c)
    print(f"scene: {len(T)} triangles, {n_rays} incoherent rays, nodes/ray {st[0]:.2f}, triangle tests/ray {st[1]:.2f}")
    for name, rays, any_hit in (("closest", c, False), ("any", a, True)):
        for balance, defs in VARIANTS:
            got, _ = b.kernel_trace(rays, any_hit=any_hit, blocks=1, balance=balance, defines=defs)
            if not any_hit:
                assert (got == want).all()
            ph = b.phase_stats.astype(np.float64)
            node_x, node_l = ph[0][0], ph[0][1]
            if balance:
                rec_x, rec_l = ph[3][0], ph[3][1]
                slots = node_x * COST["node"] + rec_x * COST["balanced"] + ph[4][0] * COST["refill"]
            else:
                rec_x, rec_l = ph[1][0], ph[1][1]
                slots = node_x * COST["node"] + rec_x * COST["record"] + ph[4][0] * COST["refill"]
            print(f"{name:8s} balance={int(balance)} {' '.join(defs):16s}: node steps/ray {node_x * 32 / n_rays:6.2f} at {node_l / max(node_x, 1):5.2f} lanes | "
                  f"record steps/ray {rec_x * 32 / n_rays:6.2f} at {rec_l / max(rec_x, 1):5.2f} lanes | refills {int(ph[4][0])} | "
                  f"est. issue slots per ray {slots / n_rays:7.1f}")


if __name__ == "__main__":
    main()
