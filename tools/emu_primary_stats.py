"""Primary (camera) rays of the bench frame on the CPU warp emulation: the static assignment that k_gbuffer uses (a warp keeps its 8 x 4
pixels until the slowest lane is done) against the dynamic ray fetch of k_trace_dyn, on the same rays in the same order.  Test
infrastructure (uses the oracle's primary-ray function for the rays); no GPU needed:  python tools/emu_primary_stats.py [tiles]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from emu_lib import EmuBvh  # noqa: E402
from oracle import oracle_py  # noqa: E402
from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.camera import camera_for, make_params  # noqa: E402
from raytracer3_b200.types import RAY, ptr  # noqa: E402

CN, CR = 258.0, 95.0     # static SASS instructions of a node step / a record step of k_trace_dyn


def main():
    n_tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    W, H = 1920, 1080
    sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    b = EmuBvh(tris=T)
    p = make_params(camera_for(sc, "cam0", W, H), W, H)
    L = oracle_py.lib()
    rng = np.random.default_rng(1)
    tiles_x, tiles_y = W // 8, H // 4
    # runs of 16 horizontally adjacent 8x4 tiles (a 128-thread block of k_gbuffer covers 4), spread over the frame
    starts = rng.choice(tiles_x * tiles_y // 16, size=max(n_tiles // 16, 1), replace=False)
    rays = np.zeros(len(starts) * 16 * 32, RAY)
    r1 = np.zeros(1, RAY)
    k = 0
    for s in starts:
        for t in range(int(s) * 16, int(s) * 16 + 16):
            tx, ty = t % tiles_x, t // tiles_x
            for j in range(32):
                L.orc_primary_ray(ptr(p), tx * 8 + (j & 7), ty * 4 + (j >> 3), ptr(r1))
                rays[k] = r1[0]
                k += 1
    want, st = b.trace_closest(rays)
    print(f"{len(rays)} primary rays ({len(rays) // 32} warps of 8x4 pixels), nodes/ray {st[0]:.2f}, triangle tests/ray {st[1]:.2f}, "
          f"hit fraction {(want['prim'] != 0xFFFFFFFF).mean():.3f}")
    got, _ = b.kernel_trace(rays, static=True, blocks=1)
    assert (got == want).all()
    ph = b.phase_stats.astype(np.float64)
    n = len(rays)
    s_static = ph[5][0] * CN + ph[6][0] * CR
    print(f"static  (traverse_wide, as k_gbuffer): node steps/ray {ph[5][0] * 32 / n:6.2f} at {ph[5][1] / ph[5][0]:5.2f} lanes | record steps/ray "
          f"{ph[6][0] * 32 / n:6.2f} at {ph[6][1] / ph[6][0]:5.2f} lanes | instruction slots/ray {s_static / n:6.1f}")
    for defs, bal in (((), False), (("RT3_TRI_WAIT=8",), False), (("RT3_TRI_WAIT=16",), True)):
        got, _ = b.kernel_trace(rays, blocks=1, defines=defs, balance=bal)
        assert (got == want).all()
        ph = b.phase_stats.astype(np.float64)
        rec = ph[3] if bal else ph[1]
        s_dyn = ph[0][0] * CN + rec[0] * (165.0 if bal else CR) + ph[4][0] * 190.0
        print(f"dynamic (k_trace_dyn {' '.join(defs) + (' BALANCE' if bal else ''):24s}): node steps/ray {ph[0][0] * 32 / n:6.2f} at {ph[0][1] / ph[0][0]:5.2f} lanes | record steps/ray "
              f"{rec[0] * 32 / n:6.2f} at {rec[1] / rec[0]:5.2f} lanes | refills {int(ph[4][0])} | instruction slots/ray {s_dyn / n:6.1f}")


if __name__ == "__main__":
    main()
