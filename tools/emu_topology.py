"""Traversal work per incoherent ray in the CPU emulator of the device builder / traversal (tests/emu), for builder variants.
Work = node visits + 0.4 x triangle tests (the cost ratio measured on B200).  No GPU needed:
    python tools/emu_topology.py [detail] [rays] [lbvh,split:BETA,ploc:RADIUS,treelet:PASSES,...]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from emu_lib import EmuBvh  # noqa: E402
from raytracer3_b200 import scenes  # noqa: E402


def main():
    detail = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    n_rays = int(sys.argv[2]) if len(sys.argv) > 2 else 40000
    variants = sys.argv[3].split(",") if len(sys.argv) > 3 else ["lbvh", "split:0.2", "ploc:8", "ploc:16", "ploc:32"]
    sc = scenes.synthetic_atrium(detail, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, a = scenes.incoherent_rays(sc, n_rays, seed=0x5EED0004, tris=T.astype(np.float64))
    base = None
    for var in variants:
        opts = dict(split_beta=0.0, ploc_radius=0, treelet_passes=0)
        for part in var.split("+"):                                     # e.g. ploc:4+treelet:2
            kind, _, arg = part.partition(":")
            if kind == "split":
                opts["split_beta"] = float(arg)
            elif kind == "ploc":
                opts["ploc_radius"] = int(arg)
            elif kind == "treelet":
                opts["treelet_passes"] = int(arg)
        t0 = time.time()
        b = EmuBvh(tris=T, **opts)
        tb = time.time() - t0
        info = b.info()
        hits, st = b.trace_closest(c)
        if base is None:
            base = hits
        same = bool((hits == base).all())
        work = st[0] + 0.4 * st[1]
        print(f"{var:22s} rounds={b.ploc_rounds():3d} tris={len(T)} refs={info['tris']} (+{info['tris'] / len(T) - 1:.1%}) wide_nodes={info['nodes']} "
              f"nodes/ray={st[0]:.3f} tris/ray={st[1]:.3f} work={work:.3f} sah_nodes={info['sah_nodes']:.0f} sah_tris={info['sah_tris']:.0f} hits_identical={same} build_s={tb:.1f}", flush=True)


if __name__ == "__main__":
    main()
