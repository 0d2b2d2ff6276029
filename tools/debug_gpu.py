"""Ad-hoc GPU diagnostics (not a pytest file)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import __graft_entry__ as g
g.build()
from raytracer3_b200 import scenes
from raytracer3_b200.host import RenderContext, RenderWorld
from oracle.oracle_py import OracleScene
from emu_lib import EmuBvh

ctx = RenderContext(0)
for name, sc in (("cornell", scenes.cornell_box()), ("soup", scenes.random_soup(4000, seed=5))):
    w = RenderWorld(ctx); w.clear(); w.load_scene(sc)
    bs = w.create_acceleration_structure()
    print(name, bs.as_dict())
    orc = OracleScene(sc)
    nodes, tris = w.download_accel()
    T = orc.world_triangles()
    prims = tris[:, 3]
    print(" prim ids unique:", len(np.unique(prims)) == len(prims), "max", prims.max(), "n", len(prims))
    v0 = tris[:, 0:3].view(np.float32)
    ok = prims < len(T)
    print(" v0 matches oracle:", np.array_equal(v0[ok], T[prims[ok], 0]))
    emu = EmuBvh(nodes=nodes, tri_records=tris)
    c, a = scenes.incoherent_rays(sc, 5000, seed=11)
    rays = np.concatenate([c, scenes.random_rays(sc, 5000, seed=12)])
    want = orc.trace_closest(rays)
    cpu_on_gpu_bvh, _ = emu.trace_closest(rays)
    got = w.trace_closest_host(rays)
    print(" CPU traversal of GPU-built BVH == oracle:", (cpu_on_gpu_bvh == want).all(), (cpu_on_gpu_bvh["prim"] != want["prim"]).sum())
    print(" GPU traversal == oracle:", (got == want).all(), (got["prim"] != want["prim"]).sum())
    print(" GPU traversal == CPU traversal of same BVH:", (got == cpu_on_gpu_bvh).all())
    bad = np.nonzero(got["prim"] != want["prim"])[0][:5]
    for i in bad:
        print("  ray", i, rays[i], "got", got[i], "want", want[i])
    print(" nodes[0]:", [hex(x) for x in nodes[0]])
