#!/usr/bin/env python
"""Stage times of rt3_accel_build on the bench scene:  python tools/build_times.py [grid] [fast | rotations]
grid = n for the n x n instanced grid (8 = BASELINE configs[4], 16.8 M triangles, flattened); fast = RT3_BUILD_FAST_BUILD.
Prints the warm builds (the first call grows the memory pool).  RT3_BUILD_DEBUG=1 adds wall-clock marks per stage."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.host import RenderContext, RenderWorld  # noqa: E402
from raytracer3_b200.types import BUILD_FAST_BUILD, BUILD_FAST_TRACE, BUILD_ROTATIONS  # noqa: E402

args = sys.argv[1:]
sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
grid = next((int(a) for a in args if a.isdigit()), 1)
if grid > 1:
    sc = scenes.instanced_grid(sc, grid)
flags = BUILD_FAST_BUILD if "fast" in args else (BUILD_ROTATIONS if "rotations" in args else BUILD_FAST_TRACE)
ctx = RenderContext(0)
w = RenderWorld(ctx)
w.load_scene(sc)
for i in range(4):
    bs = w.create_acceleration_structure(flags)
    print("first call" if i == 0 else "warm      ", {k: round(v, 3) for k, v in bs.as_dict().items() if k.startswith("ms_")}, "triangles", bs.triangles, "wide nodes", bs.wide_nodes, flush=True)
