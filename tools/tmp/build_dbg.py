import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from raytracer3_b200 import scenes
from raytracer3_b200.host import RenderContext, RenderWorld
sc = scenes.instanced_grid(scenes.synthetic_atrium(scenes.SPONZA_DETAIL), 8)
ctx = RenderContext(0); w = RenderWorld(ctx); w.load_scene(sc)
for i in range(4):
    print("=== build", i, file=sys.stderr, flush=True)
    bs = w.create_acceleration_structure(1)
    print("ms_total", bs.ms_total, file=sys.stderr, flush=True)
