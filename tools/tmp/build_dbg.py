import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from raytracer3_b200 import scenes
from raytracer3_b200.host import RenderContext, RenderWorld
sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
if len(sys.argv) > 1: sc = scenes.instanced_grid(sc, int(sys.argv[1]))
ctx = RenderContext(0); w = RenderWorld(ctx); w.load_scene(sc)
for i in range(3):
    bs = w.create_acceleration_structure(1)
    print({k: round(v, 3) for k, v in bs.as_dict().items() if k.startswith("ms_")}, bs.wide_nodes, file=sys.stderr, flush=True)
