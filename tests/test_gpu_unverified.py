"""GPU tests of opt-in features that were written after round 1's GPU minutes ran out: they have been checked on the CPU emulation of the
device logic (tests/test_emu_bvh.py, tests/test_emu_warp.py) but never run on hardware, so they do not take part in the default `-m gpu`
run.  First GPU call of the next round:  RT3_TEST_UNVERIFIED=1 python -m pytest tests/test_gpu_unverified.py -m gpu -x -q
Once green, move them into test_gpu_edges.py and delete this file."""
import os
import sys

import numpy as np
import pytest

from raytracer3_b200 import scenes
from raytracer3_b200.types import MISS

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.environ.get("RT3_TEST_UNVERIFIED"), reason="not yet run on hardware; set RT3_TEST_UNVERIFIED=1")]


def test_treelet_build_gives_identical_hits_and_the_emulators_tree(gpu_ctx):
    """RT3_BUILD_TREELET: same hits bit for bit as the default build, every triangle kept once, and exactly the tree the CPU emulation of
    the builder produces (node count and visit statistics), with less traversal work than the rotations."""
    import torch
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from emu_lib import EmuBvh
    from raytracer3_b200.host import RenderWorld
    from raytracer3_b200.types import BUILD_FAST_TRACE, BUILD_TREELET, BUILD_TWO_LEVEL
    sc = scenes.synthetic_atrium(0.05, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc)
    closest, anyr = scenes.incoherent_rays(sc, 100000, seed=9, tris=T.astype(np.float64))
    rays = torch.from_numpy(np.concatenate([closest, scenes.random_rays(sc, 20000, seed=10)]).view(np.float32).reshape(-1, 8)).cuda()
    arays = torch.from_numpy(anyr.view(np.float32).reshape(-1, 8)).cuda()
    res = {}
    for name, flags in (("rot", BUILD_FAST_TRACE), ("treelet", BUILD_FAST_TRACE | BUILD_TREELET), ("treelet2", BUILD_FAST_TRACE | BUILD_TREELET | BUILD_TWO_LEVEL)):
        bs = w.create_acceleration_structure(flags)
        st = w.trace_stats(rays)
        res[name] = (bs.wide_nodes, w.trace_closest(rays).cpu().numpy().view(np.uint32), w.trace_any(arays).cpu().numpy(), st.nodes_per_ray + 0.4 * st.tris_per_ray)
        if not flags & BUILD_TWO_LEVEL:
            nodes, tris = w.download_accel()
            assert EmuBvh(nodes=nodes, tri_records=tris).check_triangles() == 0
    for name in ("treelet", "treelet2"):
        assert (res[name][1] == res["rot"][1]).all() and (res[name][2] == res["rot"][2]).all(), name
    emu = EmuBvh(tris=T, treelet_passes=2)
    assert res["treelet"][0] == emu.info()["nodes"]                       # the device build is the emulator's build
    assert res["treelet"][3] < res["rot"][3]                               # and it pays: fewer node visits + 0.4 x triangle tests per ray
    assert (res["rot"][1][:, 3] != MISS).mean() > 0.5
