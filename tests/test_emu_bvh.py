"""CPU emulation of the DEVICE builder / traversal logic (raytracer3_b200/csrc/bvh_logic.h compiled with g++, see tests/emu/emu.cpp):
Karras hierarchy + refit + 8-wide quantised collapse + octant-ordered traversal must agree with a brute-force loop over the
same triangle records, bit for bit, and must keep every triangle exactly once."""
import numpy as np
import pytest

from emu_lib import EmuBvh
from raytracer3_b200 import scenes
from raytracer3_b200.types import MISS, RAY

CASES = {
    "cornell": lambda: scenes.cornell_box(),
    "soup": lambda: scenes.random_soup(3000, seed=4),
    "atrium": lambda: scenes.synthetic_atrium(0.03, environment=False),
    "grid": lambda: scenes.instanced_grid(scenes.synthetic_atrium(0.004, environment=False), 3),
}


@pytest.mark.parametrize("optimal", [True, False])
@pytest.mark.parametrize("name", list(CASES))
def test_wide_bvh_equals_brute_force(name, optimal):
    sc = CASES[name]()
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    b = EmuBvh(tris=T, optimal_collapse=optimal)
    info = b.info()
    assert info["tris"] == len(T) and b.check_triangles() == 0          # order-independent: every triangle kept exactly once, unmodified
    assert b.check_node_format() == 0                                   # valid24 / inner mask of every node, permutation table, record indexing
    assert info["levels"] < 64
    c, a = scenes.incoherent_rays(sc, 3000, seed=31, tris=T.astype(np.float64))
    rays = np.concatenate([c, scenes.random_rays(sc, 3000, seed=32)])
    got, st = b.trace_closest(rays)
    want, _ = b.trace_closest(rays, brute=True)
    assert (got == want).all()
    assert 1.0 <= st[0] < 60 and st[1] < 40                               # nodes / triangles tested per ray stay bounded
    assert (b.trace_any(a) == b.trace_any(a, brute=True)).all()


def test_cost_optimal_collapse_needs_fewer_nodes_and_less_work():
    """The dynamic programme (bvh_logic.h dp_node / dp_select_children) against the greedy choice on the same binary tree."""
    sc = scenes.synthetic_atrium(0.05, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, _ = scenes.incoherent_rays(sc, 4000, seed=5, tris=T.astype(np.float64))
    greedy, best = EmuBvh(tris=T, optimal_collapse=False), EmuBvh(tris=T, optimal_collapse=True)
    hg, sg = greedy.trace_closest(c)
    hb, sb = best.trace_closest(c)
    assert (hg == hb).all()                                                # hits do not depend on the topology
    assert best.info()["nodes"] < 0.85 * greedy.info()["nodes"]
    assert sb[0] + 0.4 * sb[1] < sg[0] + 0.4 * sg[1]                       # node tests + 0.4 x triangle tests per ray (measured cost ratio)


def test_degenerate_inputs():
    # all triangles identical (duplicate Morton codes -> index tie-break in lbvh_delta), zero-area triangles, one triangle, none
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32)
    rays = np.zeros(64, RAY)
    rays["o"] = [0.2, 0.2, -1]
    rays["d"] = [0, 0, 1]
    rays["tmax"] = 10
    for T, optimal in [(t, o) for t in (np.repeat(tri, 37, 0), tri, np.concatenate([tri, np.zeros((5, 3, 3), np.float32)])) for o in (True, False)]:
        b = EmuBvh(tris=T, optimal_collapse=optimal)
        assert b.check_triangles() == 0
        got, _ = b.trace_closest(rays)
        assert (got["prim"] == 0).all() and np.allclose(got["t"], 1.0)     # equal t resolves to the lowest triangle id
    b = EmuBvh(tris=np.zeros((0, 3, 3), np.float32))
    got, _ = b.trace_closest(rays)
    assert (got["prim"] == MISS).all()


def test_axis_parallel_rays_and_flat_boxes():
    # floor quad at y = 0 (flat AABB) hit by rays with exact-zero direction components
    sc = scenes.cornell_box()
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    b = EmuBvh(tris=T)
    rays = np.zeros(200, RAY)
    rng = np.random.default_rng(0)
    rays["o"] = np.stack([rng.uniform(-0.9, 0.9, 200), np.full(200, 0.9), rng.uniform(-0.9, 0.9, 200)], 1)
    rays["d"] = [0, -1, 0]
    rays["tmax"] = 100
    got, _ = b.trace_closest(rays)
    want, _ = b.trace_closest(rays, brute=True)
    assert (got == want).all() and (got["prim"] != MISS).all()


@pytest.mark.parametrize("name", ["cornell", "soup", "atrium"])
def test_alternative_topologies_find_the_same_hits(name):
    """Other topologies: treelet restructuring (bvh_logic.h, RT3_BUILD_TREELET, opt-in on the device), and two experiments (tests/emu/bvh_experiments.h) that are not
    in the device build: PLOC instead of the Karras hierarchy, triangle splitting (several references with clipped boxes
    per triangle).  Topology must never change a hit."""
    sc = CASES[name]()
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, a = scenes.incoherent_rays(sc, 2000, seed=41, tris=T.astype(np.float64))
    rays = np.concatenate([c, scenes.random_rays(sc, 2000, seed=42)])
    base = EmuBvh(tris=T)
    want, _ = base.trace_closest(rays, brute=True)
    want_any = base.trace_any(a, brute=True)
    for kw in (dict(treelet_passes=1), dict(treelet_passes=2), dict(ploc_radius=4), dict(ploc_radius=25), dict(split_beta=0.5), dict(split_beta=4.0)):
        b = EmuBvh(tris=T, **kw)
        info = b.info()
        assert b.check_triangles() == 0 and info["levels"] < 64
        if "treelet_passes" in kw:
            assert info["tris"] == len(T)
        elif "ploc_radius" in kw:
            assert info["tris"] == len(T) and 1 <= b.ploc_rounds() < 200
        else:
            assert len(T) <= info["tris"] <= len(T) * (1 + kw["split_beta"]) + 1        # the split budget bounds the extra references
        got, st = b.trace_closest(rays)
        assert (got == want).all(), kw
        assert (b.trace_any(a) == want_any).all(), kw
        assert st[0] < 60 and st[1] < 40
    assert EmuBvh(tris=T, split_beta=4.0).info()["tris"] > len(T)                      # and it does split something


def test_ploc_and_split_degenerate_inputs():
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32)
    rays = np.zeros(8, RAY)
    rays["o"] = [0.2, 0.2, -1]
    rays["d"] = [0, 0, 1]
    rays["tmax"] = 10
    for T in (np.repeat(tri, 37, 0), tri, np.concatenate([tri, np.zeros((5, 3, 3), np.float32)]),
              np.concatenate([tri, tri * 1000.0 + 5.0])):                              # identical, single, zero-area, one huge + one small
        for kw in (dict(treelet_passes=2), dict(ploc_radius=3), dict(split_beta=8.0)):
            b = EmuBvh(tris=T, **kw)
            assert b.check_triangles() == 0
            got, _ = b.trace_closest(rays)
            want, _ = b.trace_closest(rays, brute=True)
            assert (got == want).all() and (got["prim"] == 0).all()


def test_treelet_restructuring_beats_rotations():
    """RT3_BUILD_TREELET (treelet_visit) against the default rotations (refit_visit) on the same Karras hierarchy: same hits, less work."""
    sc = scenes.synthetic_atrium(0.05, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, _ = scenes.incoherent_rays(sc, 4000, seed=5, tris=T.astype(np.float64))
    rot, tre = EmuBvh(tris=T), EmuBvh(tris=T, treelet_passes=2)
    hr, sr = rot.trace_closest(c)
    ht, stt = tre.trace_closest(c)
    assert (hr == ht).all()
    assert stt[0] + 0.4 * stt[1] < 0.99 * (sr[0] + 0.4 * sr[1])
    assert tre.info()["sah_nodes"] < rot.info()["sah_nodes"]
    tiny = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[2, 0, 0], [3, 0, 0], [2, 1, 0]], [[0, 2, 0], [1, 2, 0], [0, 3, 0]]], np.float32)
    for k in (1, 2, 3):                                                    # treelets smaller than 7 leaves
        assert EmuBvh(tris=tiny[:k], treelet_passes=2).check_triangles() == 0
