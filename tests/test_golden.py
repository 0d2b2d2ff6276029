"""Frozen golden vectors (tests/golden/*.npz, produced by tests/golden/make_golden.py from the CPU oracle)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


@pytest.mark.parametrize("name", ["cornell", "atrium", "metal"])
def test_oracle_reproduces_golden(name):
    g = np.load(os.path.join(HERE, "golden", f"{name}.npz"))
    sc0 = make_golden.case(name)[0]
    _, _, out = make_golden.render(name, make_golden.scene_from_arrays(g, sc0["cameras"]), g["params"])
    for k in g.files:
        assert (_bits(out[k]) == _bits(g[k])).all(), k


def test_kat_tables():
    g = np.load(os.path.join(HERE, "golden", "kat.npz"))
    now = make_golden.kat_tables()
    for k in g.files:
        assert (_bits(now[k]) == _bits(g[k])).all(), k
    assert g["murmur3"][0] == 0x2362F9DE      # MurmurHash3_x86_32(4 zero bytes, seed 0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell", "atrium", "metal"])
def test_gpu_reproduces_golden(gpu_ctx, name):
    """The CUDA path against the committed fixtures (no oracle involved at run time)."""
    from raytracer3_b200.camera import camera_for, make_params
    from raytracer3_b200.host import Renderer, RenderWorld
    g = np.load(os.path.join(HERE, "golden", f"{name}.npz"))
    sc0, flags, (W, H), kw = make_golden.case(name)
    sc = make_golden.scene_from_arrays(g, sc0["cameras"])
    w = RenderWorld(gpu_ctx)
    w.clear()
    w.load_scene(sc)
    w.create_acceleration_structure()
    p = np.ascontiguousarray(g["params"])
    r = Renderer(gpu_ctx, w, W, H)
    light = r.frame(p, debug_ids=True).cpu().numpy()
    assert (_bits(r.rg.tensor("ids").cpu().numpy()) == _bits(g["ids"])).all()
    assert (_bits(r.rg.tensor("gbuffer_depth").cpu().numpy()) == _bits(g["depth"])).all()
    assert (_bits(r.rg.tensor("bary").cpu().numpy()) == _bits(g["bary"])).all()
    assert (_bits(r.rg.tensor("gbuffer").cpu().numpy()) == _bits(g["gbuffer"])).all()
    assert (_bits(light) == _bits(g["light"])).all()
    hits = w.trace_closest_host(g["rays"])
    assert (_bits(hits) == _bits(g["hits"])).all()
