"""Generates tests/golden/*.npz from the CPU oracle (run once; the files are committed).

The reference has no golden vectors for this path and cannot run here, so these freeze the oracle's own outputs on small seeded
cases: a drift detector for oracle/, include/rt3_shared_math.h and the scene generators, and a second target for the GPU tests."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle_py import OracleScene, lib  # noqa: E402
from raytracer3_b200 import scenes  # noqa: E402
from raytracer3_b200.camera import camera_for, make_params  # noqa: E402
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV, FLAG_NEE, FLAG_SPECULAR  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(name):
    if name == "cornell":
        return scenes.cornell_box(), 0, (64, 40), dict(frame=3, bounces=4, samples=2)
    if name == "metal":      # layered diffuse + specular material with environment NEE (RT3_FLAG_SPECULAR | NEE | ENV)
        sc = scenes.metal_cornell_box()
        sc["environment"] = scenes.analytic_sky(64, 32)
        return sc, FLAG_ENV | FLAG_NEE | FLAG_SPECULAR, (64, 40), dict(frame=5, bounces=4, samples=2)
    sc = scenes.synthetic_atrium(0.01)
    sc["environment"] = scenes.analytic_sky(64, 32)
    sc["bluenoise"] = scenes.blue_noise_tile(32, seed=7, iters=3)
    return sc, FLAG_ENV | FLAG_BLUENOISE, (80, 48), dict(frame=1, bounces=4, samples=1)


SCENE_KEYS = ("vertices", "indices", "geometries", "environment", "bluenoise")


def scene_arrays(sc):
    """The scene itself is part of the fixture: numpy's SIMD sin/cos/FFT may differ in the last ulp between hosts."""
    out = {f"scene_{k}": sc[k] for k in SCENE_KEYS if sc.get(k) is not None}
    for slot, img in (sc.get("textures") or {}).items():
        out[f"scene_texture_{slot}"] = img
    return out


def scene_from_arrays(g, cameras):
    sc = {"name": "golden", "instances": [], "textures": {}, "environment": None, "cameras": cameras}
    for k in g.files:
        if k.startswith("scene_texture_"):
            sc["textures"][int(k.rsplit("_", 1)[1])] = g[k]
        elif k.startswith("scene_"):
            sc[k[6:]] = g[k]
    return sc


def render(name, sc=None, params=None):
    sc0, flags, (W, H), kw = case(name)
    sc = sc0 if sc is None else sc
    orc = OracleScene(sc, threads=1)
    p = make_params(camera_for(sc, "cam0", W, H), W, H, flags=flags, **kw) if params is None else np.ascontiguousarray(params)
    gb, dp, ids, bary = orc.pass_gbuffer(p)
    light, cnt = orc.pass_reference(p, gb, dp)
    rays = scenes.random_rays(sc, 256, seed=99)
    hits = orc.trace_closest(rays)
    return sc, p, dict(**scene_arrays(sc), params=p, gbuffer=gb, depth=dp, ids=ids, bary=bary, light=light, rays=rays, hits=hits,
                       counters=np.array([cnt.rays_secondary, cnt.paths_started, cnt.paths_escaped], np.uint64))


def kat_tables():
    L = lib()
    seeds = [(0, 0), (1, 0), (0, 1), (0xDEADBEEF, 12345), (0xFFFFFFFF, 0xFFFFFFFF)]
    return dict(
        murmur3=np.array([L.orc_murmur3(s, i) for s, i in seeds], np.uint32),
        jenkins=np.array([L.orc_hash_jenkins(a) for a in (0, 1, 2, 0xFFFFFFFF, 0x12345678)], np.uint32),
        rng_seed=np.array([L.orc_rng_seed(x, y, f) for x, y, f in ((0, 0, 0), (5, 9, 7), (1919, 1079, 63))], np.uint32),
        uniform=np.array([L.orc_uniform_float(s, i) for s, i in seeds], np.float32),
    )


if __name__ == "__main__":
    only = sys.argv[1:]          # optional: names to (re)generate; the other fixtures stay frozen
    for name in (only or ("cornell", "atrium", "metal")):
        _, _, out = render(name)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    if not only:
        np.savez_compressed(os.path.join(HERE, "kat.npz"), **kat_tables())
    print("golden written")
