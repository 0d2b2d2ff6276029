"""The C++ host side above the C ABI (include/rt3_host.hpp — the reference's RayTracingPass / ComputePass / RenderGraph / RenderWorld API
for a host that links librt3.so directly): tests/cpp/host_demo.cpp renders frames through it, with no Python or torch in the process, and
the result is compared with the CPU oracle bit for bit."""
import os
import struct
import subprocess

import numpy as np
import pytest

from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_ENV, GEOMETRY, PARAMS, VERTEX


def _demo():
    import __graft_entry__ as g
    g.build()
    return g.build_cpp_host_demo()


def _blob(path, sc, W, H, params):
    env = sc.get("environment")
    eh, ew = env.shape[:2] if env is not None else (0, 0)
    with open(path, "wb") as f:
        f.write(b"RT3S" + struct.pack("<8I", W, H, len(sc["vertices"]), len(sc["indices"]), len(sc["geometries"]), len(params), ew, eh))
        f.write(np.ascontiguousarray(sc["vertices"], VERTEX).tobytes())
        f.write(np.ascontiguousarray(sc["indices"], np.uint32).tobytes())
        f.write(np.ascontiguousarray(sc["geometries"], GEOMETRY).tobytes())
        if env is not None:
            f.write(np.ascontiguousarray(env, np.float32).tobytes())
        for p in params:
            f.write(np.ascontiguousarray(p, PARAMS).tobytes())


def _case():
    sc = scenes.cornell_box()
    sc["environment"] = scenes.analytic_sky(64, 32)
    W, H = 96, 64
    cam = camera_for(sc, "cam0", W, H)
    params = [make_params(cam, W, H, frame=n, bounces=4, samples=2, blendfactor=1.0 / (n + 1), flags=FLAG_ENV) for n in range(3)]
    return sc, W, H, params


def test_cpp_host_builds_and_refuses_to_run_without_a_device(tmp_path):
    """CPU suite: the header compiles as plain C++17 against rt3.h + the CUDA runtime, the program links against librt3.so, and without a
    GPU it stops at rt3_create with RT3_ERR_NO_DEVICE (exit code 3): there is no CPU fallback to fall into."""
    exe = _demo()
    sc, W, H, params = _case()
    blob, out = str(tmp_path / "scene.blob"), str(tmp_path / "out.bin")
    _blob(blob, sc, W, H, params)
    import torch
    r = subprocess.run([exe, blob, out], capture_output=True, text=True, timeout=300)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
    else:
        assert r.returncode == 3 and "rt3 error -2" in r.stderr, (r.returncode, r.stderr)
        assert not os.path.exists(out) or os.path.getsize(out) == 0
    bad = subprocess.run([exe, __file__, out], capture_output=True, text=True, timeout=60)
    assert bad.returncode == 1 and "not a scene blob" in bad.stderr


@pytest.mark.gpu
def test_cpp_host_frames_match_the_oracle(gpu_ctx, tmp_path):
    from oracle.oracle_py import OracleScene
    exe = _demo()
    sc, W, H, params = _case()
    blob, out = str(tmp_path / "scene.blob"), str(tmp_path / "out.bin")
    _blob(blob, sc, W, H, params)
    r = subprocess.run([exe, blob, out], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    raw = open(out, "rb").read()
    px = W * H
    assert len(raw) == len(params) * px * 16 + px * 16 + px * 4 + px * 16 + 24
    orc = OracleScene(sc)
    gb, dp, _, _ = orc.pass_gbuffer(params[0])
    prev, o, secondary, started = None, 0, 0, 0
    for n, p in enumerate(params):
        want, cnt = orc.pass_reference(p, gb, dp, prev_light=prev)
        prev = want
        secondary += cnt.rays_secondary
        started += cnt.paths_started
        got = np.frombuffer(raw, np.uint32, px * 4, o).reshape(H, W, 4)
        o += px * 16
        assert (got == want.view(np.uint32)).all(), f"Light of frame {n} differs from the oracle"
    got_gb = np.frombuffer(raw, np.uint32, px * 4, o).reshape(H, W, 4); o += px * 16
    got_dp = np.frombuffer(raw, np.uint32, px, o).reshape(H, W); o += px * 4
    got_post = np.frombuffer(raw, np.uint32, px * 4, o).reshape(H, W, 4); o += px * 16
    hit = dp != np.float32(100000.0)
    assert (got_dp == dp.view(np.uint32)).all() and (got_gb[hit] == gb[hit]).all()
    want_post = orc.pass_postprocess(params[-1], dp, prev)
    assert (got_post == np.ascontiguousarray(want_post).view(np.uint32).reshape(H, W, 4)).all()
    rays_primary, rays_secondary, paths_started = struct.unpack_from("<3Q", raw, o)
    assert rays_primary == len(params) * px and rays_secondary == secondary and paths_started == started
