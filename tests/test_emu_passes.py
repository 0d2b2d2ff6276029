"""Whole frames of the PRODUCT's kernels on the CPU: csrc/pass_kernels.cuh (k_gbuffer, k_ref_init / k_ref_shade / k_ref_shadow / k_ref_finish,
k_postprocess) and csrc/trace_kernel.cuh compiled with g++ over the lock-step warp emulation (tests/emu/warp_emu.h, pass_emu.cpp), launched
in the order of rt3_launch_gbuffer / rt3_launch_reference, and compared bit for bit with the oracle — ids, depth, barycentrics, packed
G-buffer words, radiance, counters.  The same comparison runs on the B200 in tests/test_gpu_parity.py; this one needs no GPU, so a change
to a pass kernel can be checked before it costs GPU minutes."""
import numpy as np
import pytest

from emu_lib import PassEmu
from oracle.oracle_py import OracleScene
from raytracer3_b200 import scenes, types
from raytracer3_b200.camera import camera_for, make_params


def _check(sc, W, H, flags, frames=1, samples=1, bounces=4, postprocess=False, **build):
    orc = OracleScene(sc)
    emu = PassEmu(orc, **build)
    prev_o = prev_e = None
    for frame in range(frames):
        p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=frame, bounces=bounces, samples=samples, blendfactor=1.0 if frame == 0 else 0.5, flags=flags)
        gb, dp, ids, bary = orc.pass_gbuffer(p)
        want, cnt = orc.pass_reference(p, gb, dp, prev_light=prev_o)
        got = emu.frame(p, prev_light=prev_e, postprocess=postprocess)
        assert (got["ids"] == ids).all()                                                      # primary-hit ids: bit-exact
        assert (got["depth"].view(np.uint32) == dp.view(np.uint32)).all()
        assert (got["bary"].view(np.uint32) == bary.view(np.uint32)).all()
        assert (got["gbuffer"] == gb).all()
        assert (got["light"].view(np.uint32) == want.view(np.uint32)).all()                   # radiance: bit-exact
        c = got["counters"]
        assert (int(c[0]), int(c[1]), int(c[2]), int(c[3]), int(c[4])) == (W * H, cnt.rays_secondary, cnt.rays_shadow, cnt.paths_started, cnt.paths_escaped)
        if postprocess:
            assert (got["post"].view(np.uint32) == orc.pass_postprocess(p, dp, want).view(np.uint32)).all()
        prev_o, prev_e = want, got["light"]
    return cnt


@pytest.mark.parametrize("flag_names", [(), ("ENV", "BLUENOISE"), ("ENV", "SPECULAR"), ("ENV", "BLUENOISE", "NEE"), ("ENV", "SPECULAR", "NEE")])
def test_frames_of_the_product_kernels_match_the_oracle(flag_names):
    flags = 0
    for n in flag_names:
        flags |= getattr(types, "FLAG_" + n)
    sc = scenes.synthetic_atrium(0.02)
    sc["bluenoise"] = scenes.blue_noise_tile(64)
    cnt = _check(sc, 72, 40, flags, frames=2, samples=2, postprocess=True)
    assert cnt.rays_secondary > 0 and (cnt.rays_shadow > 0) == ("NEE" in flag_names)


def test_cornell_box_ragged_size_and_bounce_counts():
    sc = scenes.cornell_box()
    for W, H, bounces in ((37, 23, 4), (8, 4, 1), (9, 5, 2), (33, 1, 0)):                      # not multiples of the 8x4 pixel blocks; no / one secondary ray
        _check(sc, W, H, 0, bounces=bounces)


def test_treelet_build_under_the_product_kernels():
    sc = scenes.synthetic_atrium(0.02)
    sc["bluenoise"] = scenes.blue_noise_tile(64)
    _check(sc, 48, 27, types.FLAG_ENV | types.FLAG_BLUENOISE, treelet_passes=2)
