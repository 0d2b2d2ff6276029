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


def _check(sc, W, H, flags, frames=1, samples=1, bounces=4, postprocess=False, strips=1, round_lengths=None, **build):
    orc = OracleScene(sc)
    emu = PassEmu(orc, **build)
    prev_o = prev_e = None
    for frame in range(frames):
        p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=frame, bounces=bounces, samples=samples, blendfactor=1.0 if frame == 0 else 0.5, flags=flags)
        gb, dp, ids, bary = orc.pass_gbuffer(p)
        want, cnt = orc.pass_reference(p, gb, dp, prev_light=prev_o)
        got = emu.frame(p, prev_light=prev_e, postprocess=postprocess, strips=strips, round_lengths=round_lengths)
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


def test_regeneration_keeps_the_queue_full_and_the_image_identical():
    """Ray regeneration between bounces (north_star; the loop being turned inside out is refrence_mode.slang:28-57): with samples > 1 a path
    that ends is replaced, in the same shade thread, by the first bounce of its pixel's next sample.  The image stays bit-identical to the
    per-pixel loop of the oracle (_check), and the queue of round r holds every unfinished pixel instead of shrinking bounce by bounce:
    its length stays above 90 % of the first round's for the first (samples - 1) * 2 rounds."""
    sc = scenes.synthetic_atrium(0.02)
    sc["bluenoise"] = scenes.blue_noise_tile(64)
    samples, bounces = 6, 4
    lengths = np.zeros(4096, np.uint64)
    cnt = _check(sc, 64, 36, types.FLAG_ENV | types.FLAG_BLUENOISE, samples=samples, bounces=bounces, round_lengths=lengths)
    rounds = samples * (bounces - 1)
    assert int(lengths[:rounds].sum()) == cnt.rays_secondary and lengths[rounds:].sum() == 0
    assert lengths[0] > 0 and (lengths[:(samples - 1) * 2] > 0.9 * lengths[0]).all(), lengths[:rounds]
    assert lengths[rounds - 1] < 0.5 * lengths[0]                                               # the tail: only the longest pixel chains are left


@pytest.mark.parametrize("samples,bounces,flag_names", [(5, 3, ("ENV",)), (3, 1, ()), (2, 0, ()), (4, 2, ("ENV", "NEE")), (3, 4, ("ENV", "SPECULAR", "NEE"))])
def test_regeneration_sample_and_bounce_counts(samples, bounces, flag_names):
    flags = 0
    for n in flag_names:
        flags |= getattr(types, "FLAG_" + n)
    sc = scenes.metal_cornell_box() if "SPECULAR" in flag_names else scenes.cornell_box()
    if flags & types.FLAG_ENV:
        sc["environment"] = scenes.analytic_sky(64, 32)
    _check(sc, 40, 24, flags, samples=samples, bounces=bounces, frames=2)


@pytest.mark.parametrize("strips", [2, 3])
def test_strips_do_not_change_the_image(strips):
    """rt3_launch_reference runs the frame as `strips` concurrent wavefronts (units of 2048 work indices dealt round-robin): same image."""
    sc = scenes.synthetic_atrium(0.02)
    sc["bluenoise"] = scenes.blue_noise_tile(64)
    _check(sc, 200, 120, types.FLAG_ENV | types.FLAG_BLUENOISE | types.FLAG_NEE, samples=2, bounces=3, strips=strips)


def test_work_index_division_by_multiply_high_is_exact():
    """pass_types.h fastdiv: the work-index -> pixel map of every pass kernel divides by launch constants with a multiply-high"""
    from emu_lib import pass_lib
    P = pass_lib()
    P.pemu_check_fastdiv.restype = __import__("ctypes").c_uint32
    assert P.pemu_check_fastdiv(2000000) == 0
