"""The PRODUCTION traversal kernel source (raytracer3_b200/csrc/trace_kernel.cuh, what nvcc compiles into librt3.so) run on the CPU over a
lock-step warp emulation (tests/emu/warp_emu.cpp: 32 fibers per warp, ballots / shuffles / __syncwarp as rendezvous points) and compared
bit for bit with a brute-force loop: dynamic ray fetch, two-phase loop, stack protocol, queue length by value or through memory.
The emulation also reports a lane that leaves the kernel while its warp waits at a collective, or lanes meeting at different
collectives — the hangs / undefined behaviour of divergent *_sync code on hardware."""
import numpy as np
import pytest

from emu_lib import EmuBvh
from raytracer3_b200 import scenes
from raytracer3_b200.types import MISS, RAY

CASES = {
    "cornell": lambda: scenes.cornell_box(),
    "soup": lambda: scenes.random_soup(3000, seed=4),
    "atrium": lambda: scenes.synthetic_atrium(0.03, environment=False),
}
# compile-time experiments of trace_kernel.cuh (none is in the product build): they must not change a single hit either
VARIANTS = [((), False), (("RT3_TRI_REPS=1",), False), (("RT3_TRI_WAIT=8",), False), ((), True), (("RT3_TRI_WAIT=16",), True), (("RT3_SMEM_STACK=4",), False),
            (("RT3_NODE_COOP=1",), False)]


def _scene(name, n=1500):
    sc = CASES[name]()
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    c, a = scenes.incoherent_rays(sc, n, seed=51, tris=T.astype(np.float64))
    return EmuBvh(tris=T), np.concatenate([c, scenes.random_rays(sc, n, seed=52)]), a


@pytest.mark.parametrize("name", list(CASES))
def test_production_kernel_equals_brute_force(name):
    b, rays, shadow = _scene(name)
    want, _ = b.trace_closest(rays, brute=True)
    got, collectives = b.kernel_trace(rays, blocks=2)
    assert (got == want).all() and collectives > 0
    assert (b.kernel_trace(rays, d_count=True)[0] == want).all()                       # queue length read from memory, as between bounces
    assert (b.kernel_trace(shadow, any_hit=True)[0] == b.trace_any(shadow, brute=True)).all()
    ph = b.phase_stats
    assert ph[0][1] / ph[0][0] > 16                                                     # dynamic fetch keeps the node phase more than half full


@pytest.mark.parametrize("defines,balance", VARIANTS)
def test_kernel_experiments_do_not_change_hits(defines, balance):
    b, rays, shadow = _scene("atrium", 1200)
    want, _ = b.trace_closest(rays, brute=True)
    got, _ = b.kernel_trace(rays, defines=defines, balance=balance)
    assert (got == want).all()
    got_any, _ = b.kernel_trace(shadow, any_hit=True, d_count=True, defines=defines, balance=balance)
    assert (got_any == b.trace_any(shadow, brute=True)).all()


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 127, 129])
def test_ragged_queue_lengths(n):
    b, rays, _ = _scene("cornell", 100)
    r = rays[:n]
    want, _ = b.trace_closest(r, brute=True) if n else (np.zeros(0, want_dtype()), None)
    for balance in (False, True):
        got, _ = b.kernel_trace(r, blocks=3, balance=balance)                           # more warps than work: most find the queue empty
        assert (got == want).all()


def want_dtype():
    from raytracer3_b200.types import HIT
    return HIT


def test_degenerate_scene_and_rays():
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32)
    rays = np.zeros(40, RAY)
    rays["o"] = [0.2, 0.2, -1]
    rays["d"] = [0, 0, 1]
    rays["tmax"] = 10
    rays["tmax"][::3] = 0.5                                                             # too short to reach the triangle
    for T in (np.repeat(tri, 37, 0), tri, np.zeros((0, 3, 3), np.float32)):
        b = EmuBvh(tris=T)
        got, _ = b.kernel_trace(rays, balance=True)
        want, _ = b.trace_closest(rays, brute=True)
        assert (got == want).all()
        if len(T):
            assert (got["prim"][1::3] == 0).all() and (got["prim"][::3] == MISS).all()


def test_kernel_on_the_frames_own_bounce_rays():
    """The oracle can hand out the secondary rays of a frame (orc_set_ray_capture); the production kernel must answer them like brute force,
    and there must be exactly as many as the oracle counts."""
    from oracle.oracle_py import OracleScene
    from raytracer3_b200.camera import camera_for, make_params
    from raytracer3_b200.types import ptr
    sc = scenes.synthetic_atrium(0.02)
    sc["bluenoise"] = scenes.blue_noise_tile(64)
    W, H = 64, 36
    orc = OracleScene(sc)
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=1, bounces=4, samples=1, flags=3)
    gb, dp, _, _ = orc.pass_gbuffer(p)
    rays, valid = np.zeros((3, W * H), RAY), np.zeros((3, W * H), np.uint8)
    orc.L.orc_set_ray_capture(ptr(rays), ptr(valid), W * H, 3)
    try:
        _, cnt = orc.pass_reference(p, gb, dp)
    finally:
        orc.L.orc_set_ray_capture(None, None, 0, 0)
    assert int(valid.sum()) == cnt.rays_secondary and (valid.sum(1)[:-1] >= valid.sum(1)[1:]).all()      # queues only shrink from bounce to bounce
    queue = np.ascontiguousarray(rays[valid != 0])
    b = EmuBvh(tris=np.ascontiguousarray(scenes.world_triangles(sc), np.float32))
    want, _ = b.trace_closest(queue, brute=True)
    for defines in ((), ("RT3_TRI_WAIT=8",)):
        got, _ = b.kernel_trace(queue, defines=defines)
        assert (got == want).all()
