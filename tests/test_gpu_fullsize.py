"""Parity at BASELINE.json's own configurations and sizes (not the reduced scenes of test_gpu_parity.py): the CUDA path through the C ABI
against the CPU oracle on the FULL bench scene — synthetic_atrium(SPONZA_DETAIL), 262,924 triangles, the stand-in for sponza_scene.glb,
which is absent from the reference mount — with the reference's own blue-noise tile (resources/bluenoise.png).

  configs[0]  1280x720 primary visibility: hit ids, depth, barycentrics, packed G-buffer words          (gbuffer.slang:8-20)
  configs[1]  1920x1080, 4 bounces, 1 spp, environment light + blue-noise sampling, two frames with the temporal blend
                                                                                                       (refrence_mode.slang:14-66)
  configs[3]  a 2^20-ray slice of the incoherent-ray generator, closest-hit and any-hit
  configs[4]  the instanced grid (2x2 here; the oracle flattens) in both build modes at 1080p
  configs[2]  see test_multi_sample_frame_* (3840x2160 is covered by the size-independent tile property in test_gpu_edges.py)

Bars: ids bit-exact, t/u/v <= 1e-5 relative, radiance RMSE <= 1e-3 (BASELINE.json north_star); asserted together with the stronger
property the determinism contract gives (bit-identical words)."""
import numpy as np
import pytest

from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import BUILD_TWO_LEVEL, FLAG_BLUENOISE, FLAG_ENV, MISS

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def full(gpu_ctx):
    """The bench scene, uploaded and built once, plus its oracle."""
    from oracle.oracle_py import OracleScene
    from raytracer3_b200.host import RenderWorld
    sc = scenes.synthetic_atrium(scenes.SPONZA_DETAIL)
    sc["bluenoise"], src = scenes.reference_blue_noise()
    assert src.startswith("resources/bluenoise.png"), src
    orc = OracleScene(sc)
    assert orc.num_triangles == 262924

    class World:
        """(Re)loads the scene into the session's context when another test has replaced it."""
        def get(self):
            if getattr(gpu_ctx, "_scene_token", None) is not sc:
                self.w = RenderWorld(gpu_ctx)
                self.w.clear()
                self.w.load_scene(sc)
                gpu_ctx._scene_token = sc
            bs = self.w.create_acceleration_structure()
            assert bs.triangles == orc.num_triangles
            return self.w
    return sc, World(), orc


def test_cfg1_primary_visibility_720p_full_scene(gpu_ctx, full):
    from raytracer3_b200.host import Renderer
    sc, world, orc = full
    w = world.get()
    W, H = 1280, 720
    p = make_params(camera_for(sc, "cam0", W, H), W, H, bounces=1, samples=1)
    r = Renderer(gpu_ctx, w, W, H)
    r.frame(p, debug_ids=True)
    ids = r.rg.tensor("ids").cpu().numpy().view(np.uint32)
    bary, depth = r.rg.tensor("bary").cpu().numpy(), r.rg.tensor("gbuffer_depth").cpu().numpy()
    gbuf = r.rg.tensor("gbuffer").cpu().numpy().view(np.uint32)
    gb, dp, oids, obary = orc.pass_gbuffer(p)
    assert (ids == oids).all(), f"{(ids != oids).any(-1).sum()} pixels differ in (prim, geometry)"
    hit = oids[..., 0] != MISS
    assert hit.mean() > 0.9
    assert (np.abs(depth[hit] - dp[hit]) / dp[hit]).max() <= 1e-5 and np.abs(bary - obary).max() <= 1e-5
    assert (_bits(depth) == _bits(dp)).all() and (_bits(bary) == _bits(obary)).all() and (gbuf == gb).all()


def test_cfg2_path_tracing_1080p_full_scene_two_frames(gpu_ctx, full):
    """The bench headline's workload, frame for frame: ENV | BLUENOISE, 4 bounces, 1 spp; the second frame blends with the first."""
    from raytracer3_b200.host import Renderer
    sc, world, orc = full
    w = world.get()
    W, H = 1920, 1080
    r = Renderer(gpu_ctx, w, W, H)
    gpu_ctx.counters(reset=True)
    prev, rays = None, 0
    for frame in range(2):
        p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=frame, bounces=4, samples=1, blendfactor=1.0 if frame == 0 else 0.5,
                        flags=FLAG_ENV | FLAG_BLUENOISE)
        light = r.frame(p).cpu().numpy()
        gb, dp, _, _ = orc.pass_gbuffer(p)
        want, cnt = orc.pass_reference(p, gb, dp, prev_light=prev)
        prev = want
        rays += cnt.rays_secondary
        rmse = float(np.sqrt(((light - want)[..., :3] ** 2).mean()))
        assert rmse <= 1e-3, rmse
        assert (_bits(light) == _bits(want)).all(), f"frame {frame}: radiance not bit-identical (rmse {rmse})"
    c = gpu_ctx.counters(reset=True)
    assert c.rays_primary == 2 * W * H and c.rays_secondary == rays and rays > 10_000_000


def test_cfg2_multi_sample_frame_full_scene(gpu_ctx, full):
    """configs[2]'s inner loop (samples > 1 inside one frame call) at 1080p on the full scene: 4 spp, 4 bounces."""
    from raytracer3_b200.host import Renderer
    sc, world, orc = full
    w = world.get()
    W, H = 1920, 1080
    p = make_params(camera_for(sc, "cam1", W, H), W, H, frame=5, bounces=4, samples=4, flags=FLAG_ENV | FLAG_BLUENOISE)
    r = Renderer(gpu_ctx, w, W, H)
    light = r.frame(p).cpu().numpy()
    gb, dp, _, _ = orc.pass_gbuffer(p)
    want, _ = orc.pass_reference(p, gb, dp)
    assert (_bits(light) == _bits(want)).all()


def test_cfg4_incoherent_slice_full_bvh(gpu_ctx, full):
    sc, world, orc = full
    w = world.get()
    closest, anyr = scenes.incoherent_rays(sc, 1 << 20)
    got, want = w.trace_closest_host(closest), orc.trace_closest(closest)
    assert (got["prim"] == want["prim"]).all(), int((got["prim"] != want["prim"]).sum())
    hit = want["prim"] != MISS
    assert hit.mean() > 0.5
    assert (np.abs(got["t"][hit] - want["t"][hit]) / want["t"][hit]).max() <= 1e-5
    for f in ("t", "u", "v"):
        assert (_bits(got[f]) == _bits(want[f])).all(), f
    gb, wb = w.trace_any_host(anyr), orc.trace_any(anyr)
    assert (gb == wb).all()
    frac = np.unpackbits(wb.view(np.uint8)).mean()
    assert 0.2 < frac < 0.99


def test_cfg5_instanced_grid_1080p_both_build_modes(gpu_ctx):
    """2 x 2 instances of the full atrium (1.05 M triangles; the 8 x 8 grid of configs[4] is the same code path, measured in bench.py
    --workload cfg5): flattened and two-level builds against the oracle (which always flattens), 1080p, 4 bounces."""
    from oracle.oracle_py import OracleScene
    from raytracer3_b200.host import Renderer, RenderWorld
    sc = scenes.instanced_grid(scenes.synthetic_atrium(scenes.SPONZA_DETAIL), 2)
    sc["bluenoise"], _ = scenes.reference_blue_noise()
    orc = OracleScene(sc)
    W, H = 1920, 1080
    p = make_params(camera_for(sc, "cam_grid", W, H), W, H, frame=1, bounces=4, samples=1, flags=FLAG_ENV | FLAG_BLUENOISE)
    gb, dp, oids, _ = orc.pass_gbuffer(p)
    want, _ = orc.pass_reference(p, gb, dp)
    assert (oids[..., 0] != MISS).mean() > 0.2
    gpu_ctx._scene_token = None
    for flags in (1, 1 | BUILD_TWO_LEVEL):
        w = RenderWorld(gpu_ctx)
        w.clear()
        w.load_scene(sc)
        bs = w.create_acceleration_structure(flags)
        if flags == 1:
            assert bs.triangles == orc.num_triangles
        r = Renderer(gpu_ctx, w, W, H)
        light = r.frame(p, debug_ids=True).cpu().numpy()
        assert (r.rg.tensor("ids").cpu().numpy().view(np.uint32) == oids).all(), flags
        assert (_bits(r.rg.tensor("gbuffer_depth").cpu().numpy()) == _bits(dp)).all(), flags
        assert (_bits(light) == _bits(want)).all(), flags
