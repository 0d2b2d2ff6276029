"""GPU parity at the edges of the domain: window sizes that are not multiples of the 8x4 pixel block / 16-pixel probe cell / 64-pixel
tile, one-pixel frames, sub-rect launches, tile lists with clipped and empty tiles, zero bounces, progressive accumulation.
Every comparison is bit-for-bit against the CPU oracle (or against the full-frame GPU result where the oracle was already compared)."""
import ctypes as C

import numpy as np
import pytest

from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV, FLAG_NEE, FLAG_SPECULAR, MISS, Rect, ptr

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _world(gpu_ctx, sc):
    from raytracer3_b200.host import RenderWorld
    w = RenderWorld(gpu_ctx)
    w.clear()
    w.load_scene(sc)
    w.create_acceleration_structure()
    return w


def _oracle(sc):
    from oracle.oracle_py import OracleScene
    return OracleScene(sc)


@pytest.fixture(scope="module")
def cornell():
    sc = scenes.cornell_box()
    sc["environment"] = scenes.analytic_sky(64, 32)
    sc["bluenoise"] = scenes.blue_noise_tile(64, iters=2)
    return sc, _oracle(sc)


@pytest.mark.parametrize("W,H", [(1, 1), (7, 3), (9, 5), (33, 17), (65, 130), (333, 187)])
def test_ragged_window_sizes(gpu_ctx, cornell, W, H):
    """Frames smaller than / not divisible by the 8x4-pixel warp block: padding lanes must neither write nor count."""
    from raytracer3_b200.host import Renderer
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    flags = FLAG_ENV | FLAG_BLUENOISE | FLAG_NEE | FLAG_SPECULAR
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=3, bounces=3, samples=2, flags=flags)
    r = Renderer(gpu_ctx, w, W, H)
    gpu_ctx.counters(reset=True)
    light = r.frame(p, debug_ids=True).cpu().numpy()
    c = gpu_ctx.counters(reset=True)
    gb, dp, ids, bary = orc.pass_gbuffer(p)
    want, cnt = orc.pass_reference(p, gb, dp)
    assert (r.rg.tensor("ids").cpu().numpy().view(np.uint32) == ids).all()
    assert (_bits(r.rg.tensor("gbuffer_depth").cpu().numpy()) == _bits(dp)).all()
    assert (_bits(r.rg.tensor("bary").cpu().numpy()) == _bits(bary)).all()
    hit = ids[..., 0] != MISS
    assert (r.rg.tensor("gbuffer").cpu().numpy().view(np.uint32)[hit] == gb[hit]).all()
    assert (_bits(light) == _bits(want)).all()
    assert c.rays_primary == W * H and c.rays_secondary == cnt.rays_secondary and c.rays_shadow == cnt.rays_shadow


def test_sub_rect_launch_touches_only_its_rect(gpu_ctx, cornell):
    """rt3_pass_gbuffer / rt3_pass_reference on a rect that starts off the block grid: inside == the full-frame result, outside untouched."""
    import torch
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    W, H = 200, 120
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=1, bounces=3, samples=1, flags=FLAG_ENV)
    pp = np.ascontiguousarray(p)
    gb_full, dp_full, _, _ = orc.pass_gbuffer(p)
    want, _ = orc.pass_reference(p, gb_full, dp_full)
    L, h, st = gpu_ctx.L, gpu_ctx.h, gpu_ctx.stream()
    for (x0, y0, rw, rh) in ((13, 7, 50, 33), (0, 0, 1, 1), (199, 119, 1, 1), (64, 0, 136, 120)):
        gb = torch.full((H, W, 4), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
        dp = torch.full((H, W), -7.0, dtype=torch.float32, device="cuda")
        lt = torch.full((H, W, 4), -3.0, dtype=torch.float32, device="cuda")
        rect = Rect(x0, y0, rw, rh)
        gpu_ctx.check(L.rt3_pass_gbuffer(h, ptr(pp), rect, gb.data_ptr(), dp.data_ptr(), None, None, st))
        gpu_ctx.check(L.rt3_pass_reference(h, ptr(pp), rect, gb.data_ptr(), dp.data_ptr(), None, lt.data_ptr(), st))
        gbn, dpn, ltn = gb.cpu().numpy().view(np.uint32), dp.cpu().numpy(), lt.cpu().numpy()
        inside = np.zeros((H, W), bool)
        inside[y0:y0 + rh, x0:x0 + rw] = True
        assert (dpn[~inside] == -7.0).all() and (gbn[~inside] == 0x5A5A5A5A).all() and (ltn[~inside] == -3.0).all()
        assert (_bits(dpn[inside]) == _bits(dp_full[inside])).all()
        hit = inside & (dp_full != np.float32(100000.0))
        assert (gbn[hit] == gb_full[hit]).all()
        assert (_bits(ltn[hit]) == _bits(want[hit])).all()
        bg = inside & ~hit
        assert (ltn[bg] == -3.0).all()                      # refrence_mode.slang:20-23: background pixels are left alone


@pytest.mark.parametrize("W,H,N,tile", [(200, 130, 3, 64), (97, 61, 2, 32), (260, 200, 5, 64)])
def test_tile_lists_with_clipped_and_empty_tiles(gpu_ctx, cornell, W, H, N, tile):
    """N logical ranks on one GPU: slabs of clipped edge tiles and padding rects, gathered in rank order and un-tiled, equal the
    full-frame render bit for bit (the multi-GPU path minus the NCCL call, which the gloo test and bench.py's verification cover)."""
    import torch
    from raytracer3_b200.host import Renderer, TiledRenderer
    from raytracer3_b200.tiles import all_tiles
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    flags = FLAG_ENV | FLAG_BLUENOISE
    cam = camera_for(sc, "cam0", W, H)
    full = Renderer(gpu_ctx, w, W, H)
    ranks = []
    for rk in range(N):
        try:
            ranks.append(TiledRenderer(gpu_ctx, w, W, H, rk, N, tile=tile))
        except AssertionError:
            pytest.skip("a rank without a full tile: not a configuration the tile renderer accepts")
    for n in range(3):                                       # progressive accumulation, slabs stay rank-local
        p = make_params(cam, W, H, frame=n, bounces=3, samples=1, blendfactor=1.0 / (n + 1), flags=flags)
        want = full.frame(p)
        slabs = [R.frame(p) for R in ranks]
    gathered = torch.cat(slabs, 0).contiguous()
    rects = np.ascontiguousarray(all_tiles(W, H, N, tile))
    frame = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    gpu_ctx.check(gpu_ctx.L.rt3_untile_rgba32f(gpu_ctx.h, ptr(rects), len(rects), tile, tile, gathered.data_ptr(), frame.data_ptr(), W, H, gpu_ctx.stream()))
    assert torch.equal(frame.view(torch.int32), want.view(torch.int32))
    assert float(frame[..., :3].abs().sum()) > 0
    # delivery in bands by peer-memory stores (rt3_scatter_tiles_rgba32f), the "peers" being buffers on this GPU: every logical rank stores
    # its tiles straight into the band of rows of the rank that presents them; each band must equal those rows of the full-frame render
    from raytracer3_b200 import tiles as T
    bands = T.band_rows(H, N, tile)
    bh = max(1, max(b - a for a, b in bands))
    band_buf = [torch.full((bh, W, 4), -1.0, dtype=torch.float32, device="cuda") for _ in range(N)]
    for rk, R in enumerate(ranks):
        real = R.rects[: R.n_real]
        dst = np.zeros(R.n_real, np.uint64)
        for i, t in enumerate(real):
            o = T.band_of_row(int(t["y0"]), H, N, tile)
            dst[i] = band_buf[o].data_ptr() + 16 * ((int(t["y0"]) - bands[o][0]) * W + int(t["x0"]))
        gpu_ctx.check(gpu_ctx.L.rt3_scatter_tiles_rgba32f(gpu_ctx.h, ptr(R.rects), ptr(dst), R.n_real, tile, tile, slabs[rk].data_ptr(), W, gpu_ctx.stream()))
    for rk, (a, b) in enumerate(bands):
        assert torch.equal(band_buf[rk][: b - a].view(torch.int32), want[a:b].view(torch.int32)), rk
    # ... and the single-process form of TiledRenderer.exchange
    one = TiledRenderer(gpu_ctx, w, W, H, 0, 1, tile=tile)
    assert one.setup_exchange() == "local"
    for n in range(3):
        p = make_params(cam, W, H, frame=n, bounces=3, samples=1, blendfactor=1.0 / (n + 1), flags=flags)
        light = one.frame(p)
    assert torch.equal(one.exchange(light, 1).view(torch.int32), want.view(torch.int32))


def test_zero_bounces_and_progressive_accumulation(gpu_ctx, cornell):
    from raytracer3_b200.host import Renderer
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    W, H = 96, 64
    cam = camera_for(sc, "cam0", W, H)
    r = Renderer(gpu_ctx, w, W, H)
    p0 = make_params(cam, W, H, frame=0, bounces=0, samples=1)           # the bounce loop does not run: radiance 0 on every hit pixel
    light = r.frame(p0).cpu().numpy()
    gb, dp, _, _ = orc.pass_gbuffer(p0)
    want, _ = orc.pass_reference(p0, gb, dp)
    assert (_bits(light) == _bits(want)).all() and (light[..., :3] == 0).all()
    prev = None
    for n in range(6):                                                     # cfg3's schedule: blendfactor 1/(n+1), Light / PrevLight ping-pong
        p = make_params(cam, W, H, frame=n, bounces=4, samples=1, blendfactor=1.0 / (n + 1), flags=FLAG_ENV)
        light = r.frame(p).cpu().numpy()
        want, _ = orc.pass_reference(p, gb, dp, prev_light=prev)
        prev = want
        assert (_bits(light) == _bits(want)).all(), f"accumulated frame {n} differs"


@pytest.mark.parametrize("W,H", [(250, 130), (17, 16), (16, 33)])
def test_probe_passes_at_ragged_sizes(gpu_ctx, cornell, W, H):
    """Probe grid with a partial last row / column of 16x16 cells."""
    from raytracer3_b200.host import Renderer
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=2, bounces=1, samples=1, flags=FLAG_ENV)
    gb, dp, _, _ = orc.pass_gbuffer(p)
    r = Renderer(gpu_ctx, w, W, H)
    light = r.probe_frame(p).cpu().numpy()
    atlas = r.rg.tensor(f"probe_atlas{1 - r.probe_cur}").cpu().numpy()
    sh = r.rg.tensor("probe_sh").cpu().numpy()
    want_atlas = orc.pass_probe_trace(p, gb, dp)
    want_sh = orc.pass_probe_sh(p, want_atlas)
    want_light = orc.pass_probe_interpolate(p, gb, dp, want_sh)
    assert (_bits(atlas) == _bits(want_atlas)).all()
    used = np.abs(want_sh).sum(1) > 0
    assert (_bits(sh[used]) == _bits(want_sh[used])).all()
    assert (_bits(light) == _bits(want_light)).all()


def test_invalid_arguments_are_status_codes(gpu_ctx, cornell):
    from raytracer3_b200 import _lib
    sc, _ = cornell
    w = _world(gpu_ctx, sc)
    W, H = 32, 16
    p = np.ascontiguousarray(make_params(camera_for(sc, "cam0", W, H), W, H, bounces=1, samples=1))
    L, h = gpu_ctx.L, gpu_ctx.h
    bad = p.copy(); bad["samples"] = 0
    assert L.rt3_pass_reference(h, ptr(bad), Rect(0, 0, W, H), None, None, None, None, None) != 0
    assert b"samples" in L.rt3_last_error(h)
    import torch
    gb = torch.zeros((H, W + 1, 4), dtype=torch.int32, device="cuda"); dp = torch.zeros((H, W + 1), dtype=torch.float32, device="cuda")
    assert L.rt3_pass_gbuffer(h, ptr(p), Rect(0, 0, W + 1, H), gb.data_ptr(), dp.data_ptr(), None, None, None) != 0      # rect outside the window
    assert b"rect" in L.rt3_last_error(h)
    with pytest.raises(_lib.Rt3Error):
        gpu_ctx.check(L.rt3_pass_gbuffer(h, ptr(p), Rect(0, 0, W, H), None, None, None, None, None))  # NULL outputs


def test_collapse_policies_give_identical_hits(gpu_ctx):
    """RT3_BUILD_FAST_TRACE (cost-optimal children, the default) vs RT3_BUILD_FAST_BUILD (greedy): same hits bit for bit, fewer nodes,
    every triangle kept exactly once in both (order-independent check on the downloaded structure)."""
    import sys, os, torch
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from emu_lib import EmuBvh
    from raytracer3_b200.host import RenderWorld
    from raytracer3_b200.types import BUILD_FAST_BUILD, BUILD_FAST_TRACE, BUILD_TWO_LEVEL
    sc = scenes.instanced_grid(scenes.synthetic_atrium(0.05, environment=False), 2)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc)
    closest, anyr = scenes.incoherent_rays(sc, 200000, seed=9)
    rays = torch.from_numpy(np.concatenate([closest, scenes.random_rays(sc, 50000, seed=10)]).view(np.float32).reshape(-1, 8)).cuda()
    arays = torch.from_numpy(anyr.view(np.float32).reshape(-1, 8)).cuda()
    results = {}
    for name, flags in (("trace", BUILD_FAST_TRACE), ("build", BUILD_FAST_BUILD), ("trace2", BUILD_FAST_TRACE | BUILD_TWO_LEVEL), ("build2", BUILD_FAST_BUILD | BUILD_TWO_LEVEL)):
        bs = w.create_acceleration_structure(flags)
        results[name] = (bs.wide_nodes, w.trace_closest(rays).cpu().numpy().view(np.uint32), w.trace_any(arays).cpu().numpy())
        if not flags & BUILD_TWO_LEVEL:
            nodes, tris = w.download_accel()
            wrapped = EmuBvh(nodes=nodes, tri_records=tris)
            assert wrapped.check_triangles() == 0 and wrapped.check_node_format() == 0
    for name in ("build", "trace2", "build2"):
        assert (results[name][1] == results["trace"][1]).all() and (results[name][2] == results["trace"][2]).all(), name
    assert results["trace"][0] < 0.85 * results["build"][0] and results["trace2"][0] < 0.85 * results["build2"][0]
    assert (results["trace"][1][:, 3] != MISS).mean() > 0.5


def test_treelet_build_gives_identical_hits_and_the_emulators_tree(gpu_ctx):
    """The default FAST_TRACE build (treelet restructuring) against RT3_BUILD_ROTATIONS: same hits bit for bit, every triangle kept once, and
    exactly the tree the CPU emulation of the builder produces (node count), with less traversal work than the rotations.  First run on a
    B200 in round 2 (profiles/r2a_ab_experiments.log)."""
    import sys, os, torch
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from emu_lib import EmuBvh
    from raytracer3_b200.host import RenderWorld
    from raytracer3_b200.types import BUILD_FAST_TRACE, BUILD_ROTATIONS, BUILD_TREELET, BUILD_TWO_LEVEL
    sc = scenes.synthetic_atrium(0.05, environment=False)
    T = np.ascontiguousarray(scenes.world_triangles(sc), np.float32)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc)
    closest, anyr = scenes.incoherent_rays(sc, 100000, seed=9, tris=T.astype(np.float64))
    rays = torch.from_numpy(np.concatenate([closest, scenes.random_rays(sc, 20000, seed=10)]).view(np.float32).reshape(-1, 8)).cuda()
    arays = torch.from_numpy(anyr.view(np.float32).reshape(-1, 8)).cuda()
    res = {}
    for name, flags in (("rot", BUILD_ROTATIONS), ("treelet", BUILD_FAST_TRACE), ("treelet_explicit", BUILD_FAST_TRACE | BUILD_TREELET),
                        ("treelet2", BUILD_FAST_TRACE | BUILD_TWO_LEVEL)):
        bs = w.create_acceleration_structure(flags)
        st = w.trace_stats(rays)
        res[name] = (bs.wide_nodes, w.trace_closest(rays).cpu().numpy().view(np.uint32), w.trace_any(arays).cpu().numpy(), st.nodes_per_ray + 0.4 * st.tris_per_ray)
        if not flags & BUILD_TWO_LEVEL:
            nodes, tris = w.download_accel()
            wrapped = EmuBvh(nodes=nodes, tri_records=tris)
            assert wrapped.check_triangles() == 0 and wrapped.check_node_format() == 0
    for name in ("treelet", "treelet_explicit", "treelet2"):
        assert (res[name][1] == res["rot"][1]).all() and (res[name][2] == res["rot"][2]).all(), name
    emu = EmuBvh(tris=T, treelet_passes=2)
    assert res["treelet"][0] == res["treelet_explicit"][0] == emu.info()["nodes"]                                # the device build is the emulator's build
    # ... down to the visit statistics: the device restructures treelets with a warp-cooperative solver (build.cu CoopTreeletDp) that must
    # return the tables of the serial definition the emulator runs, ties included; one different choice anywhere changes these averages
    sub = np.ascontiguousarray(np.concatenate([closest, scenes.random_rays(sc, 20000, seed=10)])[:30000])
    w.create_acceleration_structure(BUILD_FAST_TRACE)
    st_gpu = w.trace_stats(torch.from_numpy(sub.view(np.float32).reshape(-1, 8)).cuda())
    _, st_emu = emu.trace_closest(sub)
    assert abs(st_gpu.nodes_per_ray - st_emu[0]) <= 1e-9 * st_emu[0] and abs(st_gpu.tris_per_ray - st_emu[1]) <= 1e-9 * st_emu[1], (st_gpu.nodes_per_ray, st_emu)
    assert res["rot"][0] == EmuBvh(tris=T).info()["nodes"]
    assert res["treelet"][3] < res["rot"][3]                               # fewer node visits + 0.4 x triangle tests per ray
    assert (res["rot"][1][:, 3] != MISS).mean() > 0.5


def test_calls_on_different_streams_are_ordered_by_the_context(gpu_ctx):
    """All pass / trace calls of one context share its wavefront queues and counters (rt3.h conventions): a call recorded on another stream
    than the previous one must wait, on the device, for the previous call.  Frames and ray queries are issued alternately on two streams
    with no host synchronisation in between; every result must equal the one computed alone."""
    import torch
    from raytracer3_b200.host import Renderer, RenderWorld
    sc = scenes.synthetic_atrium(0.03, environment=False)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    W, H = 640, 360
    r = Renderer(gpu_ctx, w, W, H)
    closest, _ = scenes.incoherent_rays(sc, 300000, seed=3)
    rays = torch.from_numpy(closest.view(np.float32).reshape(-1, 8)).cuda()
    ps = [make_params(camera_for(sc, "cam0", W, H), W, H, frame=f, bounces=4, samples=2) for f in range(4)]
    want_frames = []
    for p in ps:
        want_frames.append(r.frame(p).clone())
    want_hits = w.trace_closest(rays).clone()
    torch.cuda.synchronize()
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    got_frames, got_hits = [], []
    for k, p in enumerate(ps):
        with torch.cuda.stream(sa if k % 2 == 0 else sb):
            got_frames.append(r.frame(p).clone())
        with torch.cuda.stream(sb if k % 2 == 0 else sa):
            got_hits.append(w.trace_closest(rays, torch.empty((rays.shape[0], 4), dtype=torch.float32, device="cuda")))
    torch.cuda.synchronize()
    for k in range(len(ps)):
        assert torch.equal(got_frames[k].view(torch.int32), want_frames[k].view(torch.int32)), k
        assert torch.equal(got_hits[k].view(torch.int32), want_hits.view(torch.int32)), k


def test_failed_rebuild_invalidates_the_acceleration_structure(gpu_ctx):
    """A build that fails must not leave the previous structure usable (its buffers are being overwritten): the next pass is RT3_ERR_STATE."""
    from raytracer3_b200 import _lib
    from raytracer3_b200.host import RenderWorld
    from raytracer3_b200.types import BUILD_TWO_LEVEL
    sc = scenes.instanced_grid(scenes.cornell_box(), 2)
    f, n, m = sc["instances"][0]
    sc["instances"][0] = (f, n, np.zeros(12, np.float32))                        # singular transform: fine when flattened, rejected by the two-level build
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    rays = scenes.random_rays(sc, 100)
    assert len(w.trace_closest_host(rays)) == 100
    with pytest.raises(_lib.Rt3Error):                                           # same scene, other flags: the re-build fails midway
        w.create_acceleration_structure(1 | BUILD_TWO_LEVEL)
    with pytest.raises(_lib.Rt3Error) as e:
        w.trace_closest_host(rays)
    assert e.value.status == _lib.ERR_STATE
    w.create_acceleration_structure()                                            # and a good build makes it usable again
    assert len(w.trace_closest_host(rays)) == 100


def test_far_origins_and_tiny_objects_are_not_culled(gpu_ctx):
    """Conservative culling (bvh_logic.h RT3_PAD_FRACTION): rays that start tens of scene extents away, and a tiny object inside a huge
    scene, against BRUTE FORCE over every triangle on the CPU — no true hit may be lost to the quantised slab test."""
    from oracle.oracle_py import OracleScene
    from raytracer3_b200.host import RenderWorld
    from raytracer3_b200.types import RAY
    rng = np.random.default_rng(77)
    # (a) far origins aimed at the scene
    sc = scenes.synthetic_atrium(0.02, environment=False)
    orc = OracleScene(sc)
    lo, hi = orc.bounds()
    ext = float((hi - lo).max())
    n = 60000
    tgt = rng.uniform(lo, hi, (n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist = rng.choice([3.0, 30.0, 100.0], n) * ext
    rays = np.zeros(n, RAY)
    rays["o"] = tgt - d * dist[:, None]
    rays["d"] = d
    rays["tmin"], rays["tmax"] = 0.0, 1e9
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    got, want = w.trace_closest_host(rays), orc.trace_closest(rays, brute=True)
    assert (want["prim"] != MISS).mean() > 0.5
    assert (got["prim"] == want["prim"]).all(), int((got["prim"] != want["prim"]).sum())
    assert (_bits(got["t"]) == _bits(want["t"])).all()
    # (b) a 1 cm triangle pair inside a 10 km quad: the scene extent sets the padding, the small object must still be found
    mb = scenes.MeshBuilder()
    mb.add(*scenes.quad((-5000, 0, -5000), (0, 0, 10000), (10000, 0, 0)), base_color=(0.5, 0.5, 0.5))
    mb.add(*scenes.quad((0.3, 0.5, 0.2), (0.01, 0, 0), (0, 0.01, 0)), base_color=(0.9, 0.1, 0.1))
    sc2 = scenes._scene(mb, {"cam0": dict(position=(0, 1, -2), direction=(0, 0, 1), fov_deg=60.0)}, name="tiny_in_huge")
    orc2 = OracleScene(sc2)
    m = 20000
    rays2 = np.zeros(m, RAY)
    rays2["o"] = rng.uniform((-3, 0.2, -3), (3, 3, 3), (m, 3))
    aim = np.array([0.305, 0.505, 0.2]) + rng.uniform(-0.008, 0.008, (m, 3)) * np.array([1, 1, 0])
    dd = aim - rays2["o"]; dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    rays2["d"] = dd
    rays2["tmin"], rays2["tmax"] = 0.0, 1e9
    w.clear(); w.load_scene(sc2); w.create_acceleration_structure()
    got2, want2 = w.trace_closest_host(rays2), orc2.trace_closest(rays2, brute=True)
    assert (want2["prim"] >= 2).mean() > 0.2                                   # the tiny quad is actually hit
    assert (got2["prim"] == want2["prim"]).all(), int((got2["prim"] != want2["prim"]).sum())


def test_present_bgra8_matches_oracle(gpu_ctx, cornell):
    """rt3_pass_present_bgra8 = postprocess.slang + the swapchain format (B8G8R8A8_UNORM, swapchain.rs:27-48): one word per pixel, bit-identical
    to the oracle and consistent with the RGBA32F post-process output (|float * 255 - byte| <= 0.5)."""
    import torch
    from raytracer3_b200.host import ComputePass, Renderer, WorkSize2D
    sc, orc = cornell
    w = _world(gpu_ctx, sc)
    W, H = 200, 120
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=2, bounces=4, samples=4, flags=FLAG_ENV)
    r = Renderer(gpu_ctx, w, W, H)
    light = r.frame(p, postprocess=True)
    rg = r.rg
    rg.image("B8G8R8A8", "present")
    (ComputePass.new(rg, "present").shader("present").constants(p).read("Depth", "gbuffer_depth").read("In", f"light{1 - r.cur}").write("Out", "present")
     .dispatch(WorkSize2D.FullScreen()))
    got = rg.tensor("present").cpu().numpy().view(np.uint32)
    gb, dp, _, _ = orc.pass_gbuffer(p)
    lt, _ = orc.pass_reference(p, gb, dp)
    want = orc.pass_present_bgra8(p, dp, lt)
    assert (got == want).all()
    out = rg.tensor("out").cpu().numpy()
    for shift, ch in ((16, 0), (8, 1), (0, 2)):
        byte = ((got >> shift) & 0xFF).astype(np.float64)
        assert np.abs(np.clip(out[..., ch], 0, 1).astype(np.float64) * 255.0 - byte).max() <= 0.5 + 1e-4
    assert ((got >> 24) == 255).all() and len(np.unique(got)) > 500


def test_strip_count_does_not_change_the_image_and_options_are_checked(gpu_ctx):
    """RT3_OPT_STRIPS = 1 .. 4 concurrent wavefronts per reference-pass call (regeneration and NEE included): bit-identical images."""
    import torch
    from raytracer3_b200 import _lib
    from raytracer3_b200.host import Renderer, RenderWorld
    sc = scenes.synthetic_atrium(0.03)
    sc["bluenoise"] = scenes.blue_noise_tile(64, iters=4)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    W, H = 640, 360
    r = Renderer(gpu_ctx, w, W, H)
    cam = camera_for(sc, "cam0", W, H)
    try:
        for flags, samples in ((FLAG_ENV | FLAG_BLUENOISE, 1), (FLAG_ENV | FLAG_NEE, 3)):
            p = make_params(cam, W, H, frame=7, bounces=4, samples=samples, flags=flags)
            images = []
            for strips in (1, 2, 3, 4):
                gpu_ctx.set_strips(strips)
                images.append(r.frame(p).clone())
            for k in range(1, 4):
                assert torch.equal(images[0].view(torch.int32), images[k].view(torch.int32)), (flags, k + 1)
            assert float(images[0][..., :3].sum()) > 0
        with pytest.raises(_lib.Rt3Error):
            gpu_ctx.set_strips(5)
        with pytest.raises(_lib.Rt3Error):
            gpu_ctx.check(gpu_ctx.L.rt3_set_option(gpu_ctx.h, 99, 1))
    finally:
        gpu_ctx.set_strips(0)


def test_tail_statistics_of_a_traversal_launch(gpu_ctx):
    """rt3_trace_closest_tail_stats: the instrumented instantiation gives the same hits' worth of work and sane timings."""
    import torch
    from raytracer3_b200.host import RenderWorld
    sc = scenes.synthetic_atrium(0.03, environment=False)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    closest, _ = scenes.incoherent_rays(sc, 400000, seed=5)
    rays = torch.from_numpy(closest.view(np.float32).reshape(-1, 8)).cuda()
    ts = w.trace_tail_stats(rays)
    assert ts.warps > 0 and ts.kernel_us > 1.0
    assert 0.0 < ts.mean_warp_tail_us <= ts.mean_warp_busy_us <= ts.kernel_us * 1.01
    assert 0.0 < ts.tail_fraction < 1.0
