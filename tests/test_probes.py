"""Screen-space probe GI: octahedral mapping (packing.slang:64-86), SH L2 basis and cosine-lobe
unprojection (spherical_harmonics.slang), world_pos_from_depth (gbuffer_helpers.slang:81-83) and the three probe passes.

The reference's probe shaders are unfinished ("parity unpinned", no tests): the oracle is pinned by float64 re-derivations, by the
orthonormality of the SH basis under the pass's own 8x8 octahedral quadrature, by a closed-form furnace (constant radiance L in every
direction => irradiance pi L => pixel radiance albedo * L) and on the GPU by bit-identity with the oracle."""
import numpy as np
import pytest

from oracle.oracle_py import OracleScene, f3, lib
from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_ENV, MISS, ptr

L = lib()


def octa_decode64(f):
    f = np.asarray(f, np.float64) * 2 - 1
    n = np.array([f[0], f[1], 1 - abs(f[0]) - abs(f[1])])
    t = min(max(-n[2], 0.0), 1.0)
    n[0] -= (1.0 if n[0] >= 0 else -1.0) * t
    n[1] -= (1.0 if n[1] >= 0 else -1.0) * t
    return n / np.linalg.norm(n)


def sh9_64(d):
    x, y, z = d
    return np.array([0.28209479177387814, -0.4886025119029199 * y, 0.4886025119029199 * z, -0.4886025119029199 * x, 1.0925484305920792 * x * y,
                     1.0925484305920792 * y * z, 0.31539156525252005 * (3 * z * z - 1), 1.0925484305920792 * x * z, 0.5462742152960396 * (x * x - y * y)])


def test_octahedral_mapping_round_trip_and_float64():
    rng = np.random.default_rng(31)
    o3, o2 = np.zeros(3, np.float32), np.zeros(2, np.float32)
    for _ in range(3000):
        f = rng.random(2)
        L.orc_octa_decode(float(f[0]), float(f[1]), ptr(o3))
        assert np.abs(o3 - octa_decode64(np.float32(f))).max() < 2e-6
        assert abs(np.linalg.norm(o3) - 1) < 1e-6
        L.orc_octa_encode(ptr(f3(o3)), ptr(o2))
        assert np.abs(o2 - f).max() < 2e-5                          # encode(decode(f)) == f
    for d, want in (([0, 0, 1], [0.5, 0.5]), ([1, 0, 0], [1.0, 0.5]), ([0, -1, 0], [0.5, 0.0])):
        L.orc_octa_encode(ptr(f3(d)), ptr(o2))
        assert np.allclose(o2, want, atol=1e-6)
    L.orc_octa_encode(ptr(f3([0, 0, -1])), ptr(o2))                 # the -z pole unfolds to a corner of the square
    assert np.allclose(np.abs(o2 * 2 - 1), [1, 1], atol=1e-6)


def test_sh_basis_and_cos_lobe():
    rng = np.random.default_rng(32)
    o9 = np.zeros(9, np.float32)
    for _ in range(500):
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        L.orc_sh3_evaluate(ptr(f3(d)), ptr(o9))
        assert np.abs(o9 - sh9_64(f3(d).astype(np.float64))).max() < 1e-6
    # the basis is orthonormal under the 8x8 octahedral mid-point quadrature that spherical_harmonic_conversion.slang uses (up to
    # the quadrature error of 64 points): sum_k Y_i Y_j * 4 pi / 64 ~ delta_ij
    G = np.zeros((9, 9))
    for k in range(64):
        y = sh9_64(octa_decode64([(k % 8 + 0.5) / 8, (k // 8 + 0.5) / 8]))
        G += np.outer(y, y) * 4 * np.pi / 64
    assert np.abs(G[:4, :4] - np.eye(4)).max() < 0.15
    # cosine-lobe unprojection: constant radiance 1 has SH (sqrt(4 pi), 0, ...) and irradiance pi for every normal
    sh27 = np.zeros(27, np.float32)
    sh27[0] = sh27[9] = sh27[18] = np.sqrt(4 * np.pi)
    o3 = np.zeros(3, np.float32)
    for n in ([0, 0, 1], [0.6, -0.8, 0], [-0.3, 0.5, 0.81]):
        n = np.array(n, np.float64) / np.linalg.norm(n)
        L.orc_sh3_unproject_cos_lobe(ptr(sh27), ptr(f3(n)), ptr(o3))
        assert np.allclose(o3, np.pi, rtol=1e-5)
    # a band-1 function L(w) = max-free linear a.w integrates against the clamped cosine to (2 pi / 3) a.n
    a = np.array([0.2, -0.5, 0.7])
    c1 = np.zeros(9); c1[1], c1[2], c1[3] = -a[1] / 0.4886025119029199 * (0.4886025119029199 ** 2) * 4 * np.pi / 3 / 0.4886025119029199, 0, 0
    coeff = np.zeros(9)
    coeff[1], coeff[2], coeff[3] = -a[1], a[2], -a[0]
    coeff[1:4] *= np.sqrt(4 * np.pi / 3)                             # projection of a.w on the real SH band 1
    sh27[:] = 0; sh27[0:9] = coeff
    n = np.array([0.48, 0.6, 0.64])
    L.orc_sh3_unproject_cos_lobe(ptr(sh27), ptr(f3(n)), ptr(o3))
    assert o3[0] == pytest.approx(2 * np.pi / 3 * float(a @ n), rel=1e-4)


def _setup(W=320, H=192, frame=1, blend=1.0, flags=0, proberng=0):
    sc = scenes.cornell_box()
    if flags & FLAG_ENV:
        sc["environment"] = scenes.analytic_sky(64, 32)
    orc = OracleScene(sc)
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=frame, bounces=1, samples=1, blendfactor=blend, flags=flags)
    p["proberng"] = proberng
    gb, dp, ids, _ = orc.pass_gbuffer(p)
    return sc, orc, p, gb, dp, ids


def test_world_pos_from_depth_is_the_primary_hit_point():
    sc, orc, p, gb, dp, ids = _setup()
    o3 = np.zeros(3, np.float32)
    ray = np.zeros(1, dtype=[("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4")])
    ys, xs = np.nonzero(ids[..., 0] != MISS)
    assert len(ys) > 1000
    for k in (0, len(ys) // 3, len(ys) // 2, len(ys) - 1):
        x, y = int(xs[k]), int(ys[k])
        L.orc_world_pos_from_depth(ptr(p), x, y, float(dp[y, x]), ptr(o3))
        L.orc_primary_ray(ptr(p), x, y, ptr(ray))
        assert np.allclose(o3, ray["o"][0] + ray["d"][0] * dp[y, x], atol=1e-5)


def test_probe_passes_semantics():
    sc, orc, p, gb, dp, ids = _setup()
    W, H = 320, 192
    atlas = orc.pass_probe_trace(p, gb, dp)
    assert atlas.shape == (H // 16 * 8, W // 16 * 8, 4)
    # background probes: (0, 0, 0, BACKGROUND_DEPTH) (trace_probes.slang:28-31); others: alpha = hit distance > tmin
    for py in range(H // 16):
        for px in range(W // 16):
            blk = atlas[py * 8:py * 8 + 8, px * 8:px * 8 + 8]
            if dp[py * 16, px * 16] == 100000.0:
                assert (blk[..., :3] == 0).all() and (blk[..., 3] == 100000.0).all()
            else:
                assert (blk[..., 3] > 0.0005).all()
    # only the light emits in the Cornell box: radiance is either 0 or emission * 12 (hit_logic.slang:36)
    vals = np.unique(np.round(atlas[..., 0], 4))
    assert len(vals) == 2 and vals[0] == 0 and vals[1] == pytest.approx(1.4 * 12, rel=1e-4)
    # temporal blend of the atlas
    p2 = p.copy(); p2["blendfactor"] = 0.25; p2["frame"] = 2
    a2 = orc.pass_probe_trace(p2, gb, dp, prev_atlas=atlas)
    p2b = p2.copy(); p2b["blendfactor"] = 1.0
    raw = orc.pass_probe_trace(p2b, gb, dp)
    assert np.allclose(a2[..., :3], atlas[..., :3] + (raw[..., :3] - atlas[..., :3]) * 0.25, atol=1e-5)
    # explicit direction words equal to the uniform assignment give the same atlas; mip-1 words address a 16x16 grid
    ay, ax = np.mgrid[0:atlas.shape[0], 0:atlas.shape[1]]
    words = ((ay % 8) * 8 + (ax % 8)).astype(np.uint16)
    assert np.array_equal(orc.pass_probe_trace(p, gb, dp, directions=words), atlas)
    words1 = (np.uint16(1 << 15) | (((ay % 8) * 2) * 16 + (ax % 8) * 2).astype(np.uint16))
    a_m1 = orc.pass_probe_trace(p, gb, dp, directions=words1)
    assert np.isfinite(a_m1).all() and not np.array_equal(a_m1, atlas)
    sh = orc.pass_probe_sh(p, atlas)
    lt = orc.pass_probe_interpolate(p, gb, dp, sh)
    hit = dp != 100000.0
    # a pixel whose interpolation fails writes its red marker at the JITTERED coordinate and may itself stay unwritten (:71-76)
    assert np.isfinite(lt).all() and (lt[hit][:, 3] == 1).mean() > 0.9 and (lt[~hit][:, 3] == 1).mean() < 0.05
    assert 0.05 < lt[hit][:, :3].mean() < 5.0
    # debug view of interpolate_probes.slang:84-88: probe normals as colours, in [0, 1]
    pdbg = p.copy(); pdbg["proberng"] = 1
    dbg = orc.pass_probe_interpolate(pdbg, gb, dp, sh)
    ok = hit & ~((dbg[..., 0] == 1) & (dbg[..., 1] == 0) & (dbg[..., 2] == 0))
    nonem = ok & (lt[..., 0] < 10)                                    # away from the emitter: colour <= albedo / pi
    assert (dbg[nonem][:, :3] <= 1.0 / np.pi + 1e-5).all() and (dbg[nonem][:, :3] >= 0).all() and not np.array_equal(dbg, lt)


def test_probe_furnace_constant_sky():
    """Every probe direction sees radiance L (uniform environment, open scene): SH projection + cosine-lobe unprojection must give
    irradiance pi L, i.e. pixel radiance = albedo * L (up to the 64-direction quadrature of the pass: the L=0 term is exact)."""
    sc = scenes.random_soup(1, n_geometries=1)                       # a single triangle: nearly everything escapes
    tri = scenes.world_triangles(sc)[0]
    sc["environment"] = np.full((8, 16, 3), 0.75, np.float32)
    cen = tri.mean(0)
    nrm = np.cross(tri[1] - tri[0], tri[2] - tri[0]); nrm /= np.linalg.norm(nrm)
    sc["cameras"] = {"cam0": dict(position=tuple(cen + nrm * 0.05), direction=tuple(-nrm), fov_deg=20.0)}
    sc["geometries"]["base_color"][0] = [0.5, 0.5, 0.5, 1.0]
    sc["geometries"]["emission"][0] = 0.0
    sc["geometries"]["base_color_tex"][0] = -1
    orc = OracleScene(sc)
    W, H = 64, 64
    p = make_params(camera_for(sc, "cam0", W, H), W, H, frame=0, bounces=1, samples=1, blendfactor=1.0, flags=FLAG_ENV)
    gb, dp, ids, _ = orc.pass_gbuffer(p)
    hit = ids[..., 0] != MISS
    if hit[::16, ::16].all():
        atlas = orc.pass_probe_trace(p, gb, dp)
        sh = orc.pass_probe_sh(p, atlas)
        lt = orc.pass_probe_interpolate(p, gb, dp, sh)
        # half of the directions leave through the back of the one-sided... the triangle blocks nothing above its plane: the upper
        # hemisphere sees the sky (0.75), the lower one too (single triangle, rays start ON it): constant radiance everywhere
        good = hit & ~((lt[..., 0] == 1) & (lt[..., 1] == 0))
        assert good.sum() > 100
        assert np.allclose(lt[good][:, :3], 0.5 * 0.75, rtol=0.05)
    else:
        pytest.skip("camera does not cover the triangle at every probe anchor")


def test_structured_importance_sampling_live_lines():
    """structured_importance_sampling.slang:19-72 as it runs today: the pdf of a direction is the mean clamped cosine against the probe's 256
    G-buffer normals; `brdf_pdf < 0` never holds, the sort is ascending, so the test at :55 is `min pdf < pdf`: every thread whose pdf exceeds
    the probe's minimum writes 0x8000 | ti * 4 (mip 1), the minimum-pdf thread(s) write ti (mip 0).  Checked against a float64 restatement."""
    sc, orc, p, gb, dp, ids = _setup()
    W, H = 320, 192
    words = orc.pass_probe_sis(p, gb)
    assert words.shape == (H // 16 * 8, W // 16 * 8) and words.dtype == np.uint16
    o3 = np.zeros(3, np.float32)
    dirs = np.array([octa_decode64(np.float32([(k % 8 + 0.5) / 8.0, (k // 8 + 0.5) / 8.0])) for k in range(64)])
    n_min_total = 0
    for pcy in range(H // 16):
        for pcx in range(W // 16):
            normals = np.zeros((256, 3))
            for ti in range(64):
                for y in range(2):
                    for x in range(2):
                        px, py = (pcx * 8 + ti % 8) * 2 + x, (pcy * 8 + ti // 8) * 2 + y
                        L.orc_unpack_normal(int(gb[py, px, 1]), ptr(o3))
                        normals[ti * 4 + y * 2 + x] = o3
            pdf = np.maximum(normals @ dirs.T, 0.0).sum(0) / 256.0
            blk = words[pcy * 8:pcy * 8 + 8, pcx * 8:pcx * 8 + 8].reshape(64).astype(np.int64)
            ti = np.arange(64)
            is_min = blk == ti                                                    # ti == 0 writes 0 in the else-branch and 0x8000 otherwise
            assert (blk[~is_min] == (0x8000 | (ti[~is_min] * 4))).all()
            assert is_min.any() and (pdf[is_min] <= pdf.min() + 1e-6).all()
            assert (pdf[~is_min] >= pdf.min() - 1e-6).all()
            n_min_total += int(is_min.sum())
    # every direction in the hemisphere behind all 256 normals has pdf == 0 == the minimum: on flat surfaces about half of the threads tie at
    # the minimum and keep their own cell (mip 0), the other half — every direction the surface can see — takes the mip-1 branch
    frac_mip1 = 1.0 - n_min_total / words.size
    assert 0.3 < frac_mip1 < 0.8, frac_mip1
    # what trace_probes makes of these words: texels that no thread aims at hold the clear value (trace_probes.slang:32), the others a hit
    atlas = orc.pass_probe_trace(p, gb, dp, directions=words)
    for pcy in range(H // 16):
        for pcx in range(W // 16):
            if dp[pcy * 16, pcx * 16] == 100000.0:
                continue
            blk = words[pcy * 8:pcy * 8 + 8, pcx * 8:pcx * 8 + 8].reshape(64).astype(np.int64)
            aimed = np.zeros(64, bool)
            for ti in range(64):
                idx, size = blk[ti] & 0x7FFF, 16 if blk[ti] >> 15 else 8
                aimed[int((idx // size) / size * 8) * 8 + int((idx % size) / size * 8)] = True
            a = atlas[pcy * 8:pcy * 8 + 8, pcx * 8:pcx * 8 + 8].reshape(64, 4)
            assert (a[~aimed] == 0).all() and (a[aimed][:, 3] > 0.0005).all()
            assert 16 <= aimed.sum() <= 64


@pytest.mark.gpu
def test_gpu_structured_importance_sampling_matches_oracle(gpu_ctx):
    """rt3_pass_probe_sis (64 threads per probe, shared-memory bitonic sort) against the oracle, and the probe frame driven by its words."""
    from raytracer3_b200.host import Renderer, RenderWorld
    W, H = 320, 192
    sc, orc, p, gb, dp, ids = _setup(W, H, flags=FLAG_ENV)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    r = Renderer(gpu_ctx, w, W, H)
    prev = prev_light = None
    bits = lambda a: np.ascontiguousarray(a).view(np.uint32)
    for frame in range(2):
        p["frame"] = frame
        p["blendfactor"] = 1.0 if frame == 0 else 0.5
        light = r.probe_frame(p, sis=True).cpu().numpy()
        cur = 1 - r.probe_cur
        words = r.rg.tensor("probe_directions").cpu().numpy().view(np.uint16)
        want_words = orc.pass_probe_sis(p, gb)
        assert (words == want_words).all(), f"{(words != want_words).sum()} direction words differ"
        assert 0.3 < ((want_words >> 15) == 1).mean() < 0.8
        atlas = r.rg.tensor(f"probe_atlas{cur}").cpu().numpy()
        want_atlas = orc.pass_probe_trace(p, gb, dp, prev_atlas=prev, directions=want_words)
        want_sh = orc.pass_probe_sh(p, want_atlas)
        want_light = orc.pass_probe_interpolate(p, gb, dp, want_sh, light=prev_light)
        prev, prev_light = want_atlas, want_light
        assert (bits(atlas) == bits(want_atlas)).all(), f"atlas differs, frame {frame}"
        assert (bits(light) == bits(want_light)).all(), f"interpolated radiance differs, frame {frame}"


@pytest.mark.gpu
@pytest.mark.parametrize("flags,blend", [(0, 1.0), (FLAG_ENV, 0.5)])
def test_gpu_probe_passes_match_oracle(gpu_ctx, flags, blend):
    from raytracer3_b200.host import Renderer, RenderWorld
    W, H = 320, 192
    sc, orc, p, gb, dp, ids = _setup(W, H, flags=flags)
    w = RenderWorld(gpu_ctx)
    w.clear(); w.load_scene(sc); w.create_acceleration_structure()
    r = Renderer(gpu_ctx, w, W, H)
    prev = prev_light = None
    for frame in range(2):
        p["frame"] = frame
        p["blendfactor"] = 1.0 if frame == 0 else blend
        light = r.probe_frame(p).cpu().numpy()
        cur = 1 - r.probe_cur
        atlas = r.rg.tensor(f"probe_atlas{cur}").cpu().numpy()
        sh = r.rg.tensor("probe_sh").cpu().numpy()
        want_atlas = orc.pass_probe_trace(p, gb, dp, prev_atlas=prev)
        want_sh = orc.pass_probe_sh(p, want_atlas)
        want_light = orc.pass_probe_interpolate(p, gb, dp, want_sh, light=prev_light)   # untouched pixels keep the last frame's value
        prev, prev_light = want_atlas, want_light
        bits = lambda a: np.ascontiguousarray(a).view(np.uint32)
        assert (bits(atlas) == bits(want_atlas)).all(), f"atlas differs, frame {frame}"
        used = np.abs(want_sh).sum(1) > 0
        assert (bits(sh[used]) == bits(want_sh[used])).all(), "SH coefficients differ"
        assert (bits(light) == bits(want_light)).all(), f"interpolated radiance differs, frame {frame}"
    assert light[..., :3].mean() > 0.01
