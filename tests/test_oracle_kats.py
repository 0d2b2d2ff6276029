"""Pins the CPU oracle (oracle/) — the reference ships no tests, goldens or CPU code for this path ("parity unpinned" at the
driver boundary), so the oracle is pinned by published known answers (MurmurHash3_x86_32, Jenkins integer hash, IEEE half,
EXT_texture_shared_exponent), by independent numpy/float64 re-derivations of every restated Slang function, by brute-force ==
BVH equivalence, by an analytic furnace test and by frozen golden vectors (tests/golden/, made by tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle.oracle_py import OracleScene, f3, lib
from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_BLUENOISE, FLAG_ENV, FLAG_NEE, MISS, RAY, ptr

L = lib()
M32 = 0xFFFFFFFF


def py_jenkins(a):  # shaders/include/random.slang:5-15, re-derived independently in Python integers
    a = ((a + 0x7ED55D16) + (a << 12)) & M32
    a = ((a ^ 0xC761C23C) ^ (a >> 19)) & M32
    a = ((a + 0x165667B1) + (a << 5)) & M32
    a = ((a + 0xD3A2646C) ^ (a << 9)) & M32
    a = ((a + 0xFD7046C5) + (a << 3)) & M32
    a = ((a ^ 0xB55A4F09) ^ (a >> 16)) & M32
    return a


def py_murmur3_32(key4, seed):  # MurmurHash3_x86_32 of one little-endian 32-bit block (Wikipedia / SMHasher)
    rotl = lambda x, r: ((x << r) | (x >> (32 - r))) & M32
    k = (key4 * 0xCC9E2D51) & M32
    k = rotl(k, 15)
    k = (k * 0x1B873593) & M32
    h = seed ^ k
    h = (rotl(h, 13) * 5 + 0xE6546B64) & M32
    h ^= 4
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & M32
    h ^= h >> 16
    return h


def test_murmur3_published_vectors():
    # SMHasher-verified vectors for 4-byte keys: ("\0\0\0\0", seed 0) and ("\xff\xff\xff\xff", seed 0), ("!Ce\x87", seed 0 / 0x5082EDEE)
    assert L.orc_murmur3(0, 0) == 0x2362F9DE
    assert L.orc_murmur3(0, 0xFFFFFFFF) == 0x76293B50
    assert L.orc_murmur3(0, 0x87654321) == 0xF55B516B
    assert L.orc_murmur3(0x5082EDEE, 0x87654321) == 0x2362F9DE
    rng = np.random.default_rng(0)
    for s, i in rng.integers(0, 2**32, (200, 2)):
        assert L.orc_murmur3(int(s), int(i)) == py_murmur3_32(int(i), int(s))


def test_jenkins_zcurve_seed():
    rng = np.random.default_rng(1)
    for a in [0, 1, 2, 0xFFFFFFFF, 0x12345678] + list(rng.integers(0, 2**32, 200)):
        assert L.orc_hash_jenkins(int(a)) == py_jenkins(int(a))
    assert L.orc_zcurve(0, 0) == 0 and L.orc_zcurve(1, 0) == 1 and L.orc_zcurve(0, 1) == 2 and L.orc_zcurve(3, 3) == 15
    assert L.orc_zcurve(0xFFFF, 0) == 0x55555555 and L.orc_zcurve(0, 0xFFFF) == 0xAAAAAAAA
    assert L.orc_zcurve(0x1FFFF, 0) == 0x55555555  # only the low 16 bits of each coordinate are interleaved (math.slang:105-117)
    for x, y, f in [(0, 0, 0), (5, 9, 7), (1919, 1079, 63)]:
        z = L.orc_zcurve(x, y)
        assert L.orc_rng_seed(x, y, f) == (py_jenkins(z) + f) & M32


def test_uniform_float_mantissa_trick():
    rng = np.random.default_rng(2)
    for s, i in rng.integers(0, 2**32, (300, 2)):
        h = py_murmur3_32(int(i), int(s))
        want = np.array([(h & 0x7FFFFF) | 0x3F800000], np.uint32).view(np.float32)[0] - np.float32(1)
        got = L.orc_uniform_float(int(s), int(i))
        assert got == want and 0.0 <= got < 1.0


def test_vdc_and_r2():
    assert L.orc_vdc(1) == 0.5 and L.orc_vdc(2) == 0.25 and L.orc_vdc(3) == 0.75
    out = np.zeros(2, np.float32)
    phi = 1.32471795724474602596
    for i in (0, 1, 2, 17, 100):
        L.orc_r2(i, ptr(out))
        want = np.array([(i / phi + 0.5) % 1.0, (i / (phi * phi) + 0.5) % 1.0])
        assert np.abs(out - want).max() < 2e-5


def test_half_conversion_matches_numpy():
    rng = np.random.default_rng(3)
    vals = np.concatenate([rng.normal(0, 1, 2000), rng.normal(0, 1e-5, 500), rng.uniform(-70000, 70000, 500),
                           [0.0, -0.0, 1.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8, 6.1e-5, np.inf, -np.inf]]).astype(np.float32)
    for v in vals:
        want = int(np.array([v], np.float32).astype(np.float16).view(np.uint16)[0])
        assert L.orc_f32_to_f16(float(v)) == want, v
    for h in list(range(0, 0x7C01, 37)) + list(range(0x8000, 0xFC01, 41)):
        want = np.array([h], np.uint16).view(np.float16).astype(np.float32)[0]
        assert L.orc_f16_to_f32(h) == want


def np_rgb9e5(rgb):  # EXT_texture_shared_exponent, independent float64 derivation
    mx = (511.0 / 512.0) * 65536.0
    c = np.clip(np.asarray(rgb, np.float64), 0, mx)
    m = c.max()
    e = max(-16, int(np.floor(np.log2(m))) if m > 0 else -200) + 1 + 15
    d = 2.0 ** (e - 15 - 9)
    if int(np.floor(m / d + 0.5)) == 512:
        d *= 2
        e += 1
    r, g, b = (int(np.floor(x / d + 0.5)) for x in c)
    return (r << 23) | (g << 14) | (b << 5) | e


def test_rgb9e5():
    rng = np.random.default_rng(4)
    out = np.zeros(3, np.float32)
    for _ in range(500):
        rgb = (rng.random(3) * 10.0 ** rng.uniform(-4, 4)).astype(np.float32)
        p = L.orc_rgb9e5_pack(ptr(f3(rgb)))
        assert p == np_rgb9e5(rgb), rgb
        L.orc_rgb9e5_unpack(p, ptr(out))
        assert np.abs(out - np.clip(rgb, 0, 65408)).max() <= max(rgb.max(), 1e-9) / 256.0
    assert L.orc_rgb9e5_pack(ptr(f3([0, 0, 0]))) == 0
    assert L.orc_rgb9e5_pack(ptr(f3([16.8, 16.8, 16.8]))) == np_rgb9e5([16.8] * 3)


def test_normal_and_colour_packing():
    rng = np.random.default_rng(5)
    out = np.zeros(3, np.float32)
    for _ in range(500):
        n = rng.normal(size=3)
        n = (n / np.linalg.norm(n)).astype(np.float32)
        p = L.orc_pack_normal(ptr(f3(n)))
        want = (int(np.float32(np.float32(n[0] * np.float32(0.5) + np.float32(0.5)) * np.float32(2047) + np.float32(0.5)))
                | int(np.float32(np.float32(n[1] * np.float32(0.5) + np.float32(0.5)) * np.float32(1023) + np.float32(0.5))) << 11
                | int(np.float32(np.float32(n[2] * np.float32(0.5) + np.float32(0.5)) * np.float32(2047) + np.float32(0.5))) << 21)
        assert p == want
        L.orc_unpack_normal(p, ptr(out))
        assert abs(np.linalg.norm(out) - 1) < 1e-6 and np.abs(out - n).max() < 2.5e-3
        c = rng.random(3).astype(np.float32)
        pc = L.orc_pack_color(ptr(f3(c)))
        L.orc_unpack_color(pc, ptr(out))
        assert np.abs(np.sqrt(out) - np.sqrt(c)).max() <= 0.5 / 255 + 1e-6
    assert L.orc_pack_color(ptr(f3([1, 0, 0.25]))) == 255 | (128 << 16)


def test_gbuffer_round_trip():
    o4 = np.zeros(4, np.uint32)
    o11 = np.zeros(11, np.float32)
    L.orc_gbuffer_pack(ptr(f3([0.8, 0.2, 0.1])), ptr(f3([12.0, 6.0, 0.5])), ptr(f3([0.0, 0.6, 0.8])), 0.25, 0.5, ptr(o4))
    L.orc_gbuffer_unpack(ptr(o4), ptr(o11))
    assert np.abs(o11[0:3] - [0.8, 0.2, 0.1]).max() < 4e-3
    assert np.abs(o11[3:6] - [12.0, 6.0, 0.5]).max() < 12 / 256
    assert np.abs(o11[6:9] - [0.0, 0.6, 0.8]).max() < 2e-3
    assert abs(o11[9] - 0.25) < 1e-3 and o11[10] == 0.5
    assert (o4[2] & 0xFFFF) == 0x3800 and (o4[2] >> 16) == 0x3800  # sqrt(0.25) = 0.5 and 0.5 as halves


def test_trig_kernels_against_libm():
    u = np.concatenate([np.linspace(0, 1, 4001, endpoint=False), np.random.default_rng(6).random(4000)]).astype(np.float32)
    out = np.zeros(2, np.float32)
    err = 0
    for x in u:
        L.orc_sincos_2pi(float(x), ptr(out))
        err = max(err, abs(out[0] - np.sin(2 * np.pi * float(x))), abs(out[1] - np.cos(2 * np.pi * float(x))))
    assert err < 4e-7
    rng = np.random.default_rng(7)
    for y, x in rng.normal(size=(3000, 2)):
        assert abs(L.orc_atan2(float(y), float(x)) - np.arctan2(np.float32(y), np.float32(x))) < 2.5e-5
    for y in np.linspace(-1, 1, 2001):
        assert abs(L.orc_asin(float(y)) - np.arcsin(np.float32(y))) < 2.5e-5
    for x in 10.0 ** rng.uniform(-6, 6, 2000):
        assert abs(L.orc_log2(float(x)) - np.log2(np.float32(x))) < 2e-6 * max(1, abs(np.log2(x)))
    for x in rng.uniform(-20, 20, 2000):
        assert abs(L.orc_exp2(float(x)) / 2.0 ** np.float32(x) - 1) < 5e-7
    assert L.orc_atan2(0.0, 0.0) == 0.0 and L.orc_atan2(0.0, -1.0) == pytest.approx(np.pi, abs=1e-6)


def test_onb_and_cosine_sample():
    rng = np.random.default_rng(8)
    o9 = np.zeros(9, np.float32)
    for _ in range(500):
        n = rng.normal(size=3)
        n = (n / np.linalg.norm(n)).astype(np.float32)
        L.orc_onb(ptr(f3(n)), ptr(o9))
        b1, b2, nn = o9[0:3], o9[3:6], o9[6:9]
        assert (nn == n).all()
        for a, b in ((b1, b2), (b1, n), (b2, n)):
            assert abs(np.dot(a, b)) < 2e-6
        assert abs(np.linalg.norm(b1) - 1) < 2e-6 and abs(np.linalg.norm(b2) - 1) < 2e-6
        assert np.abs(np.cross(b1, b2) - n).max() < 3e-6
    o3 = np.zeros(3, np.float32)
    acc = np.zeros(3)
    us = rng.random((4000, 2)).astype(np.float32)
    for ux, uy in us:
        L.orc_diffuse_sample(float(ux), float(uy), ptr(o3))
        assert abs(np.linalg.norm(o3) - 1) < 1e-6 and o3[2] >= 0
        assert abs(o3[2] - np.sqrt(1 - uy)) < 1e-6            # brdf.slang:58: cos_theta = sqrt(1 - urand.y)
        assert abs(np.arctan2(o3[1], o3[0]) % (2 * np.pi) - 2 * np.pi * ux) < 2e-3 or o3[2] > 0.9999
        acc += o3
    assert abs(acc[2] / len(us) - 2.0 / 3.0) < 0.02 and np.abs(acc[:2] / len(us)).max() < 0.03   # E[cos] = 2/3 for a cosine lobe


def test_equirect_uv():
    o2 = np.zeros(2, np.float32)
    for d, want in (((1, 0, 0), (0.5, 0.5)), ((0, 1, 0), (0.5, 0.0)), ((0, 0, 1), (0.75, 0.5)), ((-1, 0, 1e-9), (1.0, 0.5)), ((0, 0, -1), (0.25, 0.5))):
        L.orc_equirect_uv(ptr(f3(d)), ptr(o2))
        assert np.abs(o2 - want).max() < 1e-5


def test_primary_ray_matches_float64_derivation():
    cam = camera_for(scenes.cornell_box(), "cam0", 640, 360)
    p = make_params(cam, 640, 360)
    from raytracer3_b200.types import RAY
    r = np.zeros(1, RAY)
    proj_inv = np.linalg.inv(cam.projection_matrix())
    view_inv = np.linalg.inv(cam.view_matrix())
    for px, py in ((0, 0), (639, 359), (320, 180), (17, 300)):
        L.orc_primary_ray(ptr(p), px, py, ptr(r))
        d = np.array([(px + 0.5) / 640 * 2 - 1, (py + 0.5) / 360 * 2 - 1, 1.0, 1.0])
        t = proj_inv @ d
        t3 = t[:3] / np.linalg.norm(t[:3])
        want = (view_inv @ np.array([*t3, 0.0]))[:3]
        assert np.abs(r["d"][0] - want).max() < 1e-6
        assert np.abs(r["o"][0] - np.array(cam.position)).max() < 1e-6
        assert r["tmin"][0] == 0.0 and r["tmax"][0] == 100000.0
    # no y flip (gbuffer_helpers.slang:85-93): pixel row 0 looks toward -y in view space
    L.orc_primary_ray(ptr(p), 320, 0, ptr(r))
    assert r["d"][0][1] < 0


def test_ray_triangle_vs_float64():
    rng = np.random.default_rng(9)
    out = np.zeros(3, np.float32)
    hits = 0
    for _ in range(3000):
        v = rng.normal(size=(3, 3)).astype(np.float32)
        o = (rng.normal(size=3) * 3).astype(np.float32)
        tgt = v.mean(0) + rng.normal(size=3) * 0.6
        d = (tgt - o).astype(np.float32)
        ok = L.orc_ray_triangle(ptr(f3(o)), ptr(f3(d)), ptr(f3(v[0])), ptr(f3(v[1])), ptr(f3(v[2])), ptr(out))
        e1, e2 = v[1].astype(np.float64) - v[0], v[2].astype(np.float64) - v[0]
        pv = np.cross(d.astype(np.float64), e2)
        det = e1 @ pv
        tv = o.astype(np.float64) - v[0]
        u = (tv @ pv) / det
        qv = np.cross(tv, e1)
        vv = (d.astype(np.float64) @ qv) / det
        t = (e2 @ qv) / det
        inside = u >= 0 and vv >= 0 and u + vv <= 1
        margin = min(u, vv, 1 - u - vv)
        if abs(margin) > 1e-4:
            assert bool(ok) == inside
        if ok and inside:
            hits += 1
            assert abs(out[0] - t) <= 2e-5 * max(1, abs(t)) and abs(out[1] - u) < 2e-5 and abs(out[2] - vv) < 2e-5
    assert hits > 300


def test_contraction_is_off():
    # a*b+c with two roundings differs from the fused result for these operands
    a, b, c = np.float32(1 + 2**-12), np.float32(1 + 2**-12), np.float32(-1)
    unfused = np.float32(np.float32(a * b) + c)
    assert L.orc_contract_probe(float(a), float(b), float(c)) == unfused
    assert float(unfused) != float(np.float64(a) * np.float64(b) + np.float64(c))


@pytest.mark.parametrize("maker", [lambda: scenes.cornell_box(), lambda: scenes.random_soup(1500, seed=2),
                                   lambda: scenes.synthetic_atrium(0.01, environment=False)])
def test_bvh_equals_brute_force(maker):
    sc = maker()
    orc = OracleScene(sc)
    c, a = scenes.incoherent_rays(sc, 4000, seed=21)
    rays = np.concatenate([c, scenes.random_rays(sc, 4000, seed=22)])
    assert (orc.trace_closest(rays) == orc.trace_closest(rays, brute=True)).all()
    assert (orc.trace_any(a) == orc.trace_any(a, brute=True)).all()


def test_instances_and_resolve_prim():
    base = scenes.synthetic_atrium(0.005, environment=False)
    grid = scenes.instanced_grid(base, 2)
    orc = OracleScene(grid)
    nt = scenes.scene_num_triangles(base)
    assert orc.num_triangles == 4 * nt
    per_geo = base["geometries"]["index_count"] // 3
    assert orc.resolve_prim(0) == (0, 0, 0)
    assert orc.resolve_prim(nt) == (1, 0, 0)
    assert orc.resolve_prim(nt + int(per_geo[0])) == (1, 1, 0)
    assert orc.resolve_prim(4 * nt - 1) == (3, len(per_geo) - 1, int(per_geo[-1]) - 1)
    T = orc.world_triangles()
    ext = base["vertices"]["p"].max(0) - base["vertices"]["p"].min(0)
    assert np.allclose(T[nt] - T[0], [1.05 * ext[0], 0, 0], atol=1e-4)
    want = scenes.world_triangles(grid)
    assert np.abs(T - want).max() < 1e-4


def test_furnace_analytic():
    """Closed box, albedo rho, emission e: L_B = 12 e * sum_{k<B} rho^k — except that bounce 0 reads the quantised G-buffer
    (888-sqrt albedo, RGB9E5 emissive), so the first factor uses the round-tripped values."""
    rho, e = 0.5, 0.25 / 12.0
    sc = scenes.furnace_box(rho, e)
    orc = OracleScene(sc)
    W, H = 64, 40
    for B in (1, 2, 4, 6):
        p = make_params(camera_for(sc, "cam0", W, H), W, H, bounces=B, samples=2)
        gb, dp, ids, _ = orc.pass_gbuffer(p)
        assert (ids[..., 0] != MISS).all()
        light, cnt = orc.pass_reference(p, gb, dp)
        o11 = np.zeros(11, np.float32)
        L.orc_gbuffer_unpack(ptr(np.ascontiguousarray(gb[0, 0])), ptr(o11))
        rho0, e0 = float(o11[0]), float(o11[3])
        want = e0 + rho0 * 12 * e * sum(rho ** k for k in range(B - 1))
        # a path vertex closer than TMin = 0.001 to a neighbouring wall can leak out of the box (the reference's only
        # self-hit guard is TMin, refrence_mode.slang:31); such paths lose their tail, everything else is exact.
        err = np.abs(light[..., :3] - want).max(-1)
        assert (err < 2e-6 * max(1, want)).mean() > 0.99 and (light[..., :3] <= want + 2e-6).all()
        assert cnt.paths_escaped < 0.002 * W * H * 2 * max(1, B - 1)
        assert W * H * 2 * (B - 1) - 3 * cnt.paths_escaped <= cnt.rays_secondary <= W * H * 2 * (B - 1)


def test_reference_pass_semantics():
    sc = scenes.cornell_box()
    orc = OracleScene(sc)
    W, H = 96, 60
    cam = camera_for(sc, "cam0", W, H)
    p = make_params(cam, W, H, bounces=4, samples=2, frame=5)
    gb, dp, ids, _ = orc.pass_gbuffer(p)
    bg = dp == 100000.0
    assert bg.any() and (~bg).any() and ((ids[..., 0] == MISS) == bg).all()
    light, _ = orc.pass_reference(p, gb, dp)
    assert (light[bg] == 0).all()                       # Light is not written for background pixels (refrence_mode.slang:18-21)
    light2 = np.full((H, W, 4), 7.0, np.float32)
    orc.pass_reference(p, gb, dp, light=light2)
    assert (light2[bg] == 7.0).all() and (light2[~bg][:, 3] == 0).all()
    # blend: lerp(prev, radiance, f) (refrence_mode.slang:61-65)
    p2 = make_params(cam, W, H, bounces=4, samples=2, frame=5, blendfactor=0.25)
    prev = np.full((H, W, 4), 2.0, np.float32)
    l3, _ = orc.pass_reference(p2, gb, dp, prev_light=prev)
    assert np.abs(l3[~bg][:, :3] - (2.0 + (light[~bg][:, :3] - 2.0) * 0.25)).max() < 1e-6
    # frame index changes the RNG stream; same frame reproduces bit-exactly
    l4, _ = orc.pass_reference(p, gb, dp)
    assert (l4 == light).all()
    l5, _ = orc.pass_reference(make_params(cam, W, H, bounces=4, samples=2, frame=6), gb, dp)
    assert (l5 != light).any()
    # rect: only pixels inside are touched
    l6 = np.full((H, W, 4), -1.0, np.float32)
    orc.pass_reference(p, gb, dp, rect=(8, 4, 40, 20), light=l6)
    inside = np.zeros((H, W), bool)
    inside[4:24, 8:48] = True
    assert (l6[~inside] == -1).all() and (l6[inside & ~bg] == light[inside & ~bg]).all()


def test_env_and_bluenoise_modes():
    sc = scenes.synthetic_atrium(0.01)
    sc["environment"] = scenes.analytic_sky(128, 64)
    sc["bluenoise"] = scenes.blue_noise_tile(64, iters=3)
    orc = OracleScene(sc)
    W, H = 96, 54
    cam = camera_for(sc, "cam0", W, H)
    base = make_params(cam, W, H, bounces=4, samples=4)
    gb, dp, _, _ = orc.pass_gbuffer(base)
    l0, c0 = orc.pass_reference(base, gb, dp)
    l1, c1 = orc.pass_reference(make_params(cam, W, H, bounces=4, samples=4, flags=FLAG_ENV), gb, dp)
    assert c0.paths_escaped > 0 and c1.paths_escaped == c0.paths_escaped      # same paths, sky adds light on escape
    assert l1[..., :3].sum() > l0[..., :3].sum() and (l1 >= l0).all()
    l2, _ = orc.pass_reference(make_params(cam, W, H, bounces=4, samples=4, flags=FLAG_ENV | FLAG_BLUENOISE), gb, dp)
    assert abs(l2[..., :3].mean() / l1[..., :3].mean() - 1) < 0.25            # same estimator, different sampler
    o2 = np.zeros(2, np.float32)
    us = []
    for i in range(0, 64, 2):
        L.orc_sample_2d(orc.h, FLAG_BLUENOISE, 10, 20, 3, i, ptr(o2))
        assert 0 <= o2[0] < 1 and 0 <= o2[1] < 1
        us.append(o2.copy())
    assert np.abs(np.mean(us, 0) - 0.5).max() < 0.2


def test_postprocess_agx():
    o3 = np.zeros(3, np.float32)

    def ref(c):  # float64 transcription of postprocess.slang:13-88 (AGX_LOOK 2)
        m = np.array([[0.842479062253094, 0.0423282422610123, 0.0423756549057051], [0.0784335999999992, 0.878468636469772, 0.0784336],
                      [0.0792237451477643, 0.0791661274605434, 0.879142973793104]])
        v = np.asarray(c, np.float64) @ m
        v = (np.clip(np.log2(np.maximum(v, 1e-300)), -12.47393, 4.026069) + 12.47393) / (4.026069 + 12.47393)
        x2, x4 = v * v, v ** 4
        v = 15.5 * x4 * x2 - 40.14 * x4 * v + 31.96 * x4 - 6.868 * x2 * v + 0.4298 * x2 + 0.1191 * v - 0.00232
        luma = v @ np.array([0.2126, 0.7152, 0.0722])
        v = luma + 1.1 * (np.maximum(v, 0) ** 1.1 - luma)
        mi = np.array([[1.19687900512017, -0.0528968517574562, -0.0529716355144438], [-0.0980208811401368, 1.15190312990417, -0.0980434501171241],
                       [-0.0990297440797205, -0.0989611768448433, 1.15107367264116]])
        return v @ mi

    for c in ([0.18, 0.18, 0.18], [1, 0.5, 0.1], [4, 4, 4], [0.01, 0.2, 0.9], [0, 0, 0]):
        L.orc_agx(ptr(f3(c)), ptr(o3))
        assert np.abs(o3 - ref(c)).max() < 2e-4, (c, o3, ref(c))


def test_next_event_estimation_is_unbiased():
    """RT3_FLAG_NEE (environment importance sampling + shadow rays + balance-heuristic MIS) must converge to the same image as
    plain BRDF sampling of the sky, with less noise."""
    sc = scenes.synthetic_atrium(0.01)
    sc["environment"] = scenes.analytic_sky(128, 64)
    orc = OracleScene(sc)
    W, H = 64, 36
    cam = camera_for(sc, "cam0", W, H)
    imgs = {}
    for flags in (FLAG_ENV, FLAG_ENV | FLAG_NEE):
        acc = np.zeros((H, W, 3))
        for f in range(6):
            p = make_params(cam, W, H, frame=f, bounces=4, samples=32, flags=flags)
            gb, dp, _, _ = orc.pass_gbuffer(p)
            lt, cnt = orc.pass_reference(p, gb, dp)
            acc += lt[..., :3] / 6
        imgs[flags] = acc
        assert (cnt.rays_shadow > 0) == bool(flags & FLAG_NEE)
    a, b = imgs[FLAG_ENV], imgs[FLAG_ENV | FLAG_NEE]
    assert abs(a.mean() - b.mean()) / a.mean() < 0.04
    p1 = make_params(cam, W, H, frame=77, bounces=4, samples=4, flags=FLAG_ENV)
    gb, dp, _, _ = orc.pass_gbuffer(p1)
    n1 = np.sqrt(((orc.pass_reference(p1, gb, dp)[0][..., :3] - b) ** 2).mean())
    n2 = np.sqrt(((orc.pass_reference(make_params(cam, W, H, frame=77, bounces=4, samples=4, flags=FLAG_ENV | FLAG_NEE), gb, dp)[0][..., :3] - b) ** 2).mean())
    assert n2 < n1


def test_env_sampling_tables():
    sc = scenes.cornell_box()
    sc["environment"] = scenes.analytic_sky(64, 32)
    orc = OracleScene(sc)
    # the sampled direction maps back to the texel it was drawn from: uv -> direction -> uv round trip
    o2 = np.zeros(2, np.float32)
    for d in ([0.3, 0.8, 0.5], [-0.7, 0.1, 0.2], [0.0, -0.9, 0.4]):
        d = np.array(d, np.float32) / np.linalg.norm(d)
        L.orc_equirect_uv(ptr(f3(d)), ptr(o2))
        assert 0 <= o2[0] <= 1 and 0 <= o2[1] <= 1


@pytest.mark.parametrize("scene_fn", [lambda: scenes.cornell_box(), lambda: scenes.random_soup(3000, seed=9), lambda: scenes.synthetic_atrium(0.03, environment=False),
                                      lambda: scenes.random_soup(1, n_geometries=1), lambda: scenes.random_soup(3, n_geometries=1)])
def test_bvh4_sse_equals_scalar_bvh2_equals_brute_force(scene_fn):
    """The three query structures of the oracle give the same bits: every triangle (the definition), the scalar BVH2, and the BVH4 with
    4-wide SSE slab and Moeller-Trumbore tests that the passes and the timed CPU arm use.  The SSE triangle test mirrors rt3m::ray_triangle
    operation for operation, so t, u, v are bit-identical, not merely close."""
    sc = scene_fn()
    orc = OracleScene(sc)
    c, a = scenes.incoherent_rays(sc, 6000, seed=17)
    rays = np.concatenate([c, scenes.random_rays(sc, 6000, seed=18)])
    want = orc.trace_closest(rays, brute=True)
    for kw in (dict(), dict(bvh2=True)):
        got = orc.trace_closest(rays, **kw)
        assert (got["prim"] == want["prim"]).all(), kw
        for f in ("t", "u", "v"):
            assert (got[f].view(np.uint32) == want[f].view(np.uint32)).all(), (kw, f)
    wa = orc.trace_any(a, brute=True)
    assert (orc.trace_any(a) == wa).all() and (orc.trace_any(a, bvh2=True) == wa).all()
