"""Pins the restated specular lobe of the reference (shaders/include/brdf.slang:95-322: Schlick Fresnel, height-correlated
Smith masking, GGX NDF, VNDF sampling) and the layered material built on it (RT3_FLAG_SPECULAR).

The reference has no tests for brdf.slang and never calls SpecularBrdf from the path ("parity unpinned"), so the oracle is pinned
by an independent float64 numpy re-derivation of the published formulas (Heitz 2018, JCGT 7(4); Heitz 2014 for the correlated
Smith term), by sample <-> evaluate consistency, by pdf normalisation and by unbiasedness of the one-sample lobe mixture."""
import numpy as np
import pytest

from oracle.oracle_py import OracleScene, f3, lib
from raytracer3_b200 import scenes
from raytracer3_b200.camera import camera_for, make_params
from raytracer3_b200.types import FLAG_ENV, FLAG_NEE, FLAG_SPECULAR, ptr

L = lib()


# ---- independent float64 derivation ------------------------------------------------------------------------------------
def np_ggx_d(a2, c):
    return a2 / (np.pi * (c * c * (a2 - 1.0) + 1.0) ** 2)


def np_g1(c, a2):
    return 2.0 / (1.0 + np.sqrt(1.0 + a2 * (1.0 - c * c) / (c * c)))


def np_g2_correlated(nv, nl, a2):
    return 2.0 * nl * nv / (nl * np.sqrt(nv * nv * (1 - a2) + a2) + nv * np.sqrt(nl * nl * (1 - a2) + a2))


def np_schlick(f0, c):
    return f0 + (1.0 - f0) * max(0.0, 1.0 - c) ** 5


def np_vndf_half_vector(alpha, wo, u1, u2):
    """Heitz 2018, listing 1 (isotropic)."""
    vh = np.array([alpha * wo[0], alpha * wo[1], wo[2]])
    vh /= np.linalg.norm(vh)
    if vh[2] < 0.9999:
        t1 = np.cross([0.0, 0.0, 1.0], vh)
        t1 /= np.linalg.norm(t1)
    else:
        t1 = np.array([1.0, 0.0, 0.0])
    t2 = np.cross(vh, t1)
    r, phi = np.sqrt(u1), 2 * np.pi * u2
    a, b = r * np.cos(phi), r * np.sin(phi)
    s = 0.5 * (1.0 + vh[2])
    b = (1.0 - s) * np.sqrt(max(0.0, 1.0 - a * a)) + s * b
    nh = a * t1 + b * t2 + np.sqrt(max(0.0, 1.0 - a * a - b * b)) * vh
    h = np.array([alpha * nh[0], alpha * nh[1], max(0.0, nh[2])])
    return h / np.linalg.norm(h)


def np_specular(alpha, f0, wo, wi):
    """value_over_pdf, value, pdf (projected solid angle) of brdf.slang:269-314, in float64."""
    a2 = alpha * alpha
    m = (wo + wi) / np.linalg.norm(wo + wi)
    f = np_schlick(f0, float(np.dot(m, wi)))
    g2 = np_g2_correlated(wo[2], wi[2], a2)
    g1 = np_g1(wo[2], a2)
    d = np_ggx_d(a2, m[2])
    pdf_h = g1 * d * max(0.0, float(np.dot(wo, m))) / wo[2]
    pdf = pdf_h / (4.0 * np.dot(wi, m)) / wi[2]
    return f * g2 / g1, f * g2 * d / (4.0 * wo[2] * wi[2]), pdf


def c_sample(alpha, f0, wo, u1, u2):
    out = np.zeros(14, np.float32)
    ok = L.orc_specular_sample(float(alpha), ptr(f3(f0)), ptr(f3(wo)), float(u1), float(u2), ptr(out))
    return ok, out


def c_eval(alpha, f0, wo, wi):
    out = np.zeros(10, np.float32)
    L.orc_specular_evaluate(float(alpha), ptr(f3(f0)), ptr(f3(wo)), ptr(f3(wi)), ptr(out))
    return out


def unit(v):
    v = np.asarray(v, np.float64)
    return v / np.linalg.norm(v)


# ---- tests -------------------------------------------------------------------------------------------------------------
def test_specular_sample_matches_float64_derivation():
    rng = np.random.default_rng(21)
    checked = 0
    for _ in range(3000):
        alpha = float(rng.uniform(0.05, 1.0))
        f0 = rng.uniform(0.02, 1.0, 3)
        wo = unit([rng.normal(), rng.normal(), abs(rng.normal()) + 0.05])
        u1, u2 = rng.random(2)
        ok, out = c_sample(alpha, f0, wo, u1, u2)
        wo32 = f3(wo).astype(np.float64)
        h = np_vndf_half_vector(np.float32(alpha).astype(np.float64), wo32, np.float64(np.float32(u1)), np.float64(np.float32(u2)))
        wi = 2.0 * np.dot(wo32, h) * h - wo32
        if h[2] <= 1e-4 or wi[2] <= 1e-3:
            continue                                         # invalid or grazing samples are covered by the validity test below
        assert ok == 1
        assert np.abs(out[10:13] - wi).max() < 2e-4, (out[10:13], wi)
        vop, val, pdf = np_specular(np.float64(np.float32(alpha)), f3(f0).astype(np.float64), wo32, wi)
        assert np.allclose(out[0:3], vop, rtol=2e-3, atol=1e-5)
        assert np.allclose(out[3:6], val, rtol=5e-3, atol=1e-5)
        assert out[6] == pytest.approx(pdf, rel=5e-3, abs=1e-5)
        assert np.allclose(out[7:10], 1.0 - np_schlick(f3(f0).astype(np.float64), float(np.dot(h, wi))), atol=2e-5)
        assert out[13] == np.float32(alpha)
        checked += 1
    assert checked > 2000


def test_specular_evaluate_agrees_with_sample():
    rng = np.random.default_rng(22)
    for _ in range(1500):
        alpha = float(rng.uniform(0.05, 1.0))
        f0 = rng.uniform(0.02, 1.0, 3)
        wo = unit([rng.normal(), rng.normal(), abs(rng.normal()) + 0.1])
        ok, s = c_sample(alpha, f0, wo, *rng.random(2))
        if not ok or s[12] < 1e-3:
            continue
        e = c_eval(alpha, f0, wo, s[10:13])
        assert np.allclose(e[0:3], s[0:3], rtol=2e-3, atol=1e-5)
        assert np.allclose(e[3:6], s[3:6], rtol=5e-3, atol=1e-5)
        assert e[6] == pytest.approx(s[6], rel=5e-3, abs=1e-5)
    # below the horizon: BrdfValue::invalid() (brdf.slang:270-272)
    e = c_eval(0.3, [1, 1, 1], unit([0.2, 0.1, 0.9]), unit([0.1, 0.3, -0.5]))
    assert (e == 0).all()
    e = c_eval(0.3, [1, 1, 1], unit([0.2, 0.1, -0.9]), unit([0.1, 0.3, 0.5]))
    assert (e == 0).all()


def test_specular_sample_validity_rules():
    # wo below BRDF_SAMPLING_MIN_COS -> BrdfSample::invalid(): wi = (0, 0, -1), everything else 0 (brdf.slang:36-44, 226-228)
    ok, s = c_sample(0.4, [0.9, 0.9, 0.9], unit([1.0, 0.0, 1e-7]), 0.3, 0.6)
    assert ok == 0 and (s[10:13] == [0, 0, -1]).all() and (s[0:7] == 0).all()
    # a mirror-like lobe reflects about the normal
    wo = unit([0.3, -0.2, 0.8])
    ok, s = c_sample(0.001, [1, 1, 1], wo, 0.5, 0.25)
    assert ok == 1 and np.abs(s[10:13] - [-wo[0], -wo[1], wo[2]]).max() < 5e-3
    assert np.allclose(s[0:3], 1.0, atol=2e-3)                # F = 1, G2 / G1 -> 1 for a smooth surface


def test_vndf_pdf_is_normalised_and_weights_conserve_energy():
    """Integral of the solid-angle pdf over the hemisphere = 1 - (mass reflected below the horizon); E[value_over_pdf] <= 1 for f0 = 1."""
    rng = np.random.default_rng(23)
    n = 40000
    for alpha, wo in ((0.2, unit([0.3, 0.1, 0.9])), (0.6, unit([0.7, -0.2, 0.5]))):
        # uniform hemisphere directions, pdf 1 / (2 pi)
        z = rng.random(n)
        ph = 2 * np.pi * rng.random(n)
        r = np.sqrt(1 - z * z)
        wis = np.stack([r * np.cos(ph), r * np.sin(ph), z], 1)
        acc = 0.0
        for wi in wis[:8000]:
            acc += c_eval(alpha, [1, 1, 1], wo, wi)[6] * wi[2]      # projected -> solid angle
        integral = acc / 8000 * 2 * np.pi
        mean_w, valid = 0.0, 0
        us = rng.random((8000, 2))
        for u1, u2 in us:
            ok, s = c_sample(alpha, [1, 1, 1], wo, u1, u2)
            if ok:
                mean_w += s[0]
                valid += 1
        assert integral == pytest.approx(valid / len(us), abs=0.05)
        assert 0.5 < mean_w / len(us) <= 1.0 + 1e-6


def test_dominant_direction():
    n, v = unit([0, 0, 1]), unit([0.6, 0.0, 0.8])
    o = np.zeros(3, np.float32)
    L.orc_specular_dominant_direction(ptr(f3(n)), ptr(f3(v)), 0.0, ptr(o))
    assert np.abs(o - [-0.6, 0.0, 0.8]).max() < 1e-6           # smooth: the mirror direction
    L.orc_specular_dominant_direction(ptr(f3(n)), ptr(f3(v)), 1.0, ptr(o))
    assert np.abs(o - n).max() < 1e-6                           # fully rough: the normal (brdf.slang:318-322)


def _layered(albedo, normal, rough, metal, ray_d, u1, u2):
    out = np.zeros(8, np.float32)
    code = L.orc_layered_sample(ptr(f3(albedo)), ptr(f3(normal)), float(rough), float(metal), ptr(f3(ray_d)), float(u1), float(u2), ptr(out))
    return code, out


def test_layered_material_is_unbiased_and_falls_back_to_diffuse():
    rng = np.random.default_rng(24)
    normal = unit([0.2, 0.9, -0.3])
    ray_d = unit(-normal + 0.6 * unit([0.5, 0.1, 0.4]))          # arrives from the front side
    for albedo, rough, metal in (([0.8, 0.5, 0.3], 0.5, 0.0), ([0.9, 0.7, 0.2], 0.3, 1.0), ([0.6, 0.6, 0.6], 0.8, 0.5)):
        # Monte-Carlo mean of the path weight == integral of value * cos over the hemisphere (quadrature with uniform directions)
        us = rng.random((12000, 2))
        acc, n_spec, n_valid = np.zeros(3), 0, 0
        for u1, u2 in us:
            code, o = _layered(albedo, normal, rough, metal, ray_d, u1, u2)
            if code & 1:
                n_valid += 1
                acc += o[3:6]
                assert np.dot(o[0:3], normal) > 0
            n_spec += (code >> 1) & 1
        mc = acc / len(us)
        p_spec = o[7]
        assert n_spec / len(us) == pytest.approx(p_spec, abs=0.02)
        z = rng.random(12000)
        ph = 2 * np.pi * rng.random(12000)
        r = np.sqrt(1 - z * z)
        t = np.cross(normal, [1.0, 0.0, 0.0]); t /= np.linalg.norm(t)
        b = np.cross(normal, t)
        quad, o4 = np.zeros(3), np.zeros(4, np.float32)
        pdf_int = 0.0
        for k in range(len(z)):
            wi = r[k] * np.cos(ph[k]) * t + r[k] * np.sin(ph[k]) * b + z[k] * normal
            L.orc_layered_evaluate(ptr(f3(albedo)), ptr(f3(normal)), float(rough), float(metal), ptr(f3(ray_d)), ptr(f3(wi)), ptr(o4))
            quad += o4[0:3] * z[k]
            pdf_int += o4[3]
        quad *= 2 * np.pi / len(z)
        assert np.allclose(mc, quad, rtol=0.08, atol=0.01), (mc, quad)
        # the mixture pdf integrates to 1 minus the VNDF mass reflected below the horizon (those samples end the path)
        assert pdf_int * 2 * np.pi / len(z) == pytest.approx(n_valid / len(us), abs=0.08)
    # seen from behind (wo.z <= 0): p_spec = 0 and the sample is exactly the reference's diffuse lobe with weight albedo * (1 - metalness)
    back = unit(normal + 0.3 * unit([0.5, 0.1, 0.4]))
    code, o = _layered([0.8, 0.5, 0.3], normal, 0.4, 0.25, back, 0.37, 0.61)
    assert code == 1 and o[7] == 0.0
    d3, o9 = np.zeros(3, np.float32), np.zeros(9, np.float32)
    L.orc_diffuse_sample(0.37, 0.61, ptr(d3))
    L.orc_onb(ptr(f3(normal)), ptr(o9))
    want = o9[0:3] * d3[0] + o9[3:6] * d3[1] + o9[6:9] * d3[2]
    assert np.abs(o[0:3] - want).max() < 1e-6
    assert np.allclose(o[3:6], np.array([0.8, 0.5, 0.3], np.float32) * np.float32(0.75), rtol=1e-6)


def _metal_cornell():
    return scenes.metal_cornell_box()


def test_specular_flag_renders_and_nee_stays_unbiased():
    sc = _metal_cornell()
    sc["environment"] = scenes.analytic_sky(64, 32)
    orc = OracleScene(sc)
    W, H = 48, 32
    cam = camera_for(sc, "cam0", W, H)
    p0 = make_params(cam, W, H, frame=2, bounces=4, samples=8, flags=0)
    gb, dp, _, _ = orc.pass_gbuffer(p0)
    base, _ = orc.pass_reference(p0, gb, dp)
    spec, cnt = orc.pass_reference(make_params(cam, W, H, frame=2, bounces=4, samples=8, flags=FLAG_SPECULAR), gb, dp)
    assert np.isfinite(spec).all() and (spec[..., :3] >= 0).all()
    assert np.abs(spec - base).max() > 1e-3                    # the material model changes the image
    assert 0.3 < spec[..., :3].mean() / base[..., :3].mean() < 3.0
    imgs = {}
    for flags in (FLAG_ENV | FLAG_SPECULAR, FLAG_ENV | FLAG_NEE | FLAG_SPECULAR):
        acc = np.zeros((H, W, 3))
        for f in range(4):
            lt, cnt = orc.pass_reference(make_params(cam, W, H, frame=f, bounces=3, samples=64, flags=flags), gb, dp)
            acc += lt[..., :3] / 4
        imgs[flags] = acc
        assert (cnt.rays_shadow > 0) == bool(flags & FLAG_NEE)
    a, b = imgs[FLAG_ENV | FLAG_SPECULAR], imgs[FLAG_ENV | FLAG_NEE | FLAG_SPECULAR]
    assert abs(a.mean() - b.mean()) / a.mean() < 0.05
