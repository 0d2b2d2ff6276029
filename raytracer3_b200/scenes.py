"""Deterministic stand-in scenes and ray sets (numpy, closed form, seeded).

`resources/sponza_scene.glb` and `resources/skybox2.exr` are absent from the reference mount
(/root/reference/.MISSING_LARGE_BLOBS:2-3), so every performance number is taken on the synthetic
atrium below and tagged scene=synthetic_atrium.  The Cornell box mirrors the
8-object scene of /root/reference/resources/refrence.png with the material palette recovered from
/root/reference/imported_assets/Default/box.glb and is used for correctness only.

A scene is a plain dict of numpy arrays in the layouts of raytracer3_b200.types:
    vertices (VERTEX), indices (u32), geometries (GEOMETRY), instances [(first_geo, count, xform12)],
    textures {slot: HxWx4 u8}, environment (HxWx3 f32) or None, cameras {name: Camera kwargs}
"""
import numpy as np

from .types import GEOMETRY, IDENTITY_3X4, RAY, VERTEX

# ----------------------------------------------------------------------------- mesh primitives


class MeshBuilder:
    """Accumulates geometries (one per material) into the global vertex / index buffers."""

    def __init__(self):
        self.v = []
        self.i = []
        self.g = []
        self.nv = 0
        self.ni = 0

    def add(self, pos, nrm, uv, tri, base_color=(0.8, 0.8, 0.8), emission=(0, 0, 0), tex=-1, roughness=1.0, metallic=0.0,
            xform=None):
        pos = np.asarray(pos, np.float32).reshape(-1, 3)
        nrm = np.asarray(nrm, np.float32).reshape(-1, 3)
        uv = np.asarray(uv, np.float32).reshape(-1, 2)
        tri = np.asarray(tri, np.uint32).reshape(-1)
        v = np.zeros(len(pos), VERTEX)
        v["p"], v["n"], v["t"] = pos, nrm, uv
        g = np.zeros(1, GEOMETRY)
        g["base_color"] = (*base_color, 1.0)
        g["base_color_tex"] = tex
        g["metallic"] = metallic
        g["roughness"] = roughness
        g["index_offset"] = self.ni
        g["vertex_offset"] = self.nv
        g["emission"] = (*emission, 0.0)
        g["index_count"] = len(tri)
        g["xform"] = IDENTITY_3X4 if xform is None else np.asarray(xform, np.float32)
        self.v.append(v)
        self.i.append(tri)
        self.g.append(g)
        self.nv += len(v)
        self.ni += len(tri)
        return len(self.g) - 1

    def finish(self):
        return (np.concatenate(self.v) if self.v else np.zeros(0, VERTEX),
                np.concatenate(self.i) if self.i else np.zeros(0, np.uint32),
                np.concatenate(self.g) if self.g else np.zeros(0, GEOMETRY))


def merge(parts):
    """Concatenate (pos, nrm, uv, tri) tuples into one."""
    P, N, U, T = [], [], [], []
    base = 0
    for p, n, u, t in parts:
        P.append(p); N.append(n); U.append(u); T.append(np.asarray(t, np.int64) + base)
        base += len(p)
    return (np.concatenate(P).astype(np.float32), np.concatenate(N).astype(np.float32), np.concatenate(U).astype(np.float32),
            np.concatenate(T).astype(np.uint32))


def quad(o, du, dv, nu=1, nv=1, normal=None, uv_scale=(1.0, 1.0), disp=None):
    """Grid of nu x nv cells spanning o + s*du + t*dv; normal defaults to normalize(du x dv)."""
    o, du, dv = (np.asarray(a, np.float64) for a in (o, du, dv))
    s, t = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    s, t = s.ravel(), t.ravel()
    pos = o + s[:, None] * du + t[:, None] * dv
    n = np.cross(du, dv)
    n = n / np.linalg.norm(n) if normal is None else np.asarray(normal, np.float64)
    nrm = np.broadcast_to(n, pos.shape).copy()
    if disp is not None:
        pos, nrm = disp(pos, nrm, s, t)
    uv = np.stack([s * uv_scale[0], t * uv_scale[1]], 1)
    ix, iy = np.meshgrid(np.arange(nu), np.arange(nv), indexing="xy")
    a = (iy * (nu + 1) + ix).ravel()
    tri = np.stack([a, a + 1, a + nu + 2, a, a + nu + 2, a + nu + 1], 1).reshape(-1, 3)
    return pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), tri.astype(np.uint32)


def box(lo, hi, inward=False, faces="xXyYzZ", sub=1):
    """Axis-aligned box, per-face normals.  faces: x/X = -x/+x side etc.  inward flips normals + winding."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    e = hi - lo
    parts = []
    spec = {
        "x": (lo, (0, 0, e[2]), (0, e[1], 0)),
        "X": ((hi[0], lo[1], lo[2]), (0, e[1], 0), (0, 0, e[2])),
        "y": (lo, (e[0], 0, 0), (0, 0, e[2])),
        "Y": ((lo[0], hi[1], lo[2]), (0, 0, e[2]), (e[0], 0, 0)),
        "z": (lo, (0, e[1], 0), (e[0], 0, 0)),
        "Z": ((lo[0], lo[1], hi[2]), (e[0], 0, 0), (0, e[1], 0)),
    }
    for f in faces:
        o, du, dv = spec[f]
        if inward:
            du, dv = dv, du
        parts.append(quad(o, du, dv, sub, sub))
    return merge(parts)


def lathe(profile_r, profile_y, center, seg, closed_caps=False):
    """Surface of revolution about +Y through `center`; profile given bottom to top."""
    r = np.asarray(profile_r, np.float64)
    y = np.asarray(profile_y, np.float64)
    m = len(r)
    th = np.linspace(0, 2 * np.pi, seg + 1)
    ct, st = np.cos(th), np.sin(th)
    pos = np.stack([np.outer(r, ct), np.repeat(y[:, None], seg + 1, 1), np.outer(r, st)], -1).reshape(-1, 3) + np.asarray(center)
    dr, dy = np.gradient(r), np.gradient(y)
    nl = np.sqrt(dr * dr + dy * dy) + 1e-12
    nr, ny = dy / nl, -dr / nl
    nrm = np.stack([np.outer(nr, ct), np.repeat(ny[:, None], seg + 1, 1), np.outer(nr, st)], -1).reshape(-1, 3)
    uv = np.stack(np.meshgrid(np.linspace(0, 1, seg + 1), np.linspace(0, 1, m), indexing="xy"), -1).reshape(-1, 2)
    ix, iy = np.meshgrid(np.arange(seg), np.arange(m - 1), indexing="xy")
    a = (iy * (seg + 1) + ix).ravel()
    tri = np.stack([a, a + seg + 1, a + seg + 2, a, a + seg + 2, a + 1], 1).reshape(-1, 3)
    return pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), tri.astype(np.uint32)


def arch(x0, x1, y_spring, z0, z1, seg):
    """Semicircular arch soffit between x0 and x1 springing at y_spring, extruded from z0 to z1,
    plus the two spandrel faces (flat wall pieces between the arc and the rectangle above it)."""
    cx, r = 0.5 * (x0 + x1), 0.5 * (x1 - x0)
    th = np.linspace(np.pi, 0, seg + 1)
    ax, ay = cx + r * np.cos(th), y_spring + r * np.sin(th)
    top = y_spring + r * 1.15
    parts = []
    # soffit (normals point toward the arc centre: visible from below)
    pos = np.concatenate([np.stack([ax, ay, np.full_like(ax, z0)], 1), np.stack([ax, ay, np.full_like(ax, z1)], 1)])
    n = np.stack([-(np.cos(th)), -(np.sin(th)), np.zeros_like(th)], 1)
    nrm = np.concatenate([n, n])
    uv = np.concatenate([np.stack([np.linspace(0, 1, seg + 1), np.zeros(seg + 1)], 1), np.stack([np.linspace(0, 1, seg + 1), np.ones(seg + 1)], 1)])
    a = np.arange(seg)
    tri = np.stack([a, a + 1, a + seg + 2, a, a + seg + 2, a + seg + 1], 1).reshape(-1, 3)
    parts.append((pos, nrm, uv, tri))
    # spandrels at z0 (facing -z) and z1 (facing +z)
    for z, nz in ((z0, -1.0), (z1, 1.0)):
        p = np.concatenate([np.stack([ax, ay, np.full_like(ax, z)], 1), np.stack([ax, np.full_like(ax, top), np.full_like(ax, z)], 1)])
        nn = np.broadcast_to(np.array([0, 0, nz]), p.shape).copy()
        u = np.concatenate([np.stack([np.linspace(0, 1, seg + 1), np.zeros(seg + 1)], 1), np.stack([np.linspace(0, 1, seg + 1), np.ones(seg + 1)], 1)])
        t = np.stack([a, a + seg + 1, a + seg + 2, a, a + seg + 2, a + 1], 1).reshape(-1, 3)
        if nz > 0:
            t = t[:, ::-1]
        parts.append((p, nn, u, t))
    return merge(parts)


def translate(mesh, d):
    p, n, u, t = mesh
    return (p + np.asarray(d, np.float32), n, u, t)


# ----------------------------------------------------------------------------- procedural images


def checker_texture(size=64, cells=8, c0=(235, 225, 200), c1=(120, 60, 50), seed=0):
    """sRGB RGBA8 checker with a little per-texel noise (exercises the bilinear / repeat sampler)."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size]
    m = ((x * cells // size) + (y * cells // size)) & 1
    img = np.where(m[..., None] == 0, np.array(c0), np.array(c1)).astype(np.int32)
    img = np.clip(img + rng.integers(-12, 13, (size, size, 1)), 0, 255)
    out = np.empty((size, size, 4), np.uint8)
    out[..., :3] = img
    out[..., 3] = 255
    return out


def analytic_sky(w=2048, h=1024, sun_dir=(0.35, 0.85, 0.4), sun_radiance=60.0):
    """Stand-in for resources/skybox2.exr: gradient sky + soft sun disc, equirect RGB float32 (closed form).
    Texel (i, j) covers uv = ((i+.5)/w, (j+.5)/h) of shaders/include/math.slang:6-12."""
    u = (np.arange(w) + 0.5) / w
    v = (np.arange(h) + 0.5) / h
    phi = (u - 0.5) * 2 * np.pi          # atan2(z, x)
    el = (0.5 - v) * np.pi               # asin(y)
    d = np.stack([np.outer(np.cos(el), np.cos(phi)), np.repeat(np.sin(el)[:, None], w, 1), np.outer(np.cos(el), np.sin(phi))], -1)
    s = np.asarray(sun_dir, np.float64)
    s = s / np.linalg.norm(s)
    up = np.clip(d[..., 1], 0, 1)
    horizon = np.array([0.85, 0.9, 1.0])
    zenith = np.array([0.25, 0.45, 0.95])
    sky = horizon + (zenith - horizon) * (up ** 0.5)[..., None]
    ground = np.array([0.18, 0.16, 0.14])
    img = np.where((d[..., 1] >= 0)[..., None], sky * 1.2, ground)
    cosang = np.clip((d * s).sum(-1), -1, 1)
    img = img + (sun_radiance * np.exp((cosang - 1.0) * 600.0))[..., None] * np.array([1.0, 0.93, 0.8])
    return np.ascontiguousarray(img, np.float32)


def blue_noise_tile(size=256, seed=7, iters=12):
    """RGBA8 tile with a high-pass ("blue") spectrum per channel and a flat histogram.
    Stand-in for resources/bluenoise.png (256x256 RGBA8, four decorrelated channels): iterated
    Gaussian high-pass + rank equalisation of white noise."""
    rng = np.random.default_rng(seed)
    fy = np.fft.fftfreq(size)[:, None]
    fx = np.fft.fftfreq(size)[None, :]
    lowpass = np.exp(-(fx * fx + fy * fy) / (2 * (1.5 / size * 8) ** 2))
    out = np.empty((size, size, 4), np.uint8)
    for c in range(4):
        a = rng.random((size, size))
        for _ in range(iters):
            blur = np.real(np.fft.ifft2(np.fft.fft2(a) * lowpass))
            a = a - blur
            rank = np.argsort(np.argsort(a.ravel(), kind="stable"), kind="stable")
            a = (rank.reshape(size, size) + 0.5) / (size * size)
        out[..., c] = np.minimum((a * 256).astype(np.int32), 255)
    return out


def reference_blue_noise(directory=None):
    """The reference's own sampler input: resources/bluenoise.png (256x256 RGBA8, /root/reference/resources/bluenoise.png; committed
    under <repo>/resources as a data fixture, byte-identical).  Looked up in `directory`, $RT3_ASSET_DIR, then <repo>/resources.
    Returns (tile, source); falls back to the generated stand-in `blue_noise_tile()` (source says so) only if the file is nowhere."""
    import os
    from . import assets
    here = os.path.dirname(os.path.abspath(__file__))
    for d in (directory, os.environ.get("RT3_ASSET_DIR"), os.path.join(here, "..", "resources")):
        if d and os.path.isfile(os.path.join(d, "bluenoise.png")):
            with open(os.path.join(d, "bluenoise.png"), "rb") as f:
                tile = assets.decode_image(f.read())
            assert tile.shape == (256, 256, 4) and tile.dtype == np.uint8
            return np.ascontiguousarray(tile), "resources/bluenoise.png (the reference's tile)"
    return blue_noise_tile(), "generated stand-in (scenes.blue_noise_tile)"


# ----------------------------------------------------------------------------- scenes


def _scene(mb, cameras, textures=None, environment=None, name=""):
    v, i, g = mb.finish()
    return {"name": name, "vertices": v, "indices": i, "geometries": g, "instances": [], "textures": textures or {},
            "environment": environment, "cameras": cameras}


def cornell_box(light_emission=1.4):
    """36 triangles: 5 walls (10) + ceiling light (2) + tall box (12) + short box (12)."""
    mb = MeshBuilder()
    grey, red, blue, green = (0.8, 0.8, 0.8), (0.8, 0.0, 0.006), (0.0, 0.0, 0.8), (0.009, 0.8, 0.0)
    mb.add(*quad((-1, -1, -1), (0, 0, 2), (2, 0, 0)), base_color=green)             # floor, normal +y
    mb.add(*quad((-1, 1, -1), (2, 0, 0), (0, 0, 2)), base_color=grey)               # ceiling, normal -y
    mb.add(*quad((-1, -1, 1), (0, 2, 0), (2, 0, 0)), base_color=grey)               # back wall z=+1, normal -z
    mb.add(*quad((-1, -1, -1), (0, 2, 0), (0, 0, 2)), base_color=blue)              # wall x=-1 (screen right), normal +x
    mb.add(*quad((1, -1, -1), (0, 0, 2), (0, 2, 0)), base_color=red)                # wall x=+1 (screen left: red, as in refrence.png), normal -x
    mb.add(*quad((-0.35, 0.998, -0.3), (0.7, 0, 0), (0, 0, 0.6)), base_color=grey, emission=(light_emission,) * 3)
    c, s = np.cos(0.3), np.sin(0.3)
    rot = np.array([c, 0, s, 0.35, 0, 1, 0, 0, -s, 0, c, 0.35], np.float32)
    mb.add(*box((-0.3, -1, -0.3), (0.3, 0.2, 0.3)), base_color=(0.5, 0.5, 0.5), xform=rot)      # tall box (geometry transform), near the red wall as in refrence.png
    c, s = np.cos(-0.3), np.sin(-0.3)
    rot2 = np.array([c, 0, s, -0.4, 0, 1, 0, 0, -s, 0, c, -0.3], np.float32)
    mb.add(*box((-0.28, -1, -0.28), (0.28, -0.44, 0.28)), base_color=(0.02, 0.02, 0.02), roughness=0.5, xform=rot2)
    # cam_ref: the framing of /root/reference/resources/refrence.png (front opening = 79 % of the image height, back wall = 75 % of the opening)
    cams = {"cam0": dict(position=(0, 0, -3.6), direction=(0, 0, 1), fov_deg=40.0),
            "cam_ref": dict(position=(0, 0, -7.0), direction=(0, 0, 1), fov_deg=23.8)}
    return _scene(mb, cams, name="cornell_box")


def metal_cornell_box():
    """Cornell box with glossy / metallic materials: exercises RT3_FLAG_SPECULAR (brdf.slang:95-318)."""
    sc = cornell_box()
    g = sc["geometries"]
    g["metallic"][:] = 0.0
    g["roughness"][:] = 0.6
    g["metallic"][-2:] = 1.0          # the two boxes: polished metal
    g["roughness"][-2:] = 0.15
    g["roughness"][0] = 0.25          # first wall: glossy dielectric
    sc["name"] = "metal_cornell"
    return sc


def furnace_box(albedo=0.5, emission=0.25 / 12.0):
    """Closed cube, uniform albedo rho and emission e (x12 in hit_logic.slang:36): every path vertex sees
    the same surface, so radiance after B bounces is exactly e*12 * sum_{k<B} rho^k."""
    mb = MeshBuilder()
    mb.add(*box((-1, -1, -1), (1, 1, 1), inward=True, sub=2), base_color=(albedo,) * 3, emission=(emission,) * 3)
    cams = {"cam0": dict(position=(0.1, 0.05, -0.2), direction=(0.2, 0.1, 1), fov_deg=65.0)}
    return _scene(mb, cams, name="furnace_box")


def random_soup(n_tris=2000, seed=1, extent=10.0, size=0.6, n_geometries=5):
    """Random triangle soup (degenerate-ish slivers included) for BVH == brute-force tests."""
    rng = np.random.default_rng(seed)
    mb = MeshBuilder()
    per = np.array_split(np.arange(n_tris), n_geometries)
    for k, ids in enumerate(per):
        m = len(ids)
        c = rng.uniform(-extent, extent, (m, 1, 3))
        p = (c + rng.normal(0, size, (m, 3, 3)) * rng.uniform(0.05, 1.0, (m, 1, 3))).reshape(-1, 3)
        e1, e2 = p[1::3] - p[0::3], p[2::3] - p[0::3]
        n = np.cross(e1, e2)
        n /= np.linalg.norm(n, axis=1, keepdims=True) + 1e-20
        nrm = np.repeat(n, 3, 0)
        uv = rng.random((3 * m, 2))
        mb.add(p, nrm, uv, np.arange(3 * m), base_color=tuple(rng.uniform(0.2, 0.9, 3)), emission=(0.3, 0.2, 0.1) if k == 0 else (0, 0, 0),
               tex=0 if k == 1 else -1)
    cams = {"cam0": dict(position=(0, 0, -2.5 * extent), direction=(0, 0, 1), fov_deg=65.0)}
    return _scene(mb, cams, textures={0: checker_texture(32, 4)}, name="random_soup")


def synthetic_atrium(detail=1.0, environment=True):
    """Sponza-like two-storey colonnaded hall, open to the sky.  detail=1.0 gives ~262k triangles
    (the triangle count SURVEY.md section 8a infers for sponza_scene.glb); AABB ~ 30 x 13 x 18."""
    d = float(detail)
    seg_col = max(6, int(round(48 * np.sqrt(d))))
    ring_col = max(2, int(round(20 * np.sqrt(d))))
    seg_arch = max(4, int(round(40 * d)))
    cloth_n = max(4, int(round(72 * np.sqrt(d))))
    floor_n = max(2, int(round(96 * np.sqrt(d))))
    vase_seg = max(6, int(round(40 * np.sqrt(d))))
    mb = MeshBuilder()
    L, W, H = 30.0, 18.0, 13.0          # x: long axis, z: width, y: height
    x0, x1, z0, z1 = -L / 2, L / 2, -W / 2, W / 2
    aisle = 3.5                          # colonnade depth on each long side
    storey = 5.5
    stone, stone2, brick = (0.72, 0.68, 0.6), (0.6, 0.57, 0.52), (0.55, 0.35, 0.28)
    # floor (textured) and the outer shell (inward-facing walls, open top)
    mb.add(*quad((x0, 0, z0), (0, 0, W), (L, 0, 0), floor_n, int(floor_n * 1.6), uv_scale=(6, 10)), base_color=(0.9, 0.9, 0.9), tex=0)
    mb.add(*box((x0, 0, z0), (x1, H, z1), inward=True, faces="xXzZ", sub=max(1, int(8 * d))), base_color=brick)
    # upper gallery floors (slabs) along both long sides and both ends
    slabs = [((x0, storey, z0), (x1, storey + 0.4, z0 + aisle)), ((x0, storey, z1 - aisle), (x1, storey + 0.4, z1)),
             ((x0, storey, z0 + aisle), (x0 + aisle, storey + 0.4, z1 - aisle)), ((x1 - aisle, storey, z0 + aisle), (x1, storey + 0.4, z1 - aisle))]
    mb.add(*merge([box(a, b, sub=max(1, int(6 * d))) for a, b in slabs]), base_color=stone2)
    # cornice / roof rim
    rim = [((x0, 2 * storey, z0), (x1, 2 * storey + 0.5, z0 + aisle)), ((x0, 2 * storey, z1 - aisle), (x1, 2 * storey + 0.5, z1))]
    mb.add(*merge([box(a, b, sub=max(1, int(4 * d))) for a, b in rim]), base_color=stone)
    # columns: 2 storeys x 2 sides x n_col, with entasis profile + square plinths
    n_col = 10
    xs = np.linspace(x0 + aisle, x1 - aisle, n_col)
    yy = np.linspace(0, 1, ring_col + 1)
    prof = 0.42 * (1 - 0.18 * yy ** 2) * (1 + 0.25 * np.exp(-((yy - 0.0) / 0.04) ** 2) + 0.3 * np.exp(-((yy - 1.0) / 0.05) ** 2))
    cols, plinths = [], []
    for level in range(2):
        yb = level * (storey + 0.4)
        hcol = storey - 0.3 - (0.0 if level == 0 else 0.5)
        for zc in (z0 + aisle, z1 - aisle):
            for xc in xs:
                cols.append(lathe(prof * (1.0 if level == 0 else 0.8), yb + 0.3 + yy * (hcol - 0.3), (xc, 0, zc), seg_col))
                plinths.append(box((xc - 0.55, yb, zc - 0.55), (xc + 0.55, yb + 0.3, zc + 0.55)))
    mb.add(*merge(cols[: len(cols) // 2]), base_color=stone)
    mb.add(*merge(cols[len(cols) // 2:]), base_color=(0.78, 0.74, 0.68))
    mb.add(*merge(plinths), base_color=stone2)
    # arches between neighbouring columns on both storeys, both sides
    arches = []
    for level in range(2):
        ys = level * (storey + 0.4) + (storey - 0.3) - 0.5 * (xs[1] - xs[0]) * 0.0 - 1.6
        for zc in (z0 + aisle, z1 - aisle):
            for a, b in zip(xs[:-1], xs[1:]):
                arches.append(arch(a + 0.4, b - 0.4, ys, zc - 0.35, zc + 0.35, seg_arch))
    mb.add(*merge(arches[: len(arches) // 2]), base_color=(0.66, 0.62, 0.55))
    mb.add(*merge(arches[len(arches) // 2:]), base_color=(0.7, 0.64, 0.56))
    # hanging cloth sheets (sine-displaced grids, double-sided by a second back-facing copy)
    def cloth(xc, zc, width, height, phase):
        def disp(pos, nrm, s, t):
            amp = 0.25 * (0.3 + t)
            off = amp * np.sin(6.0 * s + phase) * np.cos(2.0 * t + phase)
            pos = pos + nrm * off[:, None]
            ds = amp * 6.0 * np.cos(6.0 * s + phase) * np.cos(2.0 * t + phase) / width
            n2 = nrm - ds[:, None] * np.array([1.0, 0, 0])
            n2 /= np.linalg.norm(n2, axis=1, keepdims=True)
            return pos, n2
        front = quad((xc - width / 2, 2 * storey - 0.5, zc), (width, 0, 0), (0, -height, 0), cloth_n, cloth_n, disp=disp, uv_scale=(3, 3))
        p, n, u, t = front
        back = (p - n * 0.01, -n, u, t[:, ::-1])
        return merge([front, back])
    cloth_cols = [(0.75, 0.1, 0.1), (0.1, 0.45, 0.15), (0.12, 0.2, 0.6), (0.8, 0.65, 0.15)]
    for k in range(4):
        sheets = [cloth(x0 + 5 + 6.5 * k, zc, 3.2, 6.0, 1.3 * k + j) for j, zc in enumerate((-2.2, 2.2))]
        mb.add(*merge(sheets), base_color=cloth_cols[k], tex=1 if k % 2 == 0 else -1)
    # vases (lathe surfaces) along the centre line + spheres on pedestals
    py = np.linspace(0, 1, max(4, int(36 * np.sqrt(d))) + 1)
    pr = 0.18 + 0.38 * np.sin(np.pi * np.clip(py * 1.1, 0, 1)) ** 1.5 + 0.08 * (py > 0.9)
    vases = [lathe(pr, 0.0 + 1.4 * py, (x, 0, z), vase_seg) for x in np.linspace(x0 + 6, x1 - 6, 5) for z in (-0.9, 0.9)]
    mb.add(*merge(vases[:5]), base_color=(0.25, 0.5, 0.45))
    mb.add(*merge(vases[5:]), base_color=(0.6, 0.3, 0.2))
    th = np.linspace(0, np.pi, max(4, int(28 * np.sqrt(d))) + 1)
    balls = [lathe(0.45 * np.sin(th), storey + 0.4 + 0.45 - 0.45 * np.cos(th), (x, 0, z), vase_seg)
             for x in xs[::3] for z in (z0 + aisle - 1.2, z1 - aisle + 1.2)]
    mb.add(*merge(balls), base_color=(0.8, 0.8, 0.82), roughness=0.3, metallic=0.0)
    # pedestals + benches (boxes) on the ground floor
    peds = [box((x - 0.5, 0, z - 0.5), (x + 0.5, 0.9, z + 0.5)) for x in np.linspace(x0 + 4, x1 - 4, 6) for z in (z0 + 1.2, z1 - 1.2)]
    mb.add(*merge(peds), base_color=(0.35, 0.33, 0.3))
    # lanterns: small emissive boxes under the galleries (so reference-faithful mode is lit at all)
    lamps = [box((x - 0.15, storey - 0.6, z - 0.15), (x + 0.15, storey - 0.3, z + 0.15)) for x in xs[1::2] for z in (z0 + aisle * 0.5, z1 - aisle * 0.5)]
    mb.add(*merge(lamps), base_color=(0.9, 0.8, 0.6), emission=(2.0, 1.5, 0.8))
    tex = {0: checker_texture(128, 8), 1: checker_texture(64, 16, (250, 240, 230), (200, 170, 150), seed=3)}
    cams = {
        "cam0": dict(position=(-0.35 * L, 0.25 * H, 0.0), direction=(1, -0.05, 0.08), fov_deg=65.0),
        "cam1": dict(position=(0.3 * L, storey + 1.8, 0.0), direction=(-1, -0.25, 0.3), fov_deg=65.0),
    }
    return _scene(mb, cams, textures=tex, environment=analytic_sky() if environment else None, name="synthetic_atrium")


def instanced_grid(scene, n=8, spacing=1.05):
    """cfg5: n x n instances of `scene`, translation step = spacing x AABB extent in X and Z."""
    p = scene["vertices"]["p"]
    ext = p.max(0) - p.min(0)
    out = dict(scene)
    inst = []
    ng = len(scene["geometries"])
    for iz in range(n):
        for ix in range(n):
            m = IDENTITY_3X4.copy()
            m[3] = ix * spacing * ext[0]
            m[11] = iz * spacing * ext[2]
            inst.append((0, ng, m))
    out["instances"] = inst
    out["name"] = f"{scene['name']}_grid{n}x{n}"
    c = p.min(0) + np.array([n * spacing * ext[0], 0, n * spacing * ext[2]]) * 0.0
    out["cameras"] = dict(scene["cameras"])
    out["cameras"]["cam_grid"] = dict(position=tuple(c + np.array([-0.3 * ext[0], 2.5 * ext[1], -0.3 * ext[2]])), direction=(1, -0.6, 1), fov_deg=65.0)
    return out


def scene_num_triangles(scene):
    per_geo = scene["geometries"]["index_count"].astype(np.int64) // 3
    if not scene["instances"]:
        return int(per_geo.sum())
    return int(sum(per_geo[f:f + c].sum() for f, c, _ in scene["instances"]))


SPONZA_DETAIL = 0.8  # synthetic_atrium(SPONZA_DETAIL) has 262,924 triangles (~ the 262k inferred for sponza_scene.glb)


def world_triangles(scene):
    """(N,3,3) float64 world-space triangles in global-id order (instances outer, geometries, primitives).
    numpy re-derivation used only to *generate* ray sets; parity uses the oracle's own flattening."""
    v, idx, g = scene["vertices"], scene["indices"], scene["geometries"]
    inst = scene["instances"] or [(0, len(g), IDENTITY_3X4)]
    out = []
    for first, count, ix in inst:
        im = np.asarray(ix, np.float64).reshape(3, 4)
        for k in range(first, first + count):
            gi = g[k]
            ids = idx[gi["index_offset"]: gi["index_offset"] + gi["index_count"]].astype(np.int64) + int(gi["vertex_offset"])
            p = v["p"][ids].astype(np.float64)
            gm = gi["xform"].astype(np.float64).reshape(3, 4)
            p = p @ gm[:, :3].T + gm[:, 3]
            p = p @ im[:, :3].T + im[:, 3]
            out.append(p.reshape(-1, 3, 3))
    return np.concatenate(out)


def _cosine_dirs(n, rng):
    u1, u2 = rng.random(len(n)), rng.random(len(n))
    r, phi = np.sqrt(u1), 2 * np.pi * u2
    a = np.where(np.abs(n[:, :1]) > 0.9, np.array([[0, 1.0, 0]]), np.array([[1.0, 0, 0]]))
    t = np.cross(n, a)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(n, t)
    return t * (r * np.cos(phi))[:, None] + b * (r * np.sin(phi))[:, None] + n * np.sqrt(np.maximum(0, 1 - u1))[:, None]


def incoherent_rays(scene, n, seed=0x5EED0004, tris=None):
    """cfg4 ray set: area-weighted surface points, cosine-hemisphere diffuse rays
    (closest-hit set) and point-to-point shadow rays (any-hit set).  Returned unsorted = incoherent.
    Returns (closest_rays, any_rays), each a RAY array of length n."""
    rng = np.random.default_rng(seed)
    T = world_triangles(scene) if tris is None else tris
    e1, e2 = T[:, 1] - T[:, 0], T[:, 2] - T[:, 0]
    ng = np.cross(e1, e2)
    area = 0.5 * np.linalg.norm(ng, axis=1)
    ok = area > 0
    cdf = np.cumsum(np.where(ok, area, 0.0))
    cdf /= cdf[-1]

    def surface_points(m):
        k = np.minimum(np.searchsorted(cdf, rng.random(m)), len(T) - 1)
        a, b = rng.random(m), rng.random(m)
        flip = a + b > 1
        a, b = np.where(flip, 1 - a, a), np.where(flip, 1 - b, b)
        p = T[k, 0] + a[:, None] * e1[k] + b[:, None] * e2[k]
        nn = ng[k] / (np.linalg.norm(ng[k], axis=1, keepdims=True) + 1e-30)
        nn *= np.where(rng.random(m) < 0.5, -1.0, 1.0)[:, None]
        return p, nn

    p, nn = surface_points(n)
    closest = np.zeros(n, RAY)
    closest["o"] = p + 1e-3 * nn
    closest["d"] = _cosine_dirs(nn, rng)
    closest["tmin"], closest["tmax"] = 1e-3, 1e5
    q, _ = surface_points(n)
    o = p + 1e-3 * nn
    d = q - o
    dist = np.linalg.norm(d, axis=1)
    anyr = np.zeros(n, RAY)
    anyr["o"] = o
    anyr["d"] = d / np.maximum(dist, 1e-12)[:, None]
    anyr["tmin"] = 1e-3
    anyr["tmax"] = (dist * (1 - 1e-3)).astype(np.float32)
    return closest, anyr


def random_rays(scene, n, seed=3):
    """Uniform random origins in the (slightly enlarged) scene AABB, uniform directions; includes
    axis-parallel directions with exact zeros (slab-test edge case)."""
    rng = np.random.default_rng(seed)
    p = scene["vertices"]["p"]
    lo, hi = p.min(0) - 1.0, p.max(0) + 1.0
    r = np.zeros(n, RAY)
    r["o"] = rng.uniform(lo, hi, (n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = max(1, n // 16)
    axes = np.eye(3)[rng.integers(0, 3, k)] * np.where(rng.random(k) < 0.5, -1, 1)[:, None]
    d[:k] = axes
    r["d"] = d
    r["tmin"], r["tmax"] = 0.0, 1e5
    return r


# ------------------------------------------------------------------------------------------------ the real assets, when present
def default_cameras(scene):
    """SURVEY.md section 8d preset for a scene that brings no camera: eye at the AABB centre - 0.35 x the long-axis extent, at 0.25 x the
    AABB height, looking down the long (horizontal) axis, up +Y, fov 65 degrees (main.rs:69-76)."""
    t = world_triangles(scene).reshape(-1, 3)
    lo, hi = t.min(0), t.max(0)
    c, ext = 0.5 * (lo + hi), hi - lo
    axis = 0 if ext[0] >= ext[2] else 2
    eye = c.copy()
    eye[axis] -= 0.35 * ext[axis]
    eye[1] = lo[1] + 0.25 * ext[1]
    d = np.zeros(3)
    d[axis] = 1.0
    return {"cam0": dict(position=tuple(float(x) for x in eye), direction=tuple(float(x) for x in d), fov_deg=65.0)}


def load_real_assets(directory):
    """The scene resolution rule of SURVEY.md section 8d: `sponza_scene.glb` (and `skybox2.exr`) from `directory` if the files are
    there (they are not part of the reference mount, .MISSING_LARGE_BLOBS:2-3), else None.  Returns a scene dictionary named
    'sponza_glb'; callers tag every number with scene['name']."""
    import os
    from . import assets
    glb = os.path.join(directory, "sponza_scene.glb")
    if not os.path.isfile(glb):
        return None
    with open(glb, "rb") as f:
        sc = assets.load_glb(f.read(), name="sponza_glb")
    exr = os.path.join(directory, "skybox2.exr")
    if os.path.isfile(exr):
        with open(exr, "rb") as f:
            sc["environment"] = assets.load_environment_exr(f.read())
    else:
        sc["environment"] = analytic_sky()
    if not sc.get("cameras"):
        sc["cameras"] = default_cameras(sc)
    return sc
