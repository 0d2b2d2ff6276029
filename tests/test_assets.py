"""Asset path: bincode `.mesh` codec, GLB import / export, meshopt-role transforms.

Golden vector: tests/golden/ref_box_mesh.npz holds the bytes of /root/reference/imported_assets/Default/box.glb, a `.mesh` file
written by the reference's own MeshSaver (older struct revision): it pins varint, Material (5 x f16 BE + u16 varint) and Vertex
(8 x f32 BE) encodings.  glTF parsing and meshlet building have no reference artefact ("parity unpinned": gltf 1.4.1 / meshopt 0.4.1
are third-party) and are covered by round trips and by the glTF 2.0 specification's rules."""
import json
import os
import struct

import numpy as np
import pytest

from raytracer3_b200 import assets, scenes
from raytracer3_b200.types import VERTEX

HERE = os.path.dirname(os.path.abspath(__file__))


def test_reference_processed_asset_decodes():
    g = np.load(os.path.join(HERE, "golden", "ref_box_mesh.npz"))
    raw = g["raw"].tobytes()
    assert len(raw) == 6253
    mats, verts = assets.read_mesh_legacy(raw)
    assert len(mats) == 8 and len(verts) == 192                       # the 8-object Cornell scene of refrence.png, 24 vertices per box
    assert (mats["texture_offset"] == assets.NO_TEXTURE).all()        # `!0u16` written as fb ff ff
    assert (mats["metallic"] == 0).all() and list(mats["roughness"]) == [1, 1, 1, 1, 1, 1, 1, 0.5]
    f16 = lambda x: float(np.float16(x))
    assert [float(x) for x in mats["color"][0]] == [f16(0.8)] * 3 and [float(x) for x in mats["color"][5]] == [0.5] * 3
    assert float(mats["color"][2][2]) == pytest.approx(0.8, abs=1e-3) and float(mats["color"][3][0]) == pytest.approx(0.8, abs=1e-3)
    assert float(mats["color"][4][1]) == pytest.approx(0.8, abs=1e-3)                                   # blue, red, green walls
    p, n = verts["p"], verts["n"]
    assert set(np.unique(p)) == {-1.0, 1.0}                            # unit cubes
    assert np.allclose(np.linalg.norm(n, axis=1), 1) and set(np.unique(n)) == {-1.0, 0.0, 1.0}
    assert (verts["t"] >= 0).all() and (verts["t"] <= 1).all()
    # re-encoding the decoded elements with the writer reproduces the reference's bytes exactly
    w = assets.BincodeWriter()
    w.varint(0); w.varint(0); w.varint(len(mats))
    for m in mats:
        assets._write_material(w, m)
    w.varint(len(verts))
    w.f32s(np.concatenate([verts["p"], verts["n"], verts["t"]], axis=1))
    w.raw(b"\x00")
    assert w.bytes() == raw


def test_bincode_varint_boundaries():
    for v, enc in ((0, b"\x00"), (250, b"\xfa"), (251, b"\xfb\x00\xfb"), (65535, b"\xfb\xff\xff"), (65536, b"\xfc\x00\x01\x00\x00"),
                   (2 ** 32 - 1, b"\xfc\xff\xff\xff\xff"), (2 ** 32, b"\xfd\x00\x00\x00\x01\x00\x00\x00\x00")):
        w = assets.BincodeWriter(); w.varint(v)
        assert w.bytes() == enc                                       # bincode spec, VarintEncoding, big endian
        assert assets.BincodeReader(enc).varint() == v
    with pytest.raises(ValueError):
        assets.BincodeReader(b"\xfb\x00").varint()
    with pytest.raises(ValueError):
        assets.BincodeReader(b"\x02").bool()


def _geometry_mesh(sc, gi, **mat):
    g = sc["geometries"][gi]
    idx = sc["indices"][g["index_offset"]:g["index_offset"] + g["index_count"]]
    return assets.process_primitive(sc["vertices"][g["vertex_offset"]:], idx, assets.material_record(**mat)), sc["vertices"][g["vertex_offset"]:]["p"][idx.reshape(-1, 3)]


def test_mesh_round_trip_and_meshlet_limits():
    sc = scenes.synthetic_atrium(0.05)
    gi = int(np.argmax(sc["geometries"]["index_count"]))
    mesh, tris = _geometry_mesh(sc, gi, base_color=(0.25, 0.5, 1.0), metallic=0.5, roughness=0.25, texture=3)
    assert len(mesh.meshlets) > 3
    assert mesh.meshlets["vertex_count"].max() <= 64 and mesh.meshlets["triangle_count"].max() <= 124       # assets/mod.rs:267
    assert (mesh.meshlets["triangle_offset"] % 4 == 0).all()
    for m in mesh.meshlets:                                           # meshlet-local indices stay inside the meshlet's vertex list
        loc = mesh.indices[m["triangle_offset"]:m["triangle_offset"] + 3 * m["triangle_count"]]
        assert loc.max() < m["vertex_count"]
    raw = assets.write_mesh(mesh)
    back = assets.read_mesh(raw)
    assert assets.write_mesh(back) == raw
    for k in ("meshlets", "materials", "vertices", "indices"):
        assert np.array_equal(getattr(back, k), getattr(mesh, k)), k
    assert back.uploaded is False
    assert np.array_equal(back.vertices["p"][back.triangles()], tris)   # same triangles, same order (scan builder keeps index order)
    assert float(back.materials[0]["roughness"]) == 0.25
    quirk = assets.read_mesh(raw, reference_decode_quirk=True)         # assets/mod.rs:89-90: roughness decoded from the metallic bytes
    assert float(quirk.materials[0]["roughness"]) == 0.5
    with pytest.raises(ValueError):
        assets.read_mesh(raw[:-3])
    with pytest.raises(ValueError):
        assets.read_mesh(raw + b"\x00")
    empty = assets.read_mesh(assets.write_mesh(assets.MeshFile()))
    assert len(empty.meshlets) == 0 and len(empty.vertices) == 0 and empty.triangles().shape == (0, 3)


def test_optimize_vertex_fetch_orders_by_first_use():
    v = np.zeros(6, VERTEX); v["p"][:, 0] = np.arange(6)
    idx = np.array([4, 2, 5, 2, 4, 0], np.uint32)                      # vertex 1 and 3 unused
    ni, nv = assets.optimize_vertex_fetch(idx, v)
    assert list(ni) == [0, 1, 2, 1, 0, 3] and list(nv["p"][:, 0]) == [4, 2, 5, 0]


def test_glb_round_trip_full_scene():
    base = scenes.synthetic_atrium(0.01)
    sc = scenes.instanced_grid(base, 2)
    raw = assets.write_glb(sc)
    doc, blob = assets.split_glb(raw)
    assert doc["asset"]["version"] == "2.0" and len(doc["meshes"]) == len(sc["geometries"]) and len(blob) % 4 == 0
    back = assets.load_glb(raw)
    assert np.array_equal(scenes.world_triangles(back), scenes.world_triangles(sc))    # node transforms, global primitive order
    assert sorted(back["textures"]) == sorted(sc["textures"])
    for k in sc["textures"]:
        assert np.array_equal(back["textures"][k], sc["textures"][k])
    for a, b in zip(back["geometries"], sc["geometries"]):
        for f in ("base_color", "metallic", "roughness", "base_color_tex", "index_count"):
            assert np.array_equal(a[f], b[f]), f
        assert np.array_equal(a["emission"][:3], b["emission"][:3])
    first = assets.load_glb(raw, first_primitive_only=True)             # the reference's MeshTransformer view (assets/mod.rs:218)
    assert len(first["geometries"]) == 1 and first["instances"] == []


def _mini_gltf(nodes, extra=None, indices=(0, 1, 2), comp=5123):
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], "<f4").tobytes()
    dt = {5121: "u1", 5123: "<u2", 5125: "<u4"}[comp]
    ib = np.array(indices, dt).tobytes()
    blob = pos + ib + b"\x00" * (-len(ib) % 4)
    doc = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}], "nodes": nodes,
           "meshes": [{"primitives": [{"attributes": {"POSITION": 0}, "indices": 1}]}],
           "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"},
                         {"bufferView": 1, "componentType": comp, "count": len(indices), "type": "SCALAR"}],
           "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 36}, {"buffer": 0, "byteOffset": 36, "byteLength": len(ib)}],
           "buffers": [{"byteLength": len(blob)}]}
    doc.update(extra or {})
    js = json.dumps(doc).encode()
    js += b" " * (-len(js) % 4)
    return struct.pack("<III", 0x46546C67, 2, 28 + len(js) + len(blob)) + struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(blob), 0x004E4942) + blob


def test_gltf_node_hierarchy_trs_and_index_types():
    # parent: translate (1,2,3) * rotate 90 degrees about +Z * scale 2; child carries the mesh with its own translation (1,0,0)
    s = float(np.sqrt(0.5))
    nodes = [{"children": [1], "translation": [1, 2, 3], "rotation": [0, 0, s, s], "scale": [2, 2, 2]}, {"mesh": 0, "translation": [1, 0, 0]}]
    for comp in (5121, 5123, 5125):
        sc = assets.load_glb(_mini_gltf(nodes, comp=comp))
        assert len(sc["instances"]) == 1
        tri = scenes.world_triangles(sc)[0]
        # local (x,y,z) -> child (+1,0,0) -> scale 2 -> rotate z 90: (x,y)->(-y,x) -> translate
        want = np.array([[1 - 0, 2 + 2, 3], [1 - 0, 2 + 4, 3], [1 - 2, 2 + 2, 3]], np.float64)
        assert np.allclose(tri, want, atol=1e-5), tri
        assert np.allclose(sc["vertices"]["n"], [[0, 0, 1]] * 3)        # generated normals (no NORMAL attribute)
        assert sc["geometries"][0]["metallic"] == 1.0 and sc["geometries"][0]["roughness"] == 1.0   # glTF defaults
    m = np.eye(4); m[:3, 3] = [5, 6, 7]
    sc = assets.load_glb(_mini_gltf([{"mesh": 0, "matrix": [float(x) for x in m.T.reshape(-1)]}]))
    assert np.allclose(scenes.world_triangles(sc)[0][0], [5, 6, 7])


def test_gltf_errors():
    good = _mini_gltf([{"mesh": 0}])
    with pytest.raises(ValueError):
        assets.load_glb(b"glTF")
    with pytest.raises(ValueError):
        assets.load_glb(b"XXXX" + good[4:])
    with pytest.raises(ValueError):
        assets.load_glb(_mini_gltf([{"mesh": 0}], indices=(0, 1, 7)))       # index out of range
    with pytest.raises(ValueError):
        assets.load_glb(_mini_gltf([{"mesh": 0, "children": [0]}]))          # cycle
    with pytest.raises(ValueError):
        assets.load_glb(good[:len(good) - 20])


def test_reference_pipeline_glb_to_mesh():
    raw = assets.import_glb_to_mesh(assets.write_glb(scenes.cornell_box()))
    mesh = assets.read_mesh(raw)
    assert len(mesh.meshlets) == 1 and len(mesh.materials) == 1 and mesh.triangles().shape == (2, 3)
    sc = assets.mesh_to_scene(mesh)
    assert scenes.world_triangles(sc).shape == (2, 3, 3)


# ---------------------------------------------------------------------------------------------------------------- OpenEXR (skybox2.exr role)
@pytest.mark.parametrize("compression", ["NONE", "RLE", "ZIPS", "ZIP"])
@pytest.mark.parametrize("half", [False, True])
def test_exr_round_trip(compression, half):
    rng = np.random.default_rng(3)
    img = (rng.random((37, 53, 3)) * 40).astype(np.float32)
    img[5:22, 10:31] = 2.5                                           # runs, so that RLE / ZIP blocks really shrink
    img[30:] = 0
    data = assets.write_exr(img, compression, half, data_window_origin=(3, -2))
    px, names = assets.read_exr(data)
    assert names == ["B", "G", "R"] and px.shape == (37, 53, 3)
    want = img.astype(np.float16).astype(np.float32) if half else img
    assert (assets.load_environment_exr(data) == want).all()
    if compression != "NONE":
        assert len(data) < len(assets.write_exr(img, "NONE", half))


def test_exr_reader_against_hand_assembled_file():
    """Bytes assembled here straight from the published layout (magic, version, attribute list, offset table, one chunk per scanline,
    channels stored alphabetically row by row): independent of write_exr."""
    import struct
    W, H = 3, 2
    r = np.array([[1.0, 2.0, 3.0], [4.0, 5.0, 6.0]], np.float16)
    g = np.array([[0.5, 0.25, 0.125], [8.0, 16.0, 32.0]], np.float32)
    z = np.array([[7, 8, 9], [10, 11, 4000000000]], np.uint32)

    def attr(name, typ, val):
        return name + b"\0" + typ + b"\0" + struct.pack("<i", len(val)) + val
    ch = (b"G\0" + struct.pack("<iB3xii", 2, 0, 1, 1) + b"R\0" + struct.pack("<iB3xii", 1, 0, 1, 1) + b"Z\0" + struct.pack("<iB3xii", 0, 0, 1, 1) + b"\0")
    hdr = (bytes([0x76, 0x2F, 0x31, 0x01]) + struct.pack("<i", 2) + attr(b"channels", b"chlist", ch) + attr(b"compression", b"compression", b"\0")
           + attr(b"dataWindow", b"box2i", struct.pack("<4i", 0, 0, W - 1, H - 1)) + attr(b"displayWindow", b"box2i", struct.pack("<4i", 0, 0, W - 1, H - 1))
           + attr(b"lineOrder", b"lineOrder", b"\0") + attr(b"pixelAspectRatio", b"float", struct.pack("<f", 1.0))
           + attr(b"screenWindowCenter", b"v2f", struct.pack("<ff", 0, 0)) + attr(b"screenWindowWidth", b"float", struct.pack("<f", 1.0)) + b"\0")
    rows = [g[y].tobytes() + r[y].tobytes() + z[y].tobytes() for y in range(H)]
    chunks = [struct.pack("<ii", y, len(rows[y])) + rows[y] for y in range(H)]
    t0 = len(hdr) + 8 * H
    table = struct.pack("<2Q", t0 + len(chunks[1]), t0)               # file order: scanline 1 first (decreasing-y files exist)
    data = hdr + table + chunks[1] + chunks[0]
    px, names = assets.read_exr(data)
    assert names == ["G", "R", "Z"]
    assert (px[..., 0] == g).all() and (px[..., 1] == r.astype(np.float32)).all() and (px[..., 2] == z.astype(np.float32)).all()
    env = assets.load_environment_exr(data)
    assert (env[..., 0] == r.astype(np.float32)).all() and (env[..., 1] == g).all() and (env[..., 2] == 0).all()


def test_exr_unsupported_and_malformed_inputs_fail_loudly():
    img = np.ones((4, 4, 3), np.float32)
    data = bytearray(assets.write_exr(img, "NONE"))
    i = data.index(b"compression\0compression\0") + len(b"compression\0compression\0") + 4
    data[i] = 4                                                      # PIZ
    with pytest.raises(NotImplementedError, match="PIZ"):
        assets.read_exr(bytes(data))
    with pytest.raises(ValueError):
        assets.read_exr(b"\x89PNG\r\n\x1a\n" + b"\0" * 32)
    tiled = bytearray(assets.write_exr(img, "NONE"))
    tiled[5] |= 0x02                                                 # version field bit 9: tiled
    with pytest.raises(NotImplementedError):
        assets.read_exr(bytes(tiled))


def test_real_asset_hook_uses_files_when_present(tmp_path):
    """SURVEY 8d scene rule: sponza_scene.glb / skybox2.exr are used if they exist, else the stand-in (None here)."""
    assert scenes.load_real_assets(str(tmp_path)) is None
    src = scenes.synthetic_atrium(0.01, environment=False)
    (tmp_path / "sponza_scene.glb").write_bytes(assets.write_glb(src))
    sky = scenes.analytic_sky(32, 16)
    (tmp_path / "skybox2.exr").write_bytes(assets.write_exr(sky, "ZIP"))
    sc = scenes.load_real_assets(str(tmp_path))
    assert sc["name"] == "sponza_glb" and scenes.scene_num_triangles(sc) == scenes.scene_num_triangles(src)
    assert (sc["environment"] == sky).all()
    cam = sc["cameras"]["cam0"]
    t = scenes.world_triangles(sc).reshape(-1, 3)
    lo, hi = t.min(0), t.max(0)
    assert lo[0] < cam["position"][0] < hi[0] and abs(cam["position"][1] - (lo[1] + 0.25 * (hi[1] - lo[1]))) < 1e-6   # inside, a quarter up
    assert cam["direction"] == (1.0, 0.0, 0.0)                      # the atrium's long axis is x
