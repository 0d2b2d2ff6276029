"""Asset path either side of the hot path: the data formats the Rust host produces and consumes.

Mirrors /root/reference/src/assets/mod.rs:

  Mesh / Meshlet / Material / Vertex      assets/mod.rs:41-133     MeshFile, MESHLET, MATERIAL, VERTEX below
  CONFIG (bincode 2, big endian, varint)  assets/mod.rs:135-137    BincodeReader / BincodeWriter
  MeshLoader (".mesh")                    assets/mod.rs:151-169    read_mesh
  MeshSaver                               assets/mod.rs:298-314    write_mesh
  GltfMeshLoader (".glb")                 assets/mod.rs:180-201    load_glb (full scene: every mesh, primitive, node transform)
  MeshTransformer                         assets/mod.rs:208-284    process_primitive (vertex-fetch remap + meshlets 64 / 124)
  skybox2.exr (image / exr crates)        main.rs:94 (commented)   read_exr / load_environment_exr (scanline OpenEXR: NONE, RLE, ZIPS, ZIP)

Third-party arithmetic absent from /root/reference (parity unpinned, algorithms restated from their published descriptions):
bincode 2.0.1 (Cargo.lock:814-815; format spec: github.com/bincode-org/bincode docs/spec.md), gltf 1.4.1 (Cargo.lock:1741-1742;
Khronos glTF 2.0 specification), meshopt 0.4.1 (Cargo.lock:2234-2235; meshoptimizer: optimizeVertexFetch, buildMeshletsScan).
The one artefact of this pipeline that the reference ships, imported_assets/Default/box.glb (a processed `.mesh` written by an
older revision of the struct), is decoded by `read_mesh_legacy` and pinned in tests/test_assets.py.

Host-side only: numpy + struct, no CUDA.  Output of load_glb / mesh_to_scene is the scene dictionary of raytracer3_b200.scenes,
i.e. exactly what RenderWorld.load_scene pushes through the C ABI (rt3_scene_push_*).
"""
import json
import struct
import zlib

import numpy as np

from .types import GEOMETRY, VERTEX

MESHLET = np.dtype([("vertex_offset", "<u4"), ("triangle_offset", "<u4"), ("vertex_count", "<u4"), ("triangle_count", "<u4")])   # assets/mod.rs:41-49
MATERIAL = np.dtype([("metallic", "<f2"), ("roughness", "<f2"), ("color", "<f2", 3), ("texture_offset", "<u2")])                   # assets/mod.rs:51-59
NO_TEXTURE = 0xFFFF                                                                                                                # `!0u16`, assets/mod.rs:251

MAX_MESHLET_VERTICES, MAX_MESHLET_TRIANGLES = 64, 124                                                                              # assets/mod.rs:267


# ---------------------------------------------------------------------------------------------------------------- bincode
class BincodeWriter:
    """bincode 2 `standard().with_variable_int_encoding().with_big_endian()` (assets/mod.rs:135-137)."""

    def __init__(self):
        self.parts = []

    def varint(self, v):
        v = int(v)
        if v < 251:
            self.parts.append(struct.pack(">B", v))
        elif v < 1 << 16:
            self.parts.append(b"\xfb" + struct.pack(">H", v))
        elif v < 1 << 32:
            self.parts.append(b"\xfc" + struct.pack(">I", v))
        else:
            self.parts.append(b"\xfd" + struct.pack(">Q", v))

    def raw(self, b):
        self.parts.append(bytes(b))

    def f32s(self, a):
        self.parts.append(np.ascontiguousarray(a, ">f4").tobytes())

    def f16s(self, a):
        self.parts.append(np.ascontiguousarray(a, ">f2").tobytes())

    def bool(self, v):
        self.parts.append(b"\x01" if v else b"\x00")

    def bytes(self):
        return b"".join(self.parts)


class BincodeReader:
    def __init__(self, data):
        self.d, self.o = memoryview(bytes(data)), 0

    def take(self, n):
        if self.o + n > len(self.d):
            raise ValueError(f"bincode: unexpected end of data at {self.o} (+{n} of {len(self.d)})")
        b = self.d[self.o:self.o + n]
        self.o += n
        return b

    def varint(self):
        t = self.take(1)[0]
        if t < 251:
            return t
        if t == 251:
            return struct.unpack(">H", self.take(2))[0]
        if t == 252:
            return struct.unpack(">I", self.take(4))[0]
        if t == 253:
            return struct.unpack(">Q", self.take(8))[0]
        raise ValueError(f"bincode: unsupported varint marker {t} at {self.o - 1}")

    def f32s(self, n):
        return np.frombuffer(self.take(4 * n), ">f4").astype("<f4")

    def f16s(self, n):
        return np.frombuffer(self.take(2 * n), ">f2").astype("<f2")

    def bool(self):
        v = self.take(1)[0]
        if v > 1:
            raise ValueError(f"bincode: invalid bool {v} at {self.o - 1}")
        return bool(v)

    def done(self):
        return self.o == len(self.d)


# ---------------------------------------------------------------------------------------------------------------- .mesh
class MeshFile:
    """`Mesh` of assets/mod.rs:118-125: meshlets, materials, meshlet-ordered vertices, meshlet-local u8 triangle indices."""

    def __init__(self, meshlets=None, materials=None, vertices=None, indices=None, uploaded=False):
        self.meshlets = np.zeros(0, MESHLET) if meshlets is None else np.ascontiguousarray(meshlets, MESHLET)
        self.materials = np.zeros(0, MATERIAL) if materials is None else np.ascontiguousarray(materials, MATERIAL)
        self.vertices = np.zeros(0, VERTEX) if vertices is None else np.ascontiguousarray(vertices, VERTEX)
        self.indices = np.zeros(0, np.uint8) if indices is None else np.ascontiguousarray(indices, np.uint8)
        self.uploaded = bool(uploaded)

    def triangles(self):
        """(T, 3) u32 indices into `vertices`: meshlet m's triangle t is indices[triangle_offset + 3t ..] + vertex_offset
        (the addressing of shaders/passes/bindless_test2.slang:31-58)."""
        out = []
        for m in self.meshlets:
            t = self.indices[m["triangle_offset"]:m["triangle_offset"] + 3 * m["triangle_count"]].astype(np.uint32)
            out.append(t.reshape(-1, 3) + np.uint32(m["vertex_offset"]))
        return np.concatenate(out) if out else np.zeros((0, 3), np.uint32)


def _write_material(w, m):
    # custom Encode, assets/mod.rs:61-72: five big-endian f16 written raw, then texture_offset as a bincode u16 (varint)
    w.f16s([m["metallic"], m["roughness"], m["color"][0], m["color"][1], m["color"][2]])
    w.varint(int(m["texture_offset"]))


def _read_material(r, reference_decode_quirk):
    h = r.f16s(5)
    m = np.zeros((), MATERIAL)
    m["metallic"] = h[0]
    # the reference's Decode builds roughness_factor from the METALLIC bytes (assets/mod.rs:89-90, 110-111): a reader that
    # must agree with the Rust host's in-memory Mesh reproduces that; the default returns what the file actually holds
    m["roughness"] = h[0] if reference_decode_quirk else h[1]
    m["color"] = h[2:5]
    m["texture_offset"] = r.varint()
    return m


def write_mesh(mesh):
    """MeshSaver (assets/mod.rs:298-314): bincode::encode_into_writer(mesh, CONFIG).  Returns bytes."""
    w = BincodeWriter()
    w.varint(len(mesh.meshlets))
    for m in mesh.meshlets:                       # derive(Encode) on four u32 fields: each a varint
        for k in ("vertex_offset", "triangle_offset", "vertex_count", "triangle_count"):
            w.varint(int(m[k]))
    w.varint(len(mesh.materials))
    for m in mesh.materials:
        _write_material(w, m)
    w.varint(len(mesh.vertices))
    if len(mesh.vertices):                        # derive(Encode) on [f32; 3], [f32; 3], [f32; 2]: fixed arrays, no length, BE floats
        v = mesh.vertices
        w.f32s(np.concatenate([v["p"], v["n"], v["t"]], axis=1))
    w.varint(len(mesh.indices))
    w.raw(mesh.indices.tobytes())                 # Vec<u8>: length + raw bytes
    w.bool(mesh.uploaded)
    return w.bytes()


def _read_vertices(r, n):
    f = r.f32s(8 * n).reshape(n, 8)
    v = np.zeros(n, VERTEX)
    v["p"], v["n"], v["t"] = f[:, 0:3], f[:, 3:6], f[:, 6:8]
    return v


def read_mesh(data, reference_decode_quirk=False):
    """MeshLoader (assets/mod.rs:151-169): bincode::decode_from_reader(reader, CONFIG) -> Mesh."""
    r = BincodeReader(data)
    n = r.varint()
    ml = np.zeros(n, MESHLET)
    for i in range(n):
        for k in ("vertex_offset", "triangle_offset", "vertex_count", "triangle_count"):
            ml[i][k] = r.varint()
    materials = np.array([_read_material(r, reference_decode_quirk) for _ in range(r.varint())], MATERIAL).reshape(-1)
    vertices = _read_vertices(r, r.varint())
    n = r.varint()
    indices = np.frombuffer(r.take(n), np.uint8).copy()
    uploaded = r.bool()
    if not r.done():
        raise ValueError(f"bincode: {len(r.d) - r.o} trailing bytes")
    mesh = MeshFile(ml, materials, vertices, indices, uploaded)
    for m in ml:
        if m["vertex_offset"] + m["vertex_count"] > len(vertices) or m["triangle_offset"] + 3 * m["triangle_count"] > len(indices):
            raise ValueError("meshlet range outside the vertex / index arrays")
    return mesh


def read_mesh_legacy(data):
    """The processed asset the reference ships (imported_assets/Default/box.glb + .meta naming MeshLoader) was written by an
    OLDER revision of `Mesh`: two empty leading Vecs, Vec<Material>, Vec<Vertex>, one trailing zero byte (6253 bytes: 8
    materials x 13 B, 192 vertices x 32 B).  Element encodings are the current ones, so this pins the Material / Vertex /
    varint codecs against bytes produced by the reference itself.  Returns (materials, vertices)."""
    r = BincodeReader(data)
    if r.varint() != 0 or r.varint() != 0:
        raise ValueError("not the legacy layout (expected two empty leading sequences)")
    materials = np.array([_read_material(r, False) for _ in range(r.varint())], MATERIAL).reshape(-1)
    vertices = _read_vertices(r, r.varint())
    tail = bytes(r.take(len(r.d) - r.o))
    if tail not in (b"\x00", b"\x00\x00"):
        raise ValueError("unexpected trailer in legacy mesh")
    return materials, vertices


# ---------------------------------------------------------------------------------------------------------------- meshopt role
def optimize_vertex_fetch(indices, vertices):
    """meshopt_optimizeVertexFetch (meshopt::optimize_vertex_fetch_in_place, assets/mod.rs:261): vertices are renumbered in
    order of first use by the index buffer, unused vertices are dropped.  Returns (new_indices, new_vertices)."""
    indices = np.ascontiguousarray(indices, np.uint32)
    uniq, first = np.unique(indices, return_index=True)
    order = uniq[np.argsort(first, kind="stable")]                 # old vertex ids in order of first reference
    remap = np.full(len(vertices), 0xFFFFFFFF, np.uint32)
    remap[order] = np.arange(len(order), dtype=np.uint32)
    return remap[indices], np.ascontiguousarray(vertices[order])


def build_meshlets_scan(indices, n_vertices, max_vertices=MAX_MESHLET_VERTICES, max_triangles=MAX_MESHLET_TRIANGLES):
    """meshopt_buildMeshletsScan: triangles are appended in index-buffer order; a meshlet closes when the next triangle would
    exceed max_vertices or max_triangles.  (The reference calls meshopt::build_meshlets with cone_weight 0, assets/mod.rs:267,
    whose kd-tree seeded growth picks triangles in a different ORDER; limits, layout and the consumer-visible contract —
    meshlet-local u8 indices, per-meshlet vertex lists — are identical, the triangle order is "parity unpinned".)
    Returns (meshlets MESHLET[], meshlet_vertices u32[], meshlet_triangles u8[])."""
    indices = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3)
    assert max_vertices <= 256 and max_triangles <= 512
    meshlets, mverts, mtris = [], [], []
    local = {}                                                      # vertex id -> local index inside the open meshlet
    cur_v, cur_t = [], []

    def close():
        nonlocal local, cur_v, cur_t
        if cur_t:
            meshlets.append((len(mverts), len(mtris), len(cur_v), len(cur_t) // 3))
            mverts.extend(cur_v)
            mtris.extend(cur_t)
            while len(mtris) % 4:                                   # meshopt pads every meshlet's triangle bytes to 4
                mtris.append(0)
        local, cur_v, cur_t = {}, [], []

    for a, b, c in indices.tolist():
        extra = (a not in local) + (b not in local and b != a) + (c not in local and c != a and c != b)
        if len(cur_v) + extra > max_vertices or len(cur_t) // 3 >= max_triangles:
            close()
        for v in (a, b, c):
            if v not in local:
                local[v] = len(cur_v)
                cur_v.append(v)
            cur_t.append(local[v])
    close()
    return (np.array(meshlets, np.uint32).reshape(-1, 4).view(MESHLET).reshape(-1), np.array(mverts, np.uint32), np.array(mtris, np.uint8))


def process_primitive(vertices, indices, material):
    """MeshTransformer::transform for one primitive (assets/mod.rs:216-284): vertex-fetch remap, meshlets of 64 vertices / 124
    triangles, vertices duplicated into meshlet order.  (optimize_vertex_cache's result is discarded by the reference,
    assets/mod.rs:253; optimize_overdraw only reorders triangles and is left out.)"""
    idx, verts = optimize_vertex_fetch(indices, np.ascontiguousarray(vertices, VERTEX))
    ml, mv, mt = build_meshlets_scan(idx, len(verts))
    return MeshFile(ml, np.array([material], MATERIAL).reshape(-1), verts[mv], mt, False)


def material_record(base_color=(0.8, 0.8, 0.8), metallic=0.0, roughness=1.0, texture=None):
    m = np.zeros((), MATERIAL)
    m["metallic"], m["roughness"], m["color"] = metallic, roughness, base_color[:3]
    m["texture_offset"] = NO_TEXTURE if texture is None else texture
    return m


def mesh_to_scene(mesh, name="mesh", reference_decode_quirk=False):
    """A processed `.mesh` as a scene dictionary for RenderWorld.load_scene (one geometry: the reference's MeshTransformer
    keeps a single primitive and a single material, assets/mod.rs:218,245-252)."""
    tri = mesh.triangles()
    g = np.zeros(1, GEOMETRY)
    m = mesh.materials[0] if len(mesh.materials) else material_record()
    g["base_color"][0] = [float(m["color"][0]), float(m["color"][1]), float(m["color"][2]), 1.0]
    g["base_color_tex"] = -1 if int(m["texture_offset"]) == NO_TEXTURE else int(m["texture_offset"])
    g["metallic"] = float(m["metallic"])
    g["roughness"] = float(m["metallic"] if reference_decode_quirk else m["roughness"])
    g["index_offset"], g["vertex_offset"], g["index_count"] = 0, 0, tri.size
    g["xform"][0] = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
    return {"name": name, "vertices": mesh.vertices.copy(), "indices": tri.reshape(-1).astype(np.uint32), "geometries": g, "instances": [],
            "textures": {}, "environment": None, "cameras": {}}


# ---------------------------------------------------------------------------------------------------------------- images
def _png_decode(data):
    """Minimal PNG reader (8-bit gray / gray+alpha / RGB / RGBA, non-interlaced) used when Pillow is absent."""
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG")
    o, idat, hdr = 8, [], None
    while o < len(data):
        n, typ = struct.unpack(">I4s", data[o:o + 8])
        body = data[o + 8:o + 8 + n]
        o += 12 + n
        if typ == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat.append(body)
        elif typ == b"IEND":
            break
    w, h, depth, ctype, _, _, interlace = hdr
    ch = {0: 1, 2: 3, 4: 2, 6: 4}.get(ctype)
    if depth != 8 or ch is None or interlace:
        raise ValueError("unsupported PNG variant (need 8-bit non-interlaced gray/RGB/RGBA)")
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), np.uint8).reshape(h, 1 + w * ch)
    out = np.zeros((h, w * ch), np.uint8)
    prev = np.zeros(w * ch, np.int32)
    for y in range(h):
        f, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if f == 0:
            cur = line
        elif f == 2:
            cur = (line + prev) & 255
        else:                                       # Sub, Average, Paeth need the running left neighbour
            cur = np.zeros(w * ch, np.int32)
            for i in range(w * ch):
                a = cur[i - ch] if i >= ch else 0
                b = prev[i]
                c = prev[i - ch] if i >= ch else 0
                if f == 1:
                    p = a
                elif f == 3:
                    p = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    p = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[i] = (line[i] + p) & 255
        out[y] = cur
        prev = cur
    img = out.reshape(h, w, ch)
    if ch == 1:
        img = np.concatenate([img] * 3 + [np.full((h, w, 1), 255, np.uint8)], 2)
    elif ch == 2:
        img = np.concatenate([img[..., :1]] * 3 + [img[..., 1:]], 2)
    elif ch == 3:
        img = np.concatenate([img, np.full((h, w, 1), 255, np.uint8)], 2)
    return np.ascontiguousarray(img)


def png_encode(rgba8):
    """RGBA8 (h, w, 4) -> PNG bytes (filter 0 on every row)."""
    img = np.ascontiguousarray(rgba8, np.uint8)
    h, w, _ = img.shape
    raw = np.concatenate([np.zeros((h, 1), np.uint8), img.reshape(h, w * 4)], 1).tobytes()

    def chunk(t, b):
        return struct.pack(">I", len(b)) + t + b + struct.pack(">I", zlib.crc32(t + b) & 0xFFFFFFFF)
    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b"")


def decode_image(data):
    """Image bytes -> RGBA8 (h, w, 4): the `image` crate's role behind gltf::import_slice (assets/mod.rs:193)."""
    data = bytes(data)
    try:
        import io

        from PIL import Image
        return np.ascontiguousarray(np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"), np.uint8))
    except ImportError:
        return _png_decode(data)


# ---------------------------------------------------------------------------------------------------------------- glTF / GLB
_GLB_MAGIC, _CHUNK_JSON, _CHUNK_BIN = 0x46546C67, 0x4E4F534A, 0x004E4942
_COMPONENT = {5120: ("i1", 1), 5121: ("u1", 1), 5122: ("<i2", 2), 5123: ("<u2", 2), 5125: ("<u4", 4), 5126: ("<f4", 4)}
_TYPE_N = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT2": 4, "MAT3": 9, "MAT4": 16}


def split_glb(data):
    """GLB container (glTF 2.0 section 4.4) -> (json dict, bin chunk bytes or b'')."""
    data = bytes(data)
    if len(data) < 12:
        raise ValueError("GLB: file shorter than its header")
    magic, version, length = struct.unpack("<III", data[:12])
    if magic != _GLB_MAGIC or version != 2:
        raise ValueError("GLB: bad magic or version")
    if length > len(data):
        raise ValueError("GLB: declared length exceeds the data")
    o, doc, blob = 12, None, b""
    while o + 8 <= length:
        n, typ = struct.unpack("<II", data[o:o + 8])
        body = data[o + 8:o + 8 + n]
        if len(body) != n:
            raise ValueError("GLB: truncated chunk")
        if typ == _CHUNK_JSON and doc is None:
            doc = json.loads(body.decode("utf-8"))
        elif typ == _CHUNK_BIN and not blob:
            blob = body
        o += 8 + n + (-n % 4)
    if doc is None:
        raise ValueError("GLB: no JSON chunk")
    return doc, blob


def _buffers(doc, blob):
    import base64
    out = []
    for i, b in enumerate(doc.get("buffers", [])):
        uri = b.get("uri")
        if uri is None:
            if i != 0:
                raise ValueError("glTF: only buffer 0 may use the GLB BIN chunk")
            out.append(blob)
        elif uri.startswith("data:"):
            out.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            raise ValueError("glTF: external buffer files are not supported (GLB only, as the reference: assets/mod.rs:199-201)")
        if len(out[-1]) < b.get("byteLength", 0):
            raise ValueError("glTF: buffer shorter than byteLength")
    return out


def read_accessor(doc, buffers, index, as_float=False):
    """Accessor -> (count, n) array.  Normalised integers become floats (glTF 2.0 section 3.6.2.2) when as_float is set."""
    acc = doc["accessors"][index]
    if "sparse" in acc:
        raise ValueError("glTF: sparse accessors are not supported")
    dt, size = _COMPONENT[acc["componentType"]]
    n, count = _TYPE_N[acc["type"]], acc["count"]
    if "bufferView" not in acc:
        arr = np.zeros((count, n), dt)
    else:
        bv = doc["bufferViews"][acc["bufferView"]]
        start = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
        stride = bv.get("byteStride", 0) or n * size
        buf = buffers[bv["buffer"]]
        need = start + (count - 1) * stride + n * size if count else start
        if need > len(buf):
            raise ValueError("glTF: accessor reads past the end of its buffer")
        arr = np.ndarray((count, n), dt, buffer=buf, offset=start, strides=(stride, size)).copy() if count else np.zeros((0, n), dt)
    if as_float and acc["componentType"] != 5126:
        scale = {5120: 127.0, 5121: 255.0, 5122: 32767.0, 5123: 65535.0}.get(acc["componentType"])
        arr = arr.astype(np.float32)
        if acc.get("normalized") and scale:
            arr = np.maximum(arr / np.float32(scale), -1.0).astype(np.float32)
    return arr


def node_local_matrix(node):
    """4x4 row-major local transform of a node: `matrix` (column-major in the file) or T * R * S (glTF 2.0 section 3.5.3)."""
    if "matrix" in node:
        return np.array(node["matrix"], np.float64).reshape(4, 4).T
    t = np.array(node.get("translation", [0, 0, 0]), np.float64)
    x, y, z, w = np.array(node.get("rotation", [0, 0, 0, 1]), np.float64)
    s = np.array(node.get("scale", [1, 1, 1]), np.float64)
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    m = np.eye(4)
    m[:3, :3] = r * s[None, :]
    m[:3, 3] = t
    return m


def _vertex_normals(p, idx):
    """Area-weighted vertex normals for primitives without NORMAL (the reference unwraps, i.e. panics: assets/mod.rs:222-225)."""
    tri = idx.reshape(-1, 3)
    fn = np.cross(p[tri[:, 1]] - p[tri[:, 0]], p[tri[:, 2]] - p[tri[:, 0]])
    n = np.zeros_like(p, dtype=np.float64)
    for k in range(3):
        np.add.at(n, tri[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    return np.where(ln > 0, n / np.maximum(ln, 1e-30), np.array([0.0, 1.0, 0.0])).astype(np.float32)


def load_glb(data, first_primitive_only=False, name="glb"):
    """GLB -> scene dictionary (raytracer3_b200.scenes format).

    Every primitive of every mesh becomes one geometry in the global vertex / index buffers (RenderWorld's layout,
    world/mod.rs:103-117); every node that carries a mesh becomes an instance with the node's world transform (InstanceInfo,
    world/mod.rs:32-37); pbrMetallicRoughness feeds base colour / metallic / roughness / base-colour texture, emissiveFactor the
    emission; images are decoded to RGBA8 texture slots (slot = glTF texture index, the reference's texture_offset,
    assets/mod.rs:251).  first_primitive_only reproduces the reference's MeshTransformer, which keeps
    `meshes().next().primitives().next()` and ignores node transforms (assets/mod.rs:218)."""
    doc, blob = split_glb(data)
    buffers = _buffers(doc, blob)
    verts, inds, geos = [], [], []
    mesh_first_geo, mesh_geo_count = {}, {}
    n_v = n_i = 0
    materials = doc.get("materials", [])
    for mi, mesh in enumerate(doc.get("meshes", [])):
        mesh_first_geo[mi] = len(geos)
        for prim in mesh.get("primitives", []):
            if prim.get("mode", 4) != 4:
                continue                                           # points / lines / strips have no triangles for the BLAS
            att = prim["attributes"]
            p = read_accessor(doc, buffers, att["POSITION"], True).astype(np.float32)
            idx = (read_accessor(doc, buffers, prim["indices"]).reshape(-1).astype(np.uint32) if "indices" in prim
                   else np.arange(len(p), dtype=np.uint32))
            idx = idx[:len(idx) - len(idx) % 3]
            if len(idx) and idx.max() >= len(p):
                raise ValueError("glTF: index out of range")
            nrm = read_accessor(doc, buffers, att["NORMAL"], True).astype(np.float32) if "NORMAL" in att else _vertex_normals(p, idx)
            uv = read_accessor(doc, buffers, att["TEXCOORD_0"], True).astype(np.float32) if "TEXCOORD_0" in att else np.zeros((len(p), 2), np.float32)
            v = np.zeros(len(p), VERTEX)
            v["p"], v["n"], v["t"] = p[:, :3], nrm[:, :3], uv[:, :2]
            g = np.zeros((), GEOMETRY)
            mat = materials[prim["material"]] if "material" in prim else {}
            pbr = mat.get("pbrMetallicRoughness", {})
            g["base_color"] = pbr.get("baseColorFactor", [1, 1, 1, 1])
            g["metallic"] = pbr.get("metallicFactor", 1.0)
            g["roughness"] = pbr.get("roughnessFactor", 1.0)
            g["base_color_tex"] = pbr["baseColorTexture"]["index"] if "baseColorTexture" in pbr else -1
            g["emission"] = list(mat.get("emissiveFactor", [0, 0, 0])) + [0.0]
            g["index_offset"], g["vertex_offset"], g["index_count"] = n_i, n_v, len(idx)
            g["xform"] = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
            verts.append(v); inds.append(idx); geos.append(g)
            n_v += len(v); n_i += len(idx)
            if first_primitive_only:
                break
        mesh_geo_count[mi] = len(geos) - mesh_first_geo[mi]
        if first_primitive_only:
            break
    instances = []
    if not first_primitive_only:
        nodes = doc.get("nodes", [])
        scenes_ = doc.get("scenes", [])
        roots = scenes_[doc.get("scene", 0)].get("nodes", []) if scenes_ else list(range(len(nodes)))
        stack = [(r, np.eye(4)) for r in reversed(roots)]
        seen = 0
        while stack:
            ni, parent = stack.pop()
            seen += 1
            if seen > 4 * len(nodes) + 4:
                raise ValueError("glTF: node graph has a cycle")
            node = nodes[ni]
            world = parent @ node_local_matrix(node)
            if "mesh" in node and mesh_geo_count.get(node["mesh"], 0):
                instances.append((mesh_first_geo[node["mesh"]], mesh_geo_count[node["mesh"]], world[:3, :].astype(np.float32).reshape(12)))
            for ch in reversed(node.get("children", [])):
                stack.append((ch, world))
    textures = {}
    for ti, tex in enumerate(doc.get("textures", [])):
        if "source" not in tex:
            continue
        img = doc["images"][tex["source"]]
        if "bufferView" in img:
            bv = doc["bufferViews"][img["bufferView"]]
            raw = buffers[bv["buffer"]][bv.get("byteOffset", 0):bv.get("byteOffset", 0) + bv["byteLength"]]
        elif img.get("uri", "").startswith("data:"):
            import base64
            raw = base64.b64decode(img["uri"].split(",", 1)[1])
        else:
            continue
        textures[ti] = decode_image(raw)
    return {"name": name, "vertices": np.concatenate(verts) if verts else np.zeros(0, VERTEX),
            "indices": np.concatenate(inds) if inds else np.zeros(0, np.uint32),
            "geometries": np.array(geos, GEOMETRY).reshape(-1) if geos else np.zeros(0, GEOMETRY),
            "instances": instances, "textures": textures, "environment": None, "cameras": {}}


def write_glb(scene):
    """Scene dictionary -> GLB bytes (one mesh per geometry, one node per (instance, geometry) whose matrix is instance x geometry
    transform, base-colour textures as embedded PNG).  Used to make test assets: sponza_scene.glb is absent from the reference
    mount (.MISSING_LARGE_BLOBS), so GLB coverage rests on round trips of the generated scenes through this writer."""
    blob = bytearray()
    views, accessors, meshes, materials, nodes, images, textures = [], [], [], [], [], [], []

    def add_view(b, target=None):
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(b)}
        if target:
            v["target"] = target
        blob.extend(b)
        views.append(v)
        return len(views) - 1

    def add_accessor(arr, ctype, typ, target=None, minmax=False):
        a = {"bufferView": add_view(arr.tobytes(), target), "componentType": ctype, "count": len(arr), "type": typ}
        if minmax:
            a["min"], a["max"] = [float(x) for x in arr.min(0)], [float(x) for x in arr.max(0)]
        accessors.append(a)
        return len(accessors) - 1

    tex_slots = sorted((scene.get("textures") or {}).keys())
    for slot in range((max(tex_slots) + 1) if tex_slots else 0):       # texture index must equal the slot number
        img = scene["textures"].get(slot)
        if img is None:
            img = np.full((1, 1, 4), 255, np.uint8)
        images.append({"bufferView": add_view(png_encode(img)), "mimeType": "image/png"})
        textures.append({"source": len(images) - 1})
    V, I = scene["vertices"], scene["indices"]
    for g in scene["geometries"]:
        idx = np.ascontiguousarray(I[g["index_offset"]:g["index_offset"] + g["index_count"]], np.uint32)
        used = int(idx.max()) + 1 if len(idx) else 0
        v = V[g["vertex_offset"]:g["vertex_offset"] + used]
        att = {"POSITION": add_accessor(np.ascontiguousarray(v["p"], "<f4"), 5126, "VEC3", 34962, True),
               "NORMAL": add_accessor(np.ascontiguousarray(v["n"], "<f4"), 5126, "VEC3", 34962),
               "TEXCOORD_0": add_accessor(np.ascontiguousarray(v["t"], "<f4"), 5126, "VEC2", 34962)}
        pbr = {"baseColorFactor": [float(x) for x in g["base_color"]], "metallicFactor": float(g["metallic"]), "roughnessFactor": float(g["roughness"])}
        if int(g["base_color_tex"]) >= 0:
            pbr["baseColorTexture"] = {"index": int(g["base_color_tex"])}
        materials.append({"pbrMetallicRoughness": pbr, "emissiveFactor": [float(x) for x in g["emission"][:3]]})
        meshes.append({"primitives": [{"attributes": att, "indices": add_accessor(idx, 5125, "SCALAR", 34963), "material": len(materials) - 1}]})
    ident = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    inst = scene.get("instances") or [(0, len(scene["geometries"]), ident)]
    for first, count, xf in inst:
        mi = np.vstack([np.asarray(xf, np.float64).reshape(3, 4), [0, 0, 0, 1]])
        for gi in range(first, first + count):
            mg = np.vstack([np.asarray(scene["geometries"][gi]["xform"], np.float64).reshape(3, 4), [0, 0, 0, 1]])
            m = (mi @ mg).astype(np.float32)
            node = {"mesh": gi}
            if not np.array_equal(m, np.eye(4, dtype=np.float32)):
                node["matrix"] = [float(x) for x in m.T.reshape(-1)]       # column-major
            nodes.append(node)
    doc = {"asset": {"version": "2.0", "generator": "rt3-b200 assets.write_glb"}, "scene": 0, "scenes": [{"nodes": list(range(len(nodes)))}],
           "nodes": nodes, "meshes": meshes, "materials": materials, "accessors": accessors, "bufferViews": views,
           "buffers": [{"byteLength": len(blob)}]}
    if images:
        doc["images"], doc["textures"] = images, textures
    js = json.dumps(doc, separators=(",", ":")).encode("utf-8")
    js += b" " * (-len(js) % 4)
    while len(blob) % 4:
        blob.append(0)
    total = 12 + 8 + len(js) + 8 + len(blob)
    return struct.pack("<III", _GLB_MAGIC, 2, total) + struct.pack("<II", len(js), _CHUNK_JSON) + js + struct.pack("<II", len(blob), _CHUNK_BIN) + bytes(blob)


def import_glb_to_mesh(data):
    """The reference's whole asset pipeline for one file: GltfMeshLoader -> MeshTransformer -> MeshSaver
    (LoadTransformAndSave, assets/mod.rs:19-23).  Returns the bytes of the processed `.mesh`."""
    sc = load_glb(data, first_primitive_only=True)
    if not len(sc["geometries"]):
        raise ValueError("GLB has no triangle primitive")
    g = sc["geometries"][0]
    tex = None if int(g["base_color_tex"]) < 0 else int(g["base_color_tex"])
    mat = material_record(tuple(float(x) for x in g["base_color"][:3]), float(g["metallic"]), float(g["roughness"]), tex)
    return write_mesh(process_primitive(sc["vertices"], sc["indices"], mat))


# ---------------------------------------------------------------------------------------------------------------- OpenEXR
# resources/skybox2.exr (the environment light BASELINE.json names; absent from the mount) is an OpenEXR image, read by the reference
# through the `image` / `exr 1.73.0` crates (Cargo.lock:1448-1449; only in a commented-out line, main.rs:94).  Restated from the
# published file layout (openexr.com "OpenEXR File Layout"): single-part scanline images, channels HALF / FLOAT / UINT with any
# sampling of 1, compression NONE / RLE / ZIPS / ZIP.  The lossy and wavelet codecs (PIZ, PXR24, B44, DWA) are not implemented and
# fail loudly.  Parity unpinned: no EXR file ships with the reference; the tests pin the reader against a byte-level hand assembly of
# the layout and against the writer below.
EXR_MAGIC = 20000630
_EXR_LINES_PER_BLOCK = {0: 1, 1: 1, 2: 1, 3: 16}           # NONE, RLE, ZIPS, ZIP
_EXR_COMPRESSION_NAMES = {0: "NONE", 1: "RLE", 2: "ZIPS", 3: "ZIP", 4: "PIZ", 5: "PXR24", 6: "B44", 7: "B44A", 8: "DWAA", 9: "DWAB"}
_EXR_PIXEL = {0: np.dtype("<u4"), 1: np.dtype("<f2"), 2: np.dtype("<f4")}


def _exr_unfilter(buf):
    """Inverse of the byte predictor + even/odd split both RLE and ZIP blocks go through."""
    b = np.frombuffer(buf, np.uint8).astype(np.int64)
    if len(b) == 0:
        return b""
    b[1:] -= 128
    b = (np.cumsum(b) & 0xFF).astype(np.uint8)
    half = (len(b) + 1) // 2
    out = np.empty(len(b), np.uint8)
    out[0::2] = b[:half]
    out[1::2] = b[half:]
    return out.tobytes()


def _exr_filter(raw):
    b = np.frombuffer(raw, np.uint8)
    t = np.concatenate([b[0::2], b[1::2]]).astype(np.int64)
    d = t.copy()
    d[1:] = (t[1:] - t[:-1] + 128 + 256) & 0xFF
    return d.astype(np.uint8).tobytes()


def _exr_rle_decode(data, expect):
    out, i = bytearray(), 0
    while i < len(data):
        c = data[i] - 256 if data[i] > 127 else data[i]
        i += 1
        if c < 0:
            out += data[i:i - c]
            i += -c
        else:
            out += bytes([data[i]]) * (c + 1)
            i += 1
    if len(out) != expect:
        raise ValueError("EXR: RLE block decodes to the wrong size")
    return bytes(out)


def _exr_rle_encode(raw):
    out, i, n = bytearray(), 0, len(raw)
    while i < n:
        j = i + 1
        while j < n and raw[j] == raw[i] and j - i < 128:
            j += 1
        if j - i >= 3:
            out += bytes([j - i - 1, raw[i]])
            i = j
        else:
            k = i
            while k < n and k - i < 127 and not (k + 2 < n and raw[k] == raw[k + 1] == raw[k + 2]):
                k += 1
            out += bytes([(256 - (k - i)) & 0xFF]) + raw[i:k]
            i = k
    return bytes(out)


def read_exr(data):
    """-> (pixels float32 [H, W, C], channel names in file order).  Scanline OpenEXR, see the section comment for the subset."""
    if len(data) < 8 or struct.unpack_from("<i", data, 0)[0] != EXR_MAGIC:
        raise ValueError("not an OpenEXR file")
    version = struct.unpack_from("<i", data, 4)[0]
    if version & 0xFF != 2 or version & (0x200 | 0x800 | 0x1000):
        raise NotImplementedError("EXR: only version-2 single-part scanline images are supported")
    o, attrs = 8, {}

    def cstr(o):
        e = data.index(b"\0", o)
        return data[o:e].decode("latin-1"), e + 1
    while data[o] != 0:
        name, o = cstr(o)
        typ, o = cstr(o)
        size = struct.unpack_from("<i", data, o)[0]
        attrs[name] = (typ, data[o + 4:o + 4 + size])
        o += 4 + size
    o += 1
    for need in ("channels", "compression", "dataWindow"):
        if need not in attrs:
            raise ValueError(f"EXR: header lacks '{need}'")
    channels, c = [], 0
    cl = attrs["channels"][1]
    while cl[c] != 0:
        e = cl.index(b"\0", c)
        name = cl[c:e].decode("latin-1")
        ptype, _plinear, xs, ys = struct.unpack_from("<iB3xii", cl, e + 1)
        if xs != 1 or ys != 1:
            raise NotImplementedError("EXR: subsampled channels are not supported")
        channels.append((name, _EXR_PIXEL[ptype]))
        c = e + 1 + 16
    comp = attrs["compression"][1][0]
    if comp not in _EXR_LINES_PER_BLOCK:
        raise NotImplementedError(f"EXR: compression {_EXR_COMPRESSION_NAMES.get(comp, comp)} is not implemented (NONE, RLE, ZIPS, ZIP are)")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    lines = _EXR_LINES_PER_BLOCK[comp]
    n_blocks = -(-H // lines)
    offsets = struct.unpack_from(f"<{n_blocks}Q", data, o)
    row_bytes = sum(W * dt.itemsize for _, dt in channels)
    out = np.zeros((H, W, len(channels)), np.float32)
    for off in offsets:
        y, size = struct.unpack_from("<ii", data, off)
        body = data[off + 8:off + 8 + size]
        nl = min(lines, y1 - y + 1)
        expect = nl * row_bytes
        if size < expect:                                   # a block that did not shrink is stored raw
            raw = zlib.decompress(body) if comp in (2, 3) else _exr_rle_decode(body, expect)
            raw = _exr_unfilter(raw)
        else:
            raw = body
        if len(raw) != expect:
            raise ValueError("EXR: block decodes to the wrong size")
        p = 0
        for ln in range(nl):
            for ci, (_, dt) in enumerate(channels):
                out[y - y0 + ln, :, ci] = np.frombuffer(raw, dt, W, p).astype(np.float32)
                p += W * dt.itemsize
    return out, [n for n, _ in channels]


def write_exr(rgb, compression="ZIP", half=False, data_window_origin=(0, 0)):
    """float [H, W, 3] -> OpenEXR bytes, channels B, G, R (alphabetical, as the format requires), FLOAT or HALF."""
    comp = {v: k for k, v in _EXR_COMPRESSION_NAMES.items()}[compression]
    rgb = np.asarray(rgb)
    H, W = rgb.shape[:2]
    dt = np.dtype("<f2") if half else np.dtype("<f4")
    x0, y0 = data_window_origin

    def attr(name, typ, value):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(value)) + value
    chl = b"".join(n.encode() + b"\0" + struct.pack("<iB3xii", 1 if half else 2, 0, 1, 1) for n in "BGR") + b"\0"
    box = struct.pack("<4i", x0, y0, x0 + W - 1, y0 + H - 1)
    hdr = (struct.pack("<ii", EXR_MAGIC, 2) + attr("channels", "chlist", chl) + attr("compression", "compression", bytes([comp]))
           + attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0")
           + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0))
           + attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0")
    lines = _EXR_LINES_PER_BLOCK[comp]
    blocks = []
    for b0 in range(0, H, lines):
        raw = b"".join(np.ascontiguousarray(rgb[y, :, ch]).astype(dt).tobytes() for y in range(b0, min(H, b0 + lines)) for ch in (2, 1, 0))
        body = raw
        if comp:
            f = _exr_filter(raw)
            enc = zlib.compress(f, 6) if comp in (2, 3) else _exr_rle_encode(f)
            if len(enc) < len(raw):
                body = enc
        blocks.append(struct.pack("<ii", y0 + b0, len(body)) + body)
    table_at = len(hdr)
    pos, table = table_at + 8 * len(blocks), []
    for b in blocks:
        table.append(pos)
        pos += len(b)
    return hdr + struct.pack(f"<{len(blocks)}Q", *table) + b"".join(blocks)


def load_environment_exr(data):
    """skybox2.exr role: equirect RGB float32 [H, W, 3] for rt3_scene_set_environment (missing colour channels read as 0)."""
    px, names = read_exr(data)
    out = np.zeros(px.shape[:2] + (3,), np.float32)
    for k, n in enumerate("RGB"):
        if n in names:
            out[..., k] = px[..., names.index(n)]
        elif "Y" in names:
            out[..., k] = px[..., names.index("Y")]
    return out
