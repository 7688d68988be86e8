"""DRT_LOAD_ROBUST: the spec-conforming glTF ingest, checked through the C ABI against a
numpy model of what the glTF 2.0 specification says a file means. Host-only: no GPU needed.
The reference-compatible mode (load flags 0) must stay exactly what it was: the same files are loaded both ways
where the reference loader can read them at all."""
import json
import os
import struct

import numpy as np
import pytest

import oracles as O
import dustraytracer_b200 as drt
from test_abi import _png

COMP = {np.dtype(np.float32): 5126, np.dtype(np.uint32): 5125, np.dtype(np.uint16): 5123, np.dtype(np.int16): 5122,
        np.dtype(np.uint8): 5121, np.dtype(np.int8): 5120}
TYPE = {1: "SCALAR", 2: "VEC2", 3: "VEC3", 4: "VEC4"}


class Glb:
    """Minimal GLB writer: accessors with arbitrary byteOffset / byteStride / buffer."""

    def __init__(self):
        self.j = {"asset": {"version": "2.0"}, "accessors": [], "bufferViews": [], "buffers": [], "meshes": [], "nodes": [],
                  "materials": [], "textures": [], "images": []}
        self.bin = bytearray()

    def view(self, data, stride=0, pad_front=0):
        while len(self.bin) % 4:
            self.bin.append(0xEE)
        self.bin += b"\xEE" * pad_front
        off = len(self.bin)
        self.bin += data
        v = {"buffer": 0, "byteOffset": off, "byteLength": len(data)}
        if stride:
            v["byteStride"] = stride
        self.j["bufferViews"].append(v)
        return len(self.j["bufferViews"]) - 1

    def accessor(self, arr, stride=0, byte_offset=0, normalized=False, pad_front=0):
        arr = np.ascontiguousarray(arr)
        n, k = (arr.shape[0], 1) if arr.ndim == 1 else arr.shape
        elem = arr.dtype.itemsize * k
        if stride:
            raw = bytearray(b"\xEE" * (byte_offset + stride * n))
            for i in range(n):
                raw[byte_offset + i * stride: byte_offset + i * stride + elem] = arr[i].tobytes()
        else:
            raw = b"\xEE" * byte_offset + arr.tobytes()
        a = {"bufferView": self.view(bytes(raw), stride, pad_front), "componentType": COMP[arr.dtype], "count": n, "type": TYPE[k]}
        if byte_offset:
            a["byteOffset"] = byte_offset
        if normalized:
            a["normalized"] = True
        self.j["accessors"].append(a)
        return len(self.j["accessors"]) - 1

    def image(self, png_bytes):
        self.j["images"].append({"bufferView": self.view(png_bytes), "mimeType": "image/png"})
        return len(self.j["images"]) - 1

    def write(self, path):
        j = {k: v for k, v in self.j.items() if v or k == "asset"}
        j["buffers"] = [{"byteLength": len(self.bin)}]
        js = json.dumps(j).encode()
        js += b" " * ((4 - len(js) % 4) % 4)
        b = bytes(self.bin) + b"\0" * ((4 - len(self.bin) % 4) % 4)
        with open(path, "wb") as f:
            f.write(struct.pack("<III", 0x46546C67, 2, 28 + len(js) + len(b)) + struct.pack("<II", len(js), 0x4E4F534A) + js +
                    struct.pack("<II", len(b), 0x004E4942) + b)
        return str(path)


def load(path, robust=True):
    s = drt.Scene()
    s.loadGLTFmodel(path, robust=robust)
    return s


def quat_matrix(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], np.float64)


def trs(t=(0, 0, 0), q=(0, 0, 0, 1), s=(1, 1, 1)):
    M = np.eye(4)
    M[:3, :3] = quat_matrix(q) @ np.diag(s)
    M[:3, 3] = t
    return M


def expected_soup(pos, nrm, uv, idx, M):
    """[T, 3, 3] positions, [T, 3, 3] normals, [T, 3, 2] uvs after the world transform M (float64 math, cast last)."""
    p = (pos.astype(np.float64) @ M[:3, :3].T + M[:3, 3])
    nm = np.linalg.inv(M[:3, :3]).T * np.linalg.det(M[:3, :3])   # cofactor matrix
    n = nrm.astype(np.float64) @ nm.T
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    tri = idx.reshape(-1, 3)
    return p[tri].astype(np.float32), n[tri].astype(np.float32), uv[tri].astype(np.float32)


def soup_of(scene):
    t = scene.triangles()
    v = t[:, 4:28].reshape(-1, 3, 8)
    return v[:, :, 0:3], v[:, :, 3:6], v[:, :, 6:8], t[:, 28:31], t[:, 31].view(np.int32)


def quad_mesh(rng, n=5):
    """A small grid mesh with distinct vertices: positions, unit normals, uvs, triangle indices."""
    pos = rng.uniform(-2, 2, (n * n, 3)).astype(np.float32)
    nrm = rng.normal(size=(n * n, 3))
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    uv = rng.uniform(0, 1, (n * n, 2)).astype(np.float32)
    idx = []
    for y in range(n - 1):
        for x in range(n - 1):
            a = y * n + x
            idx += [a, a + 1, a + n, a + 1, a + n + 1, a + n]
    return pos, nrm, uv, np.array(idx, np.uint32)


def test_hierarchy_transforms_strides_offsets_and_index_types(tmp_path):
    rng = np.random.RandomState(11)
    pos, nrm, uv, idx = quad_mesh(rng)
    for itype in (np.uint8, np.uint16, np.uint32):
        g = Glb()
        # interleaved vertex buffer: position at +4, normal at +16, uv at +28 of a 40-byte stride
        inter = bytearray(b"\xEE" * (40 * len(pos)))
        for i in range(len(pos)):
            inter[40 * i + 4: 40 * i + 16] = pos[i].tobytes()
            inter[40 * i + 16: 40 * i + 28] = nrm[i].tobytes()
            inter[40 * i + 28: 40 * i + 36] = uv[i].tobytes()
        bv = g.view(bytes(inter), stride=40, pad_front=8)
        for off, ty in ((4, "VEC3"), (16, "VEC3"), (28, "VEC2")):
            g.j["accessors"].append({"bufferView": bv, "byteOffset": off, "componentType": 5126, "count": len(pos), "type": ty})
        ia = g.accessor(idx.astype(itype), byte_offset=4 if itype != np.uint8 else 3)
        g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": ia, "material": 0}]}]
        g.j["materials"] = [{"pbrMetallicRoughness": {"baseColorFactor": [0.1, 0.2, 0.3, 1.0]}}]
        # root (TRS) -> child (matrix, mirrored) -> grandchild (mesh again); plus an unreferenced node that must NOT be drawn
        Mroot = trs((1, 2, 3), (0.0, 0.38268343, 0.0, 0.92387953), (2, 1, 0.5))
        Mchild = np.array([[-1, 0, 0, 0.5], [0, 1, 0, 0], [0.2, 0, 1, -4], [0, 0, 0, 1]], np.float64)
        Mgrand = trs((0, -1, 0), (0.5, 0.5, 0.5, 0.5))
        g.j["nodes"] = [{"mesh": 0, "translation": [1, 2, 3], "rotation": [0.0, 0.38268343, 0.0, 0.92387953], "scale": [2, 1, 0.5], "children": [1]},
                        {"mesh": 0, "matrix": Mchild.T.reshape(-1).tolist(), "children": [2]},
                        {"mesh": 0, "translation": [0, -1, 0], "rotation": [0.5, 0.5, 0.5, 0.5]},
                        {"mesh": 0, "translation": [100, 100, 100]}]
        g.j["scenes"] = [{"nodes": [3]}, {"nodes": [0]}]
        g.j["scene"] = 1
        s = load(g.write(tmp_path / "h.glb"))
        P, N, T, FN, MAT = soup_of(s)
        nt = len(idx) // 3
        assert P.shape[0] == 3 * nt and s.stats()["meshes"] == 3
        for k, M in enumerate((Mroot, Mroot @ Mchild, Mroot @ Mchild @ Mgrand)):
            eP, eN, eT = expected_soup(pos, nrm, uv, idx, M)
            sl = slice(k * nt, (k + 1) * nt)
            assert np.allclose(P[sl], eP, rtol=0, atol=2e-6 * np.abs(eP).max())
            assert np.allclose(N[sl], eN, atol=2e-6)
            assert np.array_equal(T[sl], eT)
            # the face normal is the geometric normal, oriented to the mean (transformed) vertex normal
            gn = np.cross(eP[:, 1] - eP[:, 0], eP[:, 2] - eP[:, 0]).astype(np.float64)
            gn /= np.linalg.norm(gn, axis=1, keepdims=True)
            sign = np.sign(np.einsum("ij,ij->i", gn, eN.astype(np.float64).mean(1)))
            assert np.allclose(FN[sl], gn * sign[:, None], atol=1e-4)
        assert np.all(MAT == 0)


def test_non_indexed_strips_fans_and_missing_attributes(tmp_path):
    rng = np.random.RandomState(5)
    pos = rng.uniform(-1, 1, (9, 3)).astype(np.float32)
    g = Glb()
    pa = g.accessor(pos)
    strip = np.array([0, 1, 2, 3, 4, 5], np.uint16)
    fan = np.array([6, 0, 1, 2, 3], np.uint16)
    sa, fa = g.accessor(strip), g.accessor(fan)
    g.j["meshes"] = [{"primitives": [
        {"attributes": {"POSITION": pa}},                                  # non-indexed TRIANGLES: 3 triangles
        {"attributes": {"POSITION": pa}, "indices": sa, "mode": 5},        # strip: 4 triangles
        {"attributes": {"POSITION": pa}, "indices": fa, "mode": 6},        # fan: 3 triangles
        {"attributes": {"POSITION": pa}, "mode": 1},                       # lines: skipped
    ]}]
    g.j["nodes"] = [{"mesh": 0}]
    s = load(g.write(tmp_path / "m.glb"))
    P, N, T, FN, MAT = soup_of(s)
    want = [[0, 1, 2], [3, 4, 5], [6, 7, 8],
            [0, 1, 2], [1, 3, 2], [2, 3, 4], [3, 5, 4],
            [6, 0, 1], [6, 1, 2], [6, 2, 3]]
    assert np.array_equal(P, pos[np.array(want)])
    assert np.all(N == 0) and np.all(T == 0) and np.all(MAT == -1)
    gn = np.cross(P[:, 1] - P[:, 0], P[:, 2] - P[:, 0])
    gn /= np.linalg.norm(gn, axis=1, keepdims=True)
    assert np.allclose(FN, gn, atol=1e-5)          # no vertex normals: the winding decides


def test_normalised_integer_attributes(tmp_path):
    g = Glb()
    qpos = np.array([[0, 0, 0], [32767, 0, 0], [0, -32767, 16384]], np.int16)
    quv = np.array([[0, 255], [128, 0], [255, 255]], np.uint8)
    qn = np.array([[0, 0, 127], [0, 0, 127], [0, 0, 127]], np.int8)
    pa = g.accessor(qpos, stride=8, normalized=True)
    ua = g.accessor(quv, stride=4, normalized=True)
    na = g.accessor(qn, stride=4, normalized=True)
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa, "NORMAL": na, "TEXCOORD_0": ua}}]}]
    g.j["nodes"] = [{"mesh": 0, "scale": [2, 2, 2]}]
    g.j["extensionsUsed"] = ["KHR_mesh_quantization"]
    s = load(g.write(tmp_path / "q.glb"))
    P, N, T, FN, MAT = soup_of(s)
    assert np.allclose(P[0], np.array([[0, 0, 0], [2, 0, 0], [0, -2, 2 * 16384 / 32767]], np.float32))
    assert np.allclose(T[0], quv.astype(np.float32) / 255)
    assert np.allclose(N[0], [[0, 0, 1]] * 3)


def test_texture_indirection_and_alpha_modes(tmp_path):
    rng = np.random.RandomState(2)
    rgba = rng.randint(0, 256, (6, 5, 4)).astype(np.uint8)
    rgb = rng.randint(0, 256, (3, 3, 3)).astype(np.uint8)
    g = Glb()
    pa = g.accessor(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32))
    na = g.accessor(np.array([[0, 0, 1]] * 3, np.float32))
    ua = g.accessor(np.array([[0, 0], [1, 0], [0, 1]], np.float32))
    ia = g.accessor(np.array([0, 1, 2], np.uint16))
    i_rgb, i_rgba = g.image(_png(rgb)), g.image(_png(rgba))
    g.j["textures"] = [{"source": i_rgba}, {"source": i_rgb}]      # texture index != image index
    tex = lambda t: {"baseColorTexture": {"index": t}}
    g.j["materials"] = [
        {"pbrMetallicRoughness": tex(0)},                                               # OPAQUE (default) on an RGBA image
        {"pbrMetallicRoughness": tex(0), "alphaMode": "MASK"},                          # cutoff 0.5
        {"pbrMetallicRoughness": tex(0), "alphaMode": "MASK", "alphaCutoff": 0.25},
        {"pbrMetallicRoughness": tex(0), "alphaMode": "BLEND"},
        {"pbrMetallicRoughness": tex(1), "alphaMode": "MASK"},                          # RGB image: nothing to derive
        {"pbrMetallicRoughness": tex(0), "alphaMode": "MASK", "alphaCutoff": 0.5},      # shares the copy of material 1
        {"pbrMetallicRoughness": {"baseColorFactor": [1, 0, 0, 1]}},
    ]
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa, "NORMAL": na, "TEXCOORD_0": ua}, "indices": ia, "material": m} for m in range(7)]}]
    g.j["nodes"] = [{"mesh": 0}]
    path = g.write(tmp_path / "a.glb")
    s = load(path)
    mats = s.materials()
    ti = [m.albedo_texture_index for m in mats]
    assert s.stats()["textures"] == 5           # 2 images + opaque copy + two thresholds
    assert np.array_equal(s.texture(0), rgb) and np.array_equal(s.texture(1), rgba)
    assert ti[3] == i_rgba and ti[4] == i_rgb and ti[6] == -1 and ti[5] == ti[1] and len({ti[0], ti[1], ti[2]}) == 3
    assert np.array_equal(s.texture(ti[0]), rgba[:, :, :3])
    for m, cutoff in ((1, 0.5), (2, 0.25)):
        t = s.texture(ti[m])
        assert np.array_equal(t[:, :, :3], rgba[:, :, :3])
        assert np.array_equal(t[:, :, 3], np.where(rgba[:, :, 3] / 255.0 >= cutoff, 255, 0))
    # the reference-compatible mode keeps the reference's reading of the same file: texture index used as image index
    c = load(path, robust=False)
    assert [m.albedo_texture_index for m in c.materials()][:2] == [0, 0] and c.stats()["textures"] == 2


def test_robust_and_compat_agree_on_the_reference_assets():
    """The bundled scenes have no transforms / offsets / strides, so both modes must emit the same soup
    (cornell_box and cube have every attribute; room has a primitive without TEXCOORD_0, where the modes differ
    exactly in those UVs)."""
    for name in ("cube", "cornell_box", "room"):
        a, b = load(O.scene_path(name), robust=False), load(O.scene_path(name), robust=True)
        ta, tb = a.triangles(), b.triangles()
        degenerate = np.isnan(ta[:, 28:31]).any(1)   # zero-area triangles (room.glb has 8): NaN face normal in the reference, dropped here
        assert degenerate.sum() == (8 if name == "room" else 0)
        ta = ta[~degenerate]
        assert ta.shape == tb.shape
        uv_cols = [10, 11, 18, 19, 26, 27]
        other = [c for c in range(32) if c not in uv_cols]
        assert np.array_equal(ta[:, other].view(np.uint32), tb[:, other].view(np.uint32))
        same_uv = np.all(ta[:, uv_cols].view(np.uint32) == tb[:, uv_cols].view(np.uint32), axis=1)
        if name == "room":
            assert not same_uv.all() and np.all(tb[~same_uv][:, uv_cols] == 0)
        else:
            assert same_uv.all()


def test_robust_errors_are_reported_not_crashes(tmp_path):
    g = Glb()
    pa = g.accessor(np.zeros((3, 3), np.float32))
    ia = g.accessor(np.array([0, 1, 7], np.uint16))
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa}, "indices": ia}]}]
    g.j["nodes"] = [{"mesh": 0}]
    with pytest.raises(drt.DrtError) as e:
        load(g.write(tmp_path / "oob.glb"))
    assert e.value.code == -3 and "out of range" in str(e.value)
    # a cycle in the node graph
    g.j["meshes"][0]["primitives"][0]["indices"] = g.accessor(np.array([0, 1, 2], np.uint16))
    g.j["nodes"] = [{"children": [1]}, {"children": [0], "mesh": 0}]
    g.j["scenes"] = [{"nodes": [0]}]
    with pytest.raises(drt.DrtError) as e:
        load(g.write(tmp_path / "cycle.glb"))
    assert "cycle" in str(e.value)
    # an accessor that runs past its view
    g.j["nodes"] = [{"mesh": 0}]
    g.j["accessors"][pa]["count"] = 4
    with pytest.raises(drt.DrtError) as e:
        load(g.write(tmp_path / "past.glb"))
    assert "past its bufferView" in str(e.value)
    # required compression extensions are named
    g.j["accessors"][pa]["count"] = 3
    g.j["extensionsRequired"] = ["KHR_draco_mesh_compression"]
    with pytest.raises(drt.DrtError) as e:
        load(g.write(tmp_path / "draco.glb"))
    assert "KHR_draco_mesh_compression" in str(e.value)
    # degenerate (zero-area) triangles are dropped, the rest kept
    del g.j["extensionsRequired"]
    g.j["accessors"] = []
    g.j["bufferViews"] = []
    g.bin = bytearray()
    pa = g.accessor(np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0], [0, 1, 0]], np.float32))
    ia = g.accessor(np.array([0, 1, 2, 0, 1, 3], np.uint16))
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa}, "indices": ia}]}]
    s = load(g.write(tmp_path / "degenerate.glb"))
    assert s.stats()["triangles"] == 1


def test_crafted_byte_ranges_are_rejected_not_wrapped(tmp_path):
    """Offsets / lengths / counts that are negative, fractional-huge or whose SUM wraps size_t must fail with a format error in
    both loaders (a cast of such a double to size_t is undefined; `offset + length > size` wraps)."""
    def base():
        g = Glb()
        pa = g.accessor(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32))
        na = g.accessor(np.array([[0, 0, 1]] * 3, np.float32))
        ta = g.accessor(np.zeros((3, 2), np.float32))
        ia = g.accessor(np.array([0, 1, 2], np.uint16))
        g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa, "NORMAL": na, "TEXCOORD_0": ta}, "indices": ia}]}]
        g.j["nodes"] = [{"mesh": 0}]
        return g, pa, ia

    g, pa, ia = base()
    assert load(g.write(tmp_path / "ok.glb")).stats()["triangles"] == 1 and load(str(tmp_path / "ok.glb"), robust=False).stats()["triangles"] == 1
    cases = []
    for key, val in (("byteOffset", -4), ("byteOffset", 1.8446744073709552e19), ("byteLength", 1.8446744073709552e19), ("byteLength", -1),
                     ("byteOffset", 1e300), ("byteStride", 1e18)):
        g, pa, ia = base()
        g.j["bufferViews"][0][key] = val
        cases.append(("view_%s_%g" % (key, val), g))
    g, pa, ia = base()
    g.j["bufferViews"][0]["byteOffset"] = 2 ** 64 - 8   # offset + length wraps to a small number
    g.j["bufferViews"][0]["byteLength"] = 36
    cases.append(("view_wrap", g))
    for robust in (True, False):
        for name, g in cases:
            with pytest.raises(drt.DrtError) as e:
                load(g.write(tmp_path / (name + ".glb")), robust=robust)
            assert e.value.code == -3, (name, robust, str(e.value))
    for key, val in (("count", -3), ("count", 1e300), ("byteOffset", -12), ("byteOffset", 1.8446744073709552e19), ("count", 2 ** 62)):
        g, pa, ia = base()
        g.j["accessors"][pa][key] = val
        with pytest.raises(drt.DrtError) as e:
            load(g.write(tmp_path / "acc.glb"))
        assert e.value.code == -3, (key, val, str(e.value))


def test_attribute_view_of_a_longer_second_buffer_cannot_index_past_the_first(tmp_path):
    """The reference-compatible loader reads all attributes from the index accessor's buffer (the reference's quirk). A POSITION view
    that is valid in a SECOND, longer buffer but starts beyond the end of the first must be refused there (the remaining-length
    subtraction would wrap and the copy would read far out of bounds); the robust loader follows the accessor's own buffer."""
    import base64
    g = Glb()
    ia = g.accessor(np.array([0, 1, 2], np.uint16))
    na = g.accessor(np.array([[0, 0, 1]] * 3, np.float32))
    ta = g.accessor(np.zeros((3, 2), np.float32))
    second = bytearray(b"\0" * 8192)
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    second[8000:8036] = tri.tobytes()
    g.j["bufferViews"].append({"buffer": 1, "byteOffset": 8000, "byteLength": 36})
    g.j["accessors"].append({"bufferView": len(g.j["bufferViews"]) - 1, "componentType": 5126, "count": 3, "type": "VEC3"})
    pa = len(g.j["accessors"]) - 1
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa, "NORMAL": na, "TEXCOORD_0": ta}, "indices": ia}]}]
    g.j["nodes"] = [{"mesh": 0}]
    path = str(tmp_path / "two_buffers.glb")
    g.write(path)
    # Glb.write emits one buffer: append the second as a data URI
    raw = open(path, "rb").read()
    jlen = struct.unpack_from("<I", raw, 12)[0]
    j = json.loads(raw[20:20 + jlen])
    j["buffers"].append({"byteLength": len(second), "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(second)).decode()})
    js = json.dumps(j).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    rest = raw[20 + jlen:]
    with open(path, "wb") as f:
        f.write(struct.pack("<III", 0x46546C67, 2, 20 + len(js) + len(rest)) + struct.pack("<II", len(js), 0x4E4F534A) + js + rest)
    with pytest.raises(drt.DrtError) as e:
        load(path, robust=False)
    assert e.value.code == -3 and "beyond" in str(e.value)
    s = load(path, robust=True)
    assert s.stats()["triangles"] == 1
    assert np.array_equal(soup_of(s)[0].reshape(3, 3), tri)


@pytest.mark.gpu
def test_robust_scene_renders_like_the_same_soup_given_directly(tmp_path):
    """A transformed, strided, u32-indexed file loaded with DRT_LOAD_ROBUST traces exactly like the world-space
    soup handed to drt_scene_from_triangles (the device never sees the difference)."""
    rng = np.random.RandomState(4)
    pos, nrm, uv, idx = quad_mesh(rng, n=9)
    g = Glb()
    pa, na = g.accessor(pos, stride=20, byte_offset=8), g.accessor(nrm)
    ia = g.accessor(idx, byte_offset=12)
    g.j["meshes"] = [{"primitives": [{"attributes": {"POSITION": pa, "NORMAL": na}, "indices": ia}]}]
    g.j["nodes"] = [{"children": [1], "translation": [0, 2, -3], "rotation": [0.0, 0.38268343, 0.0, 0.92387953]},
                    {"mesh": 0, "scale": [1.5, 1.0, -1.0]}]
    g.j["scenes"] = [{"nodes": [0]}]
    a = load(g.write(tmp_path / "r.glb"))
    P = soup_of(a)[0]
    b = drt.Scene.from_triangles(P)
    hits = []
    for s in (a, b):
        drt.BVHBuilder().buildIterative(s)
        r = drt.Renderer(flags=1 | 2)
        r.ResizeBuffer(160, 120)
        hits.append(r.trace_primary(drt.Camera(), s, 1))
    assert (hits[0]["prim_id"] >= 0).sum() > 500
    for k in ("prim_id", "t", "u", "v"):
        assert np.array_equal(hits[0][k].view(np.uint32), hits[1][k].view(np.uint32))
