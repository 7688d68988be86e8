#!/usr/bin/env python3
"""Generate the scene fixtures under tests/golden/scenes/ from the reference's bundled models.

Runs ONLY in the build container (needs /root/reference and oracle/_ref/libdrt_ref.so); the outputs
are committed so that tests, smoke() and bench.py never read /root/reference at run time.

For every scene in SCENES the reference GLB is re-muxed into a new GLB:
  * JSON is rewritten with only the keys the path consumes (nodes, meshes, materials, textures,
    images, accessors, bufferViews) -- extras/extensions/cameras/animations are dropped;
  * geometry bufferViews are copied byte-for-byte (they ARE the scene), re-packed 4-byte aligned;
  * every image is decoded by the REFERENCE's decoder (stb_image 2.29 through libdrt_ref.so, the
    same call the reference makes: Texture.cu:21-30) and re-encoded as an 8-bit PNG with exactly the
    channel count stb reported, so that any conforming PNG decoder reproduces the reference's texels
    bit-for-bit (this matters for JPEG sources, where decoders legitimately differ by +-1 LSB).
The script then verifies, through the reference loader itself, that the fixture loads to a triangle
soup / material table / texel set identical to the original file's.
"""
import ctypes
import json
import os
import struct
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_MODELS = "/root/reference/models"
OUT_DIR = os.path.join(ROOT, "tests", "golden", "scenes")

SCENES = {
    # fixture name          : reference model (BASELINE.json configs + the reference's own test scenes)
    "cornell_box.glb": "cornell_box.glb",
    "dense_monkey.glb": "dense_monkey.glb",
    "room.glb": "room.glb",
    "valorant_ghost.glb": "valorant_ghost.glb",
    "animemaid.glb": "animemaid.glb",
    "bvhsplit_test.glb": "test/bvhsplitTest.glb",
    "mc_transparency_test.glb": "minecraft/mcTransparencyTest.glb",
    "cube.glb": "cube.glb",
}


def load_ref():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libdrt_ref.so"))
    lib.ref_scene_load.restype = ctypes.c_void_p
    lib.ref_scene_load.argtypes = [ctypes.c_char_p]
    for fn in ("ref_scene_tri_count", "ref_scene_material_count", "ref_scene_texture_count"):
        getattr(lib, fn).argtypes = [ctypes.c_void_p]
    lib.ref_scene_get_triangles.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.ref_scene_get_materials.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.ref_scene_get_texture_info.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.POINTER(ctypes.c_int)] * 3
    lib.ref_scene_get_texture_pixels.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.ref_scene_free.argtypes = [ctypes.c_void_p]
    return lib


def ref_scene_dump(lib, path):
    """(triangles [N,32] f32 view without padding, materials bytes, [(w,h,c,pixels)])"""
    h = ctypes.c_void_p(lib.ref_scene_load(path.encode()))
    n = lib.ref_scene_tri_count(h)
    tris = np.zeros((n, 32), np.float32)
    lib.ref_scene_get_triangles(h, tris.ctypes.data_as(ctypes.c_void_p))
    tris[:, 3] = 0  # 4 padding bytes after centroid are uninitialised in the reference struct
    m = lib.ref_scene_material_count(h)
    mats = np.zeros((m, 11), np.uint32)
    lib.ref_scene_get_materials(h, mats.ctypes.data_as(ctypes.c_void_p))
    mats[:, 8] &= 0xFF   # bool Transmission + 3 padding bytes
    mats[:, 10] &= 0xFF  # bool Metallic + 3 padding bytes
    texs = []
    for i in range(lib.ref_scene_texture_count(h)):
        w, hh, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        lib.ref_scene_get_texture_info(h, i, ctypes.byref(w), ctypes.byref(hh), ctypes.byref(c))
        px = np.zeros((hh.value, w.value, c.value), np.uint8)
        lib.ref_scene_get_texture_pixels(h, i, px.ctypes.data_as(ctypes.c_void_p))
        texs.append(px)
    lib.ref_scene_free(h)
    return tris, mats, texs


def png_encode(px):
    """Minimal PNG writer (8-bit, non-interlaced, filter 0 per row, color type by channel count)."""
    h, w, c = px.shape
    color_type = {1: 0, 2: 4, 3: 2, 4: 6}[c]

    def chunk(tag, data):
        body = tag + data
        return struct.pack(">I", len(data)) + body + struct.pack(">I", zlib.crc32(body) & 0xFFFFFFFF)

    raw = np.zeros((h, 1 + w * c), np.uint8)
    raw[:, 1:] = px.reshape(h, w * c)
    # Paeth/sub filtering would compress better; use adaptive "sub" vs "none" per image for size.
    sub = raw.copy()
    sub[:, 0] = 1
    flat = px.reshape(h, w * c).astype(np.int16)
    sub[:, 1 + c:] = ((flat[:, c:] - flat[:, :-c]) & 0xFF).astype(np.uint8)
    sub[:, 1:1 + c] = flat[:, :c].astype(np.uint8)
    a = zlib.compress(raw.tobytes(), 9)
    b = zlib.compress(sub.tobytes(), 9)
    idat = a if len(a) <= len(b) else b
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color_type, 0, 0, 0))
            + chunk(b"IDAT", idat) + chunk(b"IEND", b""))


def read_glb(path):
    d = open(path, "rb").read()
    magic, ver, total = struct.unpack("<III", d[:12])
    assert magic == 0x46546C67 and ver == 2
    jlen, jtype = struct.unpack("<II", d[12:20])
    assert jtype == 0x4E4F534A
    j = json.loads(d[20:20 + jlen])
    off = 20 + jlen
    blen, btype = struct.unpack("<II", d[off:off + 8])
    assert btype == 0x004E4942
    return j, d[off + 8:off + 8 + blen]


def write_glb(path, j, binbuf):
    js = json.dumps(j, separators=(",", ":")).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    binbuf += b"\0" * ((4 - len(binbuf) % 4) % 4)
    total = 12 + 8 + len(js) + 8 + len(binbuf)
    with open(path, "wb") as f:
        f.write(struct.pack("<III", 0x46546C67, 2, total))
        f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
        f.write(struct.pack("<II", len(binbuf), 0x004E4942) + binbuf)


def keep(d, keys):
    return {k: d[k] for k in keys if k in d}


def remux(lib, src, dst):
    j, binbuf = read_glb(src)
    assert len(j["buffers"]) == 1
    _, _, texs = ref_scene_dump(lib, src)
    image_views = {img["bufferView"]: i for i, img in enumerate(j.get("images", []))}
    new_bin = bytearray()
    new_views = []
    for vi, bv in enumerate(j["bufferViews"]):
        if vi in image_views:
            data = png_encode(texs[image_views[vi]])
        else:
            o = bv.get("byteOffset", 0)
            data = binbuf[o:o + bv["byteLength"]]
        new_bin += b"\0" * ((4 - len(new_bin) % 4) % 4)
        nv = keep(bv, ["byteStride", "target"])
        nv.update({"buffer": 0, "byteOffset": len(new_bin), "byteLength": len(data)})
        new_views.append(nv)
        new_bin += data
    out = {
        "asset": {"version": "2.0", "generator": "dustraytracer_b200 tools/make_scene_fixtures.py"},
        "scene": j.get("scene", 0),
        "scenes": [keep(s, ["nodes", "name"]) for s in j.get("scenes", [])],
        "nodes": [keep(n, ["mesh", "name", "children", "translation", "rotation", "scale", "matrix"]) for n in j["nodes"]],
        "meshes": [{"name": m.get("name", ""),
                    "primitives": [keep(p, ["attributes", "indices", "material", "mode"]) for p in m["primitives"]]}
                   for m in j["meshes"]],
        "accessors": [keep(a, ["bufferView", "byteOffset", "componentType", "count", "type", "min", "max", "normalized"])
                      for a in j["accessors"]],
        "bufferViews": new_views,
        "buffers": [{"byteLength": len(new_bin) + ((4 - len(new_bin) % 4) % 4)}],
    }
    if "materials" in j:
        out["materials"] = [keep(m, ["name", "pbrMetallicRoughness", "emissiveFactor", "alphaMode", "alphaCutoff", "doubleSided"])
                            for m in j["materials"]]
    if "textures" in j:
        out["textures"] = [keep(t, ["sampler", "source"]) for t in j["textures"]]
    if "samplers" in j:
        out["samplers"] = [keep(s, ["magFilter", "minFilter", "wrapS", "wrapT"]) for s in j["samplers"]]
    if "images" in j:
        out["images"] = [{"bufferView": img["bufferView"], "mimeType": "image/png", "name": img.get("name", "")}
                         for img in j["images"]]
    write_glb(dst, out, bytes(new_bin))


def main():
    lib = load_ref()
    os.makedirs(OUT_DIR, exist_ok=True)
    manifest = {}
    for name, rel in SCENES.items():
        src = os.path.join(REF_MODELS, rel)
        dst = os.path.join(OUT_DIR, name)
        remux(lib, src, dst)
        a = ref_scene_dump(lib, src)
        b = ref_scene_dump(lib, dst)
        assert a[0].shape == b[0].shape and np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)), name
        assert np.array_equal(a[1], b[1]), name
        assert len(a[2]) == len(b[2]) and all(np.array_equal(x, y) for x, y in zip(a[2], b[2])), name
        manifest[name] = {
            "source": "models/" + rel,
            "triangles": int(a[0].shape[0]),
            "materials": int(a[1].shape[0]),
            "textures": [[int(t.shape[1]), int(t.shape[0]), int(t.shape[2])] for t in a[2]],
            "bytes": os.path.getsize(dst),
            "tri_crc32": zlib.crc32(a[0].tobytes()) & 0xFFFFFFFF,
        }
        print(name, manifest[name])
    with open(os.path.join(OUT_DIR, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    sys.exit(main())
