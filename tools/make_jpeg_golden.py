#!/usr/bin/env python3
"""Generates the JPEG decoder fixtures: small JPEG files over a matrix of encoder settings (PIL's libjpeg-turbo: baseline /
progressive, 4:4:4 / 4:2:2 / 4:2:0 / grey, restart intervals, odd sizes, qualities; OpenCV for 4:1:1 / 4:4:0 where available) and,
for each, the CRC32 + a few probe bytes of what the REFERENCE's decoder returns for it: stbi_load_from_memory of the stb_image 2.29
the reference vendors, as compiled into oracle/_ref/libdrt_ref.so (built from /root/reference by oracle/Makefile). Also the CRCs
of the JPEG textures embedded in the reference's own models (cube_plane, sceneHierTest, test/EmissiveTest, test/sunshadowTest).
Run in the build container (needs oracle/_ref and /root/reference): python tools/make_jpeg_golden.py"""
import ctypes, io, json, os, struct, sys, zlib
import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "jpeg")
L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libdrt_ref.so"))
L.stbi_load_from_memory.restype = ctypes.POINTER(ctypes.c_ubyte)
L.stbi_load_from_memory.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.c_int]
L.stbi_image_free.argtypes = [ctypes.c_void_p]


def stb_decode(data):
    w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    p = L.stbi_load_from_memory(data, len(data), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), 0)
    if not p:
        return None
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, c.value)).copy()
    L.stbi_image_free(p)
    return a


def test_image(w, h, seed):
    rng = np.random.RandomState(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([127 + 120 * np.sin(x / 5.0 + seed) * np.cos(y / 7.0), 127 + 120 * np.sin((x + y) / 9.0), 255 * ((x // 4 + y // 4) % 2)], -1)
    img += rng.normal(0, 12, img.shape)
    img[h // 3:h // 3 + 5, :, :] = 255  # hard edges
    img[:, w // 2:w // 2 + 2, :] = 0
    return np.clip(img, 0, 255).astype(np.uint8)


def main():
    os.makedirs(OUT, exist_ok=True)
    cases = {}

    def add(name, data):
        ref = stb_decode(data)
        assert ref is not None, name
        open(os.path.join(OUT, name + ".jpg"), "wb").write(data)
        cases[name] = {"shape": list(ref.shape), "crc": zlib.crc32(ref.tobytes()) & 0xFFFFFFFF,
                       "probe": [int(v) for v in ref.reshape(-1)[:: max(1, ref.size // 16)][:16]]}

    sizes = [(64, 48), (33, 17), (1, 1), (8, 8), (17, 40), (250, 131)]
    n = 0
    for (w, h) in sizes:
        img = Image.fromarray(test_image(w, h, n))
        for sub, sname in ((0, "444"), (1, "422"), (2, "420")):
            for prog in (False, True):
                for q in ((90,) if (w, h) != (64, 48) else (25, 75, 98)):
                    buf = io.BytesIO()
                    img.save(buf, "JPEG", quality=q, subsampling=sub, progressive=prog, optimize=(n % 2 == 0))
                    add("c%dx%d_%s_%s_q%d" % (w, h, sname, "prog" if prog else "base", q), buf.getvalue())
                    n += 1
        buf = io.BytesIO()
        img.convert("L").save(buf, "JPEG", quality=85, progressive=(n % 2 == 0))
        add("c%dx%d_grey" % (w, h), buf.getvalue())
    # restart intervals (PIL: restart_marker_blocks / restart_marker_rows), 4:1:1-like via cv2 sampling factors when present
    img = Image.fromarray(test_image(120, 88, 99))
    for prog in (False, True):
        for sub in (0, 2):
            buf = io.BytesIO()
            img.save(buf, "JPEG", quality=80, subsampling=sub, progressive=prog, restart_marker_blocks=3)
            add("rst3_%d_%s" % (sub, "prog" if prog else "base"), buf.getvalue())
            buf = io.BytesIO()
            img.save(buf, "JPEG", quality=80, subsampling=sub, progressive=prog, restart_marker_rows=1)
            add("rstrow_%d_%s" % (sub, "prog" if prog else "base"), buf.getvalue())
    try:
        import cv2
        bgr = np.ascontiguousarray(test_image(97, 61, 7)[:, :, ::-1])
        for name, sf in (("411", getattr(cv2, "IMWRITE_JPEG_SAMPLING_FACTOR_411", None)), ("440", getattr(cv2, "IMWRITE_JPEG_SAMPLING_FACTOR_440", None))):
            if sf is None:
                continue
            ok, enc = cv2.imencode(".jpg", bgr, [cv2.IMWRITE_JPEG_QUALITY, 88, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sf])
            if ok:
                add("cv_" + name, enc.tobytes())
    except Exception as e:  # optional
        print("cv2 cases skipped:", e)
    # CMYK / YCCK (Adobe marker) from PIL
    cmyk = Image.fromarray(test_image(40, 24, 5)).convert("CMYK")
    buf = io.BytesIO(); cmyk.save(buf, "JPEG", quality=90); add("cmyk", buf.getvalue())
    # the reference's own models: CRC of every JPEG image as the reference decodes it
    models = {}
    for name in ("cube_plane", "sceneHierTest", "test/EmissiveTest", "test/sunshadowTest"):
        d = open("/root/reference/models/%s.glb" % name, "rb").read()
        jl = struct.unpack("<I", d[12:16])[0]
        j = json.loads(d[20:20 + jl])
        bin0 = d[20 + jl + 8:]
        for i, im in enumerate(j.get("images", [])):
            bv = j["bufferViews"][im["bufferView"]]
            b = bin0[bv.get("byteOffset", 0):bv.get("byteOffset", 0) + bv["byteLength"]]
            if b[:2] != b"\xff\xd8":
                continue
            ref = stb_decode(b)
            models["%s#%d" % (name, i)] = {"shape": list(ref.shape), "crc": zlib.crc32(ref.tobytes()) & 0xFFFFFFFF, "bytes": len(b), "file_crc": zlib.crc32(b) & 0xFFFFFFFF}
    json.dump({"cases": cases, "reference_models": models, "decoder": "stb_image 2.29 (reference thirdparty/stb_image), g++ x86-64 build in oracle/_ref/libdrt_ref.so"},
              open(os.path.join(OUT, "golden.json"), "w"), indent=1, sort_keys=True)
    print(len(cases), "cases,", len(models), "model images; bytes:", sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT)))


if __name__ == "__main__":
    main()
