"""The JPEG decoder of the scene loader (csrc/drt_jpeg.cpp, drt_decode_image) against the REFERENCE's decoder: stb_image 2.29 as
vendored by the reference and compiled into oracle/_ref (fixtures by tools/make_jpeg_golden.py). The bar is bit equality of every
decoded byte (stb's IDCT, upsampling and colour conversion are deterministic integer code). No GPU needed."""
import ctypes
import glob
import json
import os
import struct
import zlib

import numpy as np
import pytest

import oracles as O
import dustraytracer_b200 as drt
from dustraytracer_b200 import api

JDIR = os.path.join(O.ROOT, "tests", "golden", "jpeg")
GOLD = json.load(open(os.path.join(JDIR, "golden.json")))


@pytest.mark.parametrize("name", sorted(GOLD["cases"]))
def test_decoded_bytes_equal_the_reference_decoder(name):
    g = GOLD["cases"][name]
    data = open(os.path.join(JDIR, name + ".jpg"), "rb").read()
    img = api.decode_image(data)
    assert list(img.shape) == g["shape"]
    assert [int(v) for v in img.reshape(-1)[:: max(1, img.size // 16)][:16]] == g["probe"]
    assert zlib.crc32(img.tobytes()) & 0xFFFFFFFF == g["crc"]


def test_matrix_covers_the_modes():
    names = set(GOLD["cases"])
    for frag in ("_444_", "_422_", "_420_", "_prog_", "_base_", "_grey", "rst3_", "rstrow_", "cmyk", "c1x1_", "c33x17_"):
        assert any(frag in n for n in names), frag
    assert len(GOLD["reference_models"]) == 4


@pytest.mark.skipif(not O.RefOracle.available(), reason="host-compiled reference (oracle/_ref) not built on this machine")
def test_live_against_the_reference_decoder():
    """Every fixture again, decoded NOW by the reference's stb build (not only the stored CRCs), byte for byte."""
    L = ctypes.CDLL(O.REF_LIB)
    L.stbi_load_from_memory.restype = ctypes.POINTER(ctypes.c_ubyte)
    L.stbi_load_from_memory.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    L.stbi_image_free.argtypes = [ctypes.c_void_p]
    for path in sorted(glob.glob(os.path.join(JDIR, "*.jpg"))):
        data = open(path, "rb").read()
        w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        p = L.stbi_load_from_memory(data, len(data), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), 0)
        assert p, path
        ref = np.ctypeslib.as_array(p, shape=(h.value, w.value, c.value)).copy()
        L.stbi_image_free(p)
        got = api.decode_image(data)
        assert got.shape == ref.shape and np.array_equal(got, ref), (os.path.basename(path), int((got != ref).sum()))


REF_MODELS = "/root/reference/models"


@pytest.mark.skipif(not os.path.isdir(REF_MODELS), reason="the reference's model directory exists in the build container only")
def test_every_bundled_reference_model_loads_and_jpeg_texels_match():
    """All GLBs the reference ships load through the reference-compatible loader (four of them embed a JPEG, which round 1
    rejected), and the JPEG texels equal what the reference's decoder returns for the same bytes."""
    paths = sorted(glob.glob(os.path.join(REF_MODELS, "**", "*.glb"), recursive=True))
    assert len(paths) >= 30
    loaded = 0
    for p in paths:
        s = drt.Scene()
        s.loadGLTFmodel(p)
        loaded += 1
        rel = os.path.relpath(p, REF_MODELS)[:-4]
        for key, g in GOLD["reference_models"].items():
            name, idx = key.split("#")
            if name == rel:
                tex = s.texture(int(idx))
                assert list(tex.shape) == g["shape"] and zlib.crc32(tex.tobytes()) & 0xFFFFFFFF == g["crc"], key
    assert loaded == len(paths)


def test_corrupt_and_truncated_streams_fail_cleanly():
    data = open(os.path.join(JDIR, "c64x48_420_base_q75.jpg"), "rb").read()
    with pytest.raises(drt.DrtError):
        api.decode_image(b"\xff\xd8\xff\xe0" + b"\x00" * 20)
    with pytest.raises(drt.DrtError):
        api.decode_image(b"not an image at all")
    # a truncated entropy-coded segment still decodes (zeros are fed, like stb), to the right size
    img = api.decode_image(data[: len(data) * 2 // 3])
    assert img.shape == (48, 64, 3)
    for cut in (2, 3, 10, 100, 200):
        try:
            api.decode_image(data[:cut])
        except drt.DrtError:
            pass
