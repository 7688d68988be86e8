"""The CPU oracle (oracle/drt_oracle.c) against the golden vectors generated from the reference's own
sources, and -- when the host-compiled reference (oracle/_ref) is present -- against the reference live.
No GPU, no product code."""
import ctypes
import os

import numpy as np
import pytest

import oracles as O

GOLDEN = os.path.join(O.ROOT, "tests", "golden")
SCENES = ["cornell_box", "room", "dense_monkey", "valorant_ghost", "animemaid", "mc_transparency_test", "bvhsplit_test", "cube", "sunshadow_test"]


@pytest.fixture(scope="module")
def port():
    return O.PortOracle()


@pytest.fixture(scope="module")
def kat():
    return np.load(os.path.join(GOLDEN, "kat.npz"))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_pcg_and_random_float(port, kat):
    L = port.L
    for s, want, rf, rs in zip(kat["seeds"], kat["pcg"], kat["random_float"], kat["random_float_seed"]):
        assert L.orc_pcg_hash(int(s)) == int(want)
        c = ctypes.c_uint(int(s))
        got = np.float32(L.orc_random_float(ctypes.byref(c)))
        assert bits(got) == bits(np.float32(rf)) and c.value == int(rs)


def test_unit_vectors_follow_the_host_argument_order(port, kat):
    L = port.L
    for s, v, vs, sp, sps in zip(kat["seeds"], kat["unit_vec3"], kat["unit_vec3_seed"], kat["unit_sphere_vec3"], kat["unit_sphere_vec3_seed"]):
        o = (ctypes.c_float * 3)()
        c = ctypes.c_uint(int(s))
        L.orc_random_unit_vec3(ctypes.byref(c), 1, o)
        assert np.array_equal(bits(np.array(o, np.float32)), bits(v)) and c.value == int(vs)
        c = ctypes.c_uint(int(s))
        L.orc_random_unit_sphere_vec3(ctypes.byref(c), 1, o)
        assert np.array_equal(bits(np.array(o, np.float32)), bits(sp)) and c.value == int(sps)
        # left-to-right order (nvcc) = same draws with x and z exchanged before normalisation
        c = ctypes.c_uint(int(s))
        L.orc_random_unit_vec3(ctypes.byref(c), 0, o)
        assert c.value == int(vs) and np.allclose(np.array(o)[::-1], v, atol=1e-6)


def test_slab_known_answers(port, kat):
    for b, o, d, want in zip(kat["slab_box"], kat["slab_o"], kat["slab_d"], kat["slab_t"]):
        got = np.float32(port.L.orc_bounds_intersect(O.P(b.copy()), O.P(o.copy()), O.P(d.copy())))
        assert bits(got) == bits(np.float32(want)) or (np.isnan(got) and np.isnan(want))


def test_moller_trumbore_known_answers(port, kat):
    for tri, o, d, h, t, uvw in zip(kat["mt_tri"], kat["mt_o"], kat["mt_d"], kat["mt_hit"], kat["mt_t"], kat["mt_uvw"]):
        tt = ctypes.c_float()
        out = (ctypes.c_float * 3)()
        got = port.L.orc_intersection(O.P(tri.copy()), O.P(o.copy()), O.P(d.copy()), ctypes.byref(tt), out)
        assert got == int(h)
        if h:
            assert bits(np.float32(tt.value)) == bits(np.float32(t))
            assert np.array_equal(bits(np.array(out, np.float32)), bits(uvw))


def _port_scene(port, name):
    """Builds the port's scene from the product's HOST loader output (no GPU involved)."""
    import dustraytracer_b200 as drt
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    tris = s.triangles()
    mats = np.frombuffer(b"".join(bytes(m) for m in s.materials()), np.uint32).reshape(-1, 11).copy() if s.stats()["materials"] else np.zeros((0, 11), np.uint32)
    mats[:, 8] &= 0xFF
    mats[:, 10] &= 0xFF
    tex = [s.texture(i) for i in range(s.stats()["textures"])]
    return port.create(tris, mats, tex), tris, mats, tex


@pytest.mark.parametrize("name", SCENES)
def test_scene_against_golden(port, name):
    g = np.load(os.path.join(GOLDEN, "scene_%s.npz" % name))
    h, tris, mats, tex = _port_scene(port, name)
    # the loader reproduces the reference's triangle soup, materials and texels bit for bit
    assert O.np.uint32(__import__("zlib").crc32(tris.tobytes())) == g["tri_crc"]
    assert np.uint32(__import__("zlib").crc32(mats.tobytes())) == g["mat_crc"]
    assert [np.uint32(__import__("zlib").crc32(t.tobytes())) for t in tex] == list(g["tex_crc"])
    # the restated builder reproduces the reference's node table and leaf order
    assert np.array_equal(port.nodes(h), g["ref_nodes"])
    assert np.array_equal(port.canonical_ids(h, tris.shape[0]), g["ref_leaf_order"])
    pw, ph = g["primary_size"]
    pid, t, u, v = port.trace_primary(h, g["camera"], int(pw), int(ph), 1)
    assert np.array_equal(pid, g["prim_id"])
    import zlib
    assert zlib.crc32(t.tobytes()) == int(g["crc_t"]) and zlib.crc32(u.tobytes()) == int(g["crc_u"]) and zlib.crc32(v.tobytes()) == int(g["crc_v"])
    rw, rh, frames, sun, bounces = [int(x) for x in g["render_size"]]
    acc, rays, _ = port.render(h, g["camera"], g["settings"], rw, rh, 1, frames)
    assert np.array_equal(bits(acc), bits(g["acc"]))
    assert list(rays) == [int(x) for x in g["rays"]]
    port.free(h)


@pytest.mark.skipif(not O.RefOracle.available(), reason="host-compiled reference (oracle/_ref) not built on this machine")
@pytest.mark.parametrize("name,sun,bounces", [("room", True, 8), ("animemaid", True, 2), ("cornell_box", False, 4)])
def test_port_equals_reference_live(port, name, sun, bounces):
    R = O.RefOracle()
    rh = R.load(O.scene_path(name))
    tris, mats, tex = R.triangles_load_order(rh), R.materials(rh), R.textures(rh)
    ph = port.create(tris, mats, tex)
    assert np.array_equal(R.nodes(rh), port.nodes(ph))
    cam = O.default_camera_bytes()
    cam[8:11] = (0.5, 1.5, 4.0)     # a camera and frame range the golden files do not hold
    cam[11:14] = (-0.2, -0.1, -1.0)
    a = R.trace_primary(rh, cam, 80, 60, 3)
    b = port.trace_primary(ph, cam, 80, 60, 3)
    for x, y in zip(a, b):
        assert np.array_equal(bits(x), bits(y))
    st = O.settings_bytes(sun=sun, bounces=bounces)
    ra, rr, _ = R.render(rh, cam, st, 64, 48, 5, 3)
    pa, pr, _ = port.render(ph, cam, st, 64, 48, 5, 3)
    assert np.array_equal(bits(ra), bits(pa)) and rr == pr
    R.free(rh)
    port.free(ph)


def test_port_camera_rays_match_the_reference_kats(port):
    """orc_camera_rays against the camera vectors tools/make_golden.py took from the reference's Camera::GetRay
    (pin-hole and thin-lens cases, Camera.cu:82-123)."""
    k = np.load(os.path.join(GOLDEN, "kat.npz"))
    for (x, y, W, H, f, defocus), ray in zip(k["cam_cases"], k["cam_rays"]):
        cam = k["cam_defocus"] if defocus else k["cam_default"]
        r = port.camera_rays(cam, int(W), int(H), int(f), rtl=True)
        assert np.array_equal(bits(r[x + y * W]), bits(ray))


@pytest.mark.parametrize("name", ["cube", "room", "cornell_box"])
def test_brute_force_checker_agrees_with_the_tree(port, name):
    """The tree-free checker used for the soups (orc_brute_force_closest) against the port's own BVH traversal (itself pinned to
    the reference's golden hits): same t, u, v bits; the id may differ only where two triangles are hit at exactly the same t
    (the tree keeps the first it meets, the checker the lowest index)."""
    h, tris, mats, tex = _port_scene(port, name)
    cam = O.default_camera_bytes()
    W, H = 64, 36
    pid, t, u, v = port.trace_primary(h, cam, W, H, 2)
    rays = port.camera_rays(cam, W, H, 2, rtl=True)
    tri9 = np.ascontiguousarray(tris.reshape(-1, 32)[:, [4, 5, 6, 12, 13, 14, 20, 21, 22]])
    bid, bt, bu, bv = port.brute_force_closest(tri9, rays)
    assert np.array_equal(bits(bt), bits(t))
    same = bid == pid
    assert same.mean() > 0.99
    assert np.array_equal(bits(bu[same]), bits(u[same])) and np.array_equal(bits(bv[same]), bits(v[same]))
    port.free(h)
