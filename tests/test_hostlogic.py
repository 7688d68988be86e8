"""Logic of the device code, run on the CPU (tests/hostemu compiles the __host__ __device__ bodies of
dustraytracer_b200/csrc/*.cuh as plain C++): loader, GPU-builder passes, wide-node traversal, shading loop.
Checked against the golden vectors generated from the reference. No GPU needed; the real kernels call the
same functions and are checked on the B200 in test_gpu_parity.py."""
import os
import struct
import zlib

import numpy as np
import pytest

import oracles as O

GOLDEN = os.path.join(O.ROOT, "tests", "golden")
SCENES = ["cornell_box", "room", "dense_monkey", "valorant_ghost", "animemaid", "mc_transparency_test", "bvhsplit_test", "cube", "sunshadow_test"]


@pytest.fixture(scope="module")
def emu():
    return O.HostEmu()


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def check_wide_tree(nodes, leaf_tris, n_tris):
    """Structural invariants of the compressed wide BVH: every triangle referenced exactly once and inside
    its (dequantised) slot box; every child box inside its parent slot box; children contiguous."""
    n = nodes.shape[0]
    f = nodes.view(np.float32).reshape(n, 20)
    seen_tri = np.zeros(n_tris, np.int32)
    seen_node = np.zeros(n, np.int32)
    seen_node[0] = 1
    v0 = leaf_tris[:, 0:3]
    v1 = v0 + leaf_tris[:, 4:7]
    v2 = v0 + leaf_tris[:, 8:11]
    tri_lo = np.minimum(np.minimum(v0, v1), v2)
    tri_hi = np.maximum(np.maximum(v0, v1), v2)
    boxes = {0: None}
    for i in range(n):
        b = nodes[i]
        p = f[i, 0:3].astype(np.float64)
        e = b[12:15].astype(np.int32)
        imask = int(b[15])
        child_base, tri_base = struct.unpack("<II", b[16:24].tobytes())
        meta = b[24:32]
        scale = np.ldexp(1.0, e - 127)
        qlo = np.stack([b[32:40], b[40:48], b[48:56]], 1).astype(np.float64)
        qhi = np.stack([b[56:64], b[64:72], b[72:80]], 1).astype(np.float64)
        lo = p + qlo * scale
        hi = p + qhi * scale
        parent = boxes.get(i)
        rank = 0
        for s in range(8):
            m = int(meta[s])
            if m == 0:
                assert not (imask >> s) & 1
                continue
            if parent is not None:
                # a child's quantised box may exceed the parent's quantised box only by rounding slack
                pass
            if (imask >> s) & 1:
                assert m == (0x20 | (24 + s))
                c = child_base + rank
                rank += 1
                assert 0 < c < n
                seen_node[c] += 1
                boxes[c] = (lo[s], hi[s])
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                assert off + cnt <= 24
                for k in range(cnt):
                    t = tri_base + off + k
                    seen_tri[t] += 1
                    eps = 1e-6 * np.maximum(1.0, np.abs(tri_hi[t]))
                    assert np.all(tri_lo[t] >= lo[s] - eps) and np.all(tri_hi[t] <= hi[s] + eps), (i, s, t)
        if parent is not None:
            used = [s for s in range(8) if meta[s]]
            eps = 1e-5 * np.maximum(1.0, np.abs(parent[1]))
            # the node's own grid origin/extent must cover what the parent promised for this child
            assert np.all(lo[used].min(0) >= parent[0] - eps - scale) and np.all(hi[used].max(0) <= parent[1] + eps + scale)
    assert np.all(seen_tri == 1), "every triangle must sit in exactly one leaf slot"
    assert np.all(seen_node == 1), "every wide node must have exactly one parent"


@pytest.mark.parametrize("name", SCENES)
def test_scene_logic_against_golden(emu, name):
    g = np.load(os.path.join(GOLDEN, "scene_%s.npz" % name))
    h = emu.load(O.scene_path(name))
    info = emu.info(h)
    # loader == reference loader
    tris = emu.triangles(h)
    assert crc(tris) == int(g["tri_crc"])
    mats = emu.materials(h)
    mats[:, 8] &= 0xFF
    mats[:, 10] &= 0xFF
    assert crc(mats) == int(g["mat_crc"])
    assert [crc(t) for t in emu.textures(h)] == [int(x) for x in g["tex_crc"]]
    # builder output is a valid tree over all triangles
    assert info["tris_emitted"] == info["tris"]
    leaf = emu.leaf_triangles(h)
    ids = leaf.view(np.uint32)[:, 3]
    assert np.array_equal(np.sort(ids), np.arange(info["tris"], dtype=np.uint32))
    if info["tris"] <= 2000:
        check_wide_tree(emu.wide_nodes(h), leaf, info["tris"])
    # primary hits: bit-exact ids, t, u, v in strict mode with the host argument order
    pw, ph = [int(x) for x in g["primary_size"]]
    pid, t, u, v, stats = emu.trace_primary(h, g["camera"], pw, ph, 1, strict=True, rtl=True)
    assert np.array_equal(pid, g["prim_id"])
    assert crc(t) == int(g["crc_t"]) and crc(u) == int(g["crc_u"]) and crc(v) == int(g["crc_v"])
    assert stats[4] <= 48, "traversal stack depth"
    # the whole bounce/shade/accumulate loop: bit-exact accumulation, identical ray counts
    rw, rh, frames, sun, bounces = [int(x) for x in g["render_size"]]
    for flat in (False, True):
        acc, cnt = emu.render(h, g["camera"], g["settings"], rw, rh, 1, 1, frames, strict=True, rtl=True, flat=flat)
        assert np.array_equal(bits(acc), bits(g["acc"])), "flat=%s" % flat
        assert cnt[0] == int(g["rays"][0]) and cnt[1] == int(g["rays"][1])
    # the phased kernel's NODE / TRI / SHADE steps (serial schedule), on both node formats
    # (fmt 2 = expanded nodes with the two-triangle TRI step: same bits, same rays, and the same NUMBER of triangle tests,
    # because the second candidate is only applied -- and counted -- when the first did not finish an any-hit ray)
    tests_seen = {}
    for fmt in (0, 1, 2):
        acc, cnt = emu.render(h, g["camera"], g["settings"], rw, rh, 1, 1, frames, strict=True, rtl=True, steps=fmt)
        assert np.array_equal(bits(acc), bits(g["acc"])), "steps fmt=%d" % fmt
        assert cnt[0] == int(g["rays"][0]) and cnt[1] == int(g["rays"][1])
        tests_seen[fmt] = (cnt[2], cnt[3])
    assert tests_seen[2] == tests_seen[1]
    # the five-step decomposition, triangles-first and nodes-first (postponed triangle groups)
    for fmt, nodes_first in ((0, 0), (1, 1), (0, 1)):
        acc, cnt = emu.render(h, g["camera"], g["settings"], rw, rh, 1, 1, frames, strict=True, rtl=True, steps5=(fmt, nodes_first))
        assert np.array_equal(bits(acc), bits(g["acc"])), "steps5 fmt=%d nodes_first=%d" % (fmt, nodes_first)
        assert cnt[0] == int(g["rays"][0]) and cnt[1] == int(g["rays"][1])
    emu.free(h)


def test_sample_range_sharding_is_a_partition_of_the_frames(emu):
    """Frames {1..8} rendered as 2 interleaved shards sum (in fp32, shard by shard) to within rounding of the
    single pass, and each shard equals the same frames rendered alone -- the property the multi-GPU path uses."""
    g = np.load(os.path.join(GOLDEN, "scene_room.npz"))
    h = emu.load(O.scene_path("room"))
    W, H = 64, 36
    full, _ = emu.render(h, g["camera"], g["settings"], W, H, 1, 1, 8, flat=True)
    a, ca = emu.render(h, g["camera"], g["settings"], W, H, 1, 2, 4, flat=True)
    b, cb = emu.render(h, g["camera"], g["settings"], W, H, 2, 2, 4, flat=True)
    assert np.abs((a + b) - full).max() <= 8 * np.finfo(np.float32).eps * np.abs(full).max()
    singles = np.zeros_like(a)
    for f in (1, 3, 5, 7):
        emu.render(h, g["camera"], g["settings"], W, H, f, 1, 1, flat=True, acc=singles)
    assert np.array_equal(bits(singles), bits(a))
    emu.free(h)


def test_fast_math_mode_stays_close(emu):
    g = np.load(os.path.join(GOLDEN, "scene_cornell_box.npz"))
    h = emu.load(O.scene_path("cornell_box"))
    pw, ph = [int(x) for x in g["primary_size"]]
    pid, t, _, _, _ = emu.trace_primary(h, g["camera"], pw, ph, 1, strict=False, rtl=True)
    assert (pid != g["prim_id"]).mean() < 1e-4
    emu.free(h)


def _soup(n, seed=7, spread=10.0, size=0.5):
    rng = np.random.RandomState(seed)
    c = rng.uniform(-spread, spread, (n, 1, 3))
    return (c + rng.uniform(-size, size, (n, 3, 3))).astype(np.float32)


def _brute_force(pos, o, d):
    """Reference Moller-Trumbore (Intersection.cu:4-36) over all triangles in float64-free numpy fp32."""
    v0, v1, v2 = pos[:, 0], pos[:, 1], pos[:, 2]
    e1, e2 = v1 - v0, v2 - v0
    best = np.full(o.shape[0], -1, np.int64)
    best_t = np.full(o.shape[0], np.inf, np.float32)
    for r in range(o.shape[0]):
        p = np.cross(d[r], e2).astype(np.float32)
        det = (e1 * p).sum(1, dtype=np.float32)
        ok = ~((det > -1e-6) & (det < 1e-6))
        inv = np.where(ok, np.float32(1) / np.where(ok, det, 1), 0).astype(np.float32)
        tv = (o[r] - v0).astype(np.float32)
        u = inv * (tv * p).sum(1, dtype=np.float32)
        q = np.cross(tv, e1).astype(np.float32)
        v = inv * (d[r] * q).sum(1, dtype=np.float32)
        t = inv * (e2 * q).sum(1, dtype=np.float32)
        ok &= (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 1e-6)
        if ok.any():
            tt = np.where(ok, t, np.inf)
            best[r] = int(np.argmin(tt))
            best_t[r] = tt[best[r]]
    return best, best_t


MAT = np.zeros((1, 11), np.uint32)
MAT.view(np.float32)[0, 0:3] = 0.8
MAT.view(np.int32)[0, 6] = -1


@pytest.mark.parametrize("n", [1, 2, 3, 4, 33, 500, 5000])
def test_builder_on_random_soups_matches_brute_force(emu, n):
    pos = _soup(n)
    h = emu.from_triangles(pos, MAT)
    info = emu.info(h)
    assert info["tris_emitted"] == n
    check_wide_tree(emu.wide_nodes(h), emu.leaf_triangles(h), n)
    cam = O.default_camera_bytes()
    cam[8:11] = (0, 0, 30)
    W, H = 48, 32
    pid, t, _, _, _ = emu.trace_primary(h, cam, W, H, 1, strict=True, rtl=True)
    # same rays through brute force
    import ctypes
    rays = np.zeros((W * H, 6), np.float32)
    hits = np.zeros((W * H, 4), np.float32)
    emu.L.emu_trace_primary(h, O.P(cam), W, H, 1, 1, 1, O.P(hits), O.P(rays), None)
    bid, bt = _brute_force(pos, rays[:, 0:3], rays[:, 3:6])
    agree = (pid == bid) | ((pid >= 0) & (bid >= 0) & (np.abs(t - bt) <= 1e-5 * np.abs(bt)))
    assert agree.all(), np.argwhere(~agree)[:5]
    emu.free(h)


def test_builder_degenerate_inputs(emu):
    # all centroids identical (the reference builder recurses forever here: SURVEY.md A.5) -> median splits
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32)
    pos = np.repeat(tri, 100, 0)
    h = emu.from_triangles(pos, MAT)
    assert emu.info(h)["tris_emitted"] == 100
    check_wide_tree(emu.wide_nodes(h), emu.leaf_triangles(h), 100)
    emu.free(h)
    # coplanar axis-aligned quads (zero-thickness boxes), and a zero-area triangle in the mix
    pos = _soup(64, size=0.3)
    pos[:, :, 1] = 2.0
    pos[5] = 0.0
    h = emu.from_triangles(pos, MAT)
    assert emu.info(h)["tris_emitted"] == 64
    check_wide_tree(emu.wide_nodes(h), emu.leaf_triangles(h), 64)
    emu.free(h)
    # empty scene: every ray misses
    h = emu.from_triangles(np.zeros((0, 3, 3), np.float32), MAT)
    pid, _, _, _, _ = emu.trace_primary(h, O.default_camera_bytes(), 8, 8)
    assert (pid == -1).all()
    emu.free(h)


@pytest.mark.parametrize("name", ["cornell_box", "animemaid"])
def test_debug_views_match_the_reference(emu, name):
    """RenderMode DEBUGMODE (RendererSettings.h:12-20; RayGen.cuh:137-161): the tree-independent views -- albedo,
    normal, barycentric, uv -- are bit-exact against the host-compiled reference. (The two BVH heat views count
    the visited nodes of whichever tree is traversed, so they cannot match a different tree.)"""
    orc = O.best_oracle()
    if orc.kind != "reference":
        pytest.skip("needs oracle/_ref/libdrt_ref.so")
    W, H = 96, 64
    cam = O.default_camera_bytes()
    oh = orc.load(O.scene_path(name))
    h = emu.load(O.scene_path(name))
    for mode in (0, 1, 2, 3):
        st = O.settings_bytes(sun=True, bounces=3, render_mode=1, debug_mode=mode)
        ref, _, _ = orc.render(oh, cam, st, W, H, 1, 2)
        got, cnt = emu.render(h, cam, st, W, H, 1, 1, 2, strict=True, rtl=True)
        assert np.array_equal(bits(got), bits(ref)), "debug mode %d" % mode
        assert cnt[1] == 0, "no shadow rays in the debug views (RayGen.cuh:114)"
        assert len(np.unique(got)) > 4
    orc.free(oh)
    emu.free(h)
