"""Parity of the CUDA path (through the C ABI, libdrt.so) against the oracle. Needs a B200: run with -m gpu.

Bars (BASELINE.json north_star):
  * primary-ray hit triangle ids: bit-exact against the reference (strict mode also reproduces t, u, v and
    whole accumulated images bit for bit; exact-t ties are counted and logged);
  * production (fused-multiply-add, hardware rsqrt) mode: ids equal except logged near-tie pixels
    (|dt| <= 1e-4 * t), converged radiance within RMSE_TOL per pixel of the oracle at equal sample count.
"""
import os
import zlib

import numpy as np
import pytest

import oracles as O

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(O.ROOT, "tests", "golden")
SCENES = ["cornell_box", "room", "dense_monkey", "valorant_ghost", "animemaid", "mc_transparency_test", "bvhsplit_test", "cube", "sunshadow_test"]
RMSE_TOL = 0.05      # LDR units (display values in [0,1]) at 64 spp; two independent 64-spp estimates differ by ~0.03
MEAN_TOL = 0.004     # bias check: mean over the image of the signed difference
STRICT = 1 | 2       # DRT_FLAG_STRICT | DRT_FLAG_HOST_ARG_ORDER (the oracle is g++-compiled)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


@pytest.fixture(scope="module")
def drt():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import dustraytracer_b200 as d
    return d


_scene_cache = {}


def gpu_scene(drt, name):
    if name not in _scene_cache:
        s = drt.Scene()
        s.loadGLTFmodel(O.scene_path(name))
        drt.BVHBuilder().buildIterative(s)
        _scene_cache[name] = s
    return _scene_cache[name]


def camera_from(drt, cam_f32):
    import ctypes
    cam = drt.Camera()
    ctypes.memmove(ctypes.addressof(cam), cam_f32.tobytes(), 88)
    return cam


def settings_from(drt, st_bytes):
    import ctypes
    st = drt.RendererSettings()
    ctypes.memmove(ctypes.addressof(st), st_bytes.tobytes(), 64)
    return st


@pytest.mark.parametrize("name", SCENES)
def test_primary_hits_bit_exact(drt, name):
    g = np.load(os.path.join(GOLDEN, "scene_%s.npz" % name))
    scene = gpu_scene(drt, name)
    st = scene.stats()
    assert st["wide_nodes"] >= 1 and st["build_ms"] > 0
    pw, ph = [int(x) for x in g["primary_size"]]
    r = drt.Renderer(flags=STRICT)
    r.ResizeBuffer(pw, ph)
    hits = r.trace_primary(camera_from(drt, g["camera"]), scene, 1)
    assert np.array_equal(hits["prim_id"], g["prim_id"]), "%d id mismatches" % (hits["prim_id"] != g["prim_id"]).sum()
    assert crc(hits["t"]) == int(g["crc_t"]) and crc(hits["u"]) == int(g["crc_u"]) and crc(hits["v"]) == int(g["crc_v"])
    c = r.counters()
    print("%s: %d exact-t tie events over %d rays, max stack %d" % (name, c["tie_events"], pw * ph, c["max_stack"]))
    # production arithmetic: same ids except logged near-ties
    r.set_flags(2)
    fast = r.trace_primary(camera_from(drt, g["camera"]), scene, 1)
    diff = np.flatnonzero(fast["prim_id"] != g["prim_id"])
    tg = np.full(pw * ph, -1, np.float32)
    tg.reshape(ph, pw)[::4, ::4] = g["tuv_stride4"][0]
    print("%s: production-mode id differences: %d of %d pixels" % (name, diff.size, pw * ph))
    assert diff.size <= max(2, int(2e-5 * pw * ph))
    sub = fast["t"].reshape(ph, pw)[::4, ::4]
    same = (fast["prim_id"].reshape(ph, pw)[::4, ::4] == g["prim_id"].reshape(ph, pw)[::4, ::4]) & (g["tuv_stride4"][0] > 0)
    assert np.all(np.abs(sub[same] - g["tuv_stride4"][0][same]) <= 1e-4 * g["tuv_stride4"][0][same])


@pytest.mark.parametrize("variant", [4, 2, 3, 0, 1, 9, 10, 11, 14, 15, 16, 17, 22, 23])
@pytest.mark.parametrize("name", SCENES)
def test_accumulated_image_bit_exact(drt, name, variant):
    g = np.load(os.path.join(GOLDEN, "scene_%s.npz" % name))
    scene = gpu_scene(drt, name)
    rw, rh, frames, sun, bounces = [int(x) for x in g["render_size"]]
    r = drt.Renderer(flags=STRICT)
    r.set_tuning(variant, 0)
    r.m_RendererSettings = settings_from(drt, g["settings"])
    r.ResizeBuffer(rw, rh)
    cam = camera_from(drt, g["camera"])
    r.Render(cam, scene, frames)           # one batched launch
    acc = r.readback(drt.DRT_ACCUM_RGB32F).reshape(-1)
    assert np.array_equal(bits(acc), bits(g["acc"])), "max abs diff %g" % np.abs(acc - g["acc"]).max()
    c = r.counters()
    assert c["closest_rays"] == int(g["rays"][0]) and c["shadow_rays"] == int(g["rays"][1])
    assert r.getSampleCount() == frames + 1
    disp = r.readback(drt.DRT_RGBA32F)
    assert np.array_equal(bits(disp[:, :, :3].reshape(-1)), bits(g["acc"] / np.float32(frames))) and np.all(disp[:, :, 3] == 1)
    # frame by frame, like the reference's one launch per frame
    r.resetAccumulationBuffer()
    assert r.getSampleCount() == 1
    for _ in range(frames):
        r.Render(cam, scene)
    assert np.array_equal(bits(r.readback(drt.DRT_ACCUM_RGB32F).reshape(-1)), bits(g["acc"]))


@pytest.mark.parametrize("name,sun,bounces", [("room", True, 8), ("animemaid", True, 2), ("dense_monkey", False, 4)])
def test_production_mode_converged_radiance(drt, name, sun, bounces):
    """Production arithmetic + nvcc argument order against the oracle at equal sample count."""
    W, H, spp = 128, 72, 64
    orc = O.best_oracle()
    cam_f = O.default_camera_bytes()
    st_b = O.settings_bytes(sun=sun, bounces=bounces)
    if orc.kind == "reference":
        oh = orc.load(O.scene_path(name))
    else:
        import dustraytracer_b200 as d
        s = d.Scene()
        s.loadGLTFmodel(O.scene_path(name))
        mats = np.frombuffer(b"".join(bytes(m) for m in s.materials()), np.uint32).reshape(-1, 11).copy()
        oh = orc.create(s.triangles(), mats, [s.texture(i) for i in range(s.stats()["textures"])])
    ref, _, _ = orc.render(oh, cam_f, st_b, W, H, 1, spp)
    ref = ref.reshape(H, W, 3) / spp
    scene = gpu_scene(drt, name)
    r = drt.Renderer()
    r.m_RendererSettings = settings_from(drt, st_b)
    r.ResizeBuffer(W, H)
    r.Render(camera_from(drt, cam_f), scene, spp)
    img = r.readback(drt.DRT_RGBA32F)[:, :, :3]
    rmse = float(np.sqrt(np.mean((img - ref) ** 2)))
    bias = float(np.mean(img - ref))
    print("%s: production vs %s oracle at %d spp: RMSE %.4f, mean diff %+.5f" % (name, orc.kind, spp, rmse, bias))
    assert rmse <= RMSE_TOL and abs(bias) <= MEAN_TOL
    orc.free(oh)


@pytest.mark.parametrize("name,size", [("room", (101, 67)), ("animemaid", (37, 5)), ("cornell_box", (8, 4)), ("dense_monkey", (250, 131))])
def test_pixel_queue_kernel_matches_fixed_pixel_kernel(drt, name, size):
    """The kernels that keep the SHADE-only lane state in shared memory (variants 11-14) and their persistent pixel-queue
    form (variants 9 / 10, what one-sample launches run) against the all-register fixed-pixel kernel in production arithmetic: ragged image sizes, 1..3 samples per launch, several launches, same bits and ray counts."""
    scene = gpu_scene(drt, name)
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)

    def run(variant, batches):
        r = drt.Renderer()
        r.set_tuning(variant, 0)
        r.m_RendererSettings = st
        r.ResizeBuffer(*size)
        for b in batches:
            r.Render(cam, scene, b)
        return r.readback(drt.DRT_ACCUM_RGB32F), r.readback(drt.DRT_RGBA32F), r.counters()

    for batches in ([1, 1, 1, 1, 1], [3, 2], [5]):
        a4, d4, c4 = run(4, batches)
        a3, d3, c3 = run(3, batches)
        for variant in (9, 10, 11, 12, 13, 14, 15, 17, 18, 22, 23):
            a, d, c = run(variant, batches)
            ref_a, ref_d, ref_c = (a3, d3, c3) if variant in (10, 14) else (a4, d4, c4)
            assert np.array_equal(bits(a), bits(ref_a)) and np.array_equal(bits(d), bits(ref_d)), (variant, batches)
            assert c == ref_c, (variant, batches)


@pytest.mark.parametrize("name", ["room", "animemaid"])
def test_launch_autotuner_is_invisible_in_the_results(drt, name):
    """A renderer left on automatic kernel selection times the fixed-pixel and the pixel-queue kernels on its first
    one-sample launches and keeps the faster one: the accumulated image must not depend on what it picked. (animemaid has more
    than 1024 triangles: automatic selection then also runs the two-triangle TRI step, in both kernel families.)"""
    scene = gpu_scene(drt, name)
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8, max_samples=100000)

    def run(explicit, calls):
        r = drt.Renderer()
        if explicit is not None:
            r.set_tuning(explicit, 0)
        r.m_RendererSettings = st
        r.ResizeBuffer(480, 270)
        for n in calls:
            r.Render(cam, scene, n)
        return r.readback(drt.DRT_ACCUM_RGB32F), r.counters(), r.kernel_launches()

    calls = [1] * 20 + [3, 3, 3, 3, 3, 3, 3, 3] + [16]
    a, ca, la = run(None, calls)
    b, cb, lb = run(4, calls)
    assert np.array_equal(bits(a), bits(b)) and ca == cb and la == lb


@pytest.mark.parametrize("variant", [4, 3, 9, 10, 0, 16, 22])
def test_tile_sharding_is_a_partition_of_the_image(drt, variant):
    """drt_renderer_set_tile_schedule: three renderers own the 8x4 tiles 0,3,6.. / 1,4,7.. / 2,5,8.. of a ragged image;
    each touches exactly the pixels of tile_owner_mask, and the SUM of their accumulation buffers has the bits of the
    unsharded render (x + 0 + 0 = x), ray counts included."""
    scene = gpu_scene(drt, "room")
    W, H, world = 203, 117, 3
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)

    def run(first, stride, batches):
        r = drt.Renderer(flags=STRICT)
        r.set_tuning(variant, 0)
        r.m_RendererSettings = st
        r.ResizeBuffer(W, H)
        r.set_tile_schedule(first, stride)
        for b in batches:
            r.Render(cam, scene, b)
        return r.readback(drt.DRT_ACCUM_RGB32F), r.readback(drt.DRT_RGBA32F), r.counters()

    for batches in ([1, 1, 1], [3]):
        full, dfull, cfull = run(0, 1, batches)
        total = np.zeros_like(full)
        rays = 0
        for rank in range(world):
            a, d, c = run(rank, world, batches)
            own = drt.tile_owner_mask(W, H, rank, world)
            assert np.array_equal(bits(a[own]), bits(full[own])) and not a[~own].any() and not d[~own].any()
            assert np.array_equal(bits(d[own]), bits(dfull[own]))
            total += a
            rays += c["closest_rays"] + c["shadow_rays"]
        assert np.array_equal(bits(total), bits(full))
        assert rays == cfull["closest_rays"] + cfull["shadow_rays"]


def test_empty_scene_renders_the_sky_in_every_kernel(drt):
    """Zero triangles (the edge the reference never guards: its builder dereferences an empty vector): every ray misses, every
    sample is one closest-hit ray and the sky; all kernel families agree bit for bit, ragged size included."""
    scene = drt.Scene.from_triangles(np.zeros((0, 3, 3), np.float32))
    drt.BVHBuilder().buildIterative(scene)
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=3, max_samples=1000)
    W, H, spp = 61, 35, 3
    ref = None
    for variant in (4, 0, 1, 3, 9, 10, 11):
        r = drt.Renderer(flags=STRICT)
        r.set_tuning(variant, 0)
        r.m_RendererSettings = st
        r.ResizeBuffer(W, H)
        r.Render(cam, scene, spp)
        r.Render(cam, scene)
        a, c = r.readback(drt.DRT_ACCUM_RGB32F), r.counters()
        assert c["closest_rays"] == W * H * (spp + 1) and c["shadow_rays"] == 0
        assert np.isfinite(a).all() and a.min() > 0
        if ref is None:
            ref = a
        assert np.array_equal(bits(a), bits(ref)), variant


def _ipc_child(handle, W, H, first, stride, batches, q):
    """Another PROCESS (spawned): imports the parent's frame buffers and renders its tiles into them."""
    try:
        import dustraytracer_b200 as d
        scene = d.Scene()
        scene.loadGLTFmodel(O.scene_path("room"))
        d.BVHBuilder().buildIterative(scene)
        r = d.Renderer(flags=STRICT)
        r.m_RendererSettings = d.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)
        r.ResizeBuffer(W, H)
        r.set_tile_schedule(first, stride)
        r.import_frame_buffers(handle)
        for b in batches:
            r.Render(d.Camera(), scene, b)
        c = r.counters()
        q.put(("ok", c["closest_rays"] + c["shadow_rays"]))
    except Exception as e:  # noqa: BLE001
        q.put(("error", repr(e)))


def test_tiles_rendered_straight_into_a_shared_frame_buffer(drt):
    """Image-tile sharding without a collective: renderers that own the other tiles store them directly into the root
    renderer's buffers -- attached in the same process, and imported over CUDA IPC by another process. The root then
    holds the bits of the unsharded render, display included, with nothing to reduce or resolve."""
    import multiprocessing as mp
    scene = gpu_scene(drt, "room")
    W, H = 203, 117
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=100000)
    batches = [1, 2, 1]

    def make(first, stride):
        r = drt.Renderer(flags=STRICT)
        r.m_RendererSettings = st
        r.ResizeBuffer(W, H)
        r.set_tile_schedule(first, stride)
        return r

    full = make(0, 1)
    for b in batches:
        full.Render(cam, scene, b)
    ref_a, ref_d, ref_c = full.readback(drt.DRT_ACCUM_RGB32F), full.readback(drt.DRT_RGBA32F), full.counters()

    # same process: rank 0 owns the buffers, "ranks" 1 and 2 are attached to them
    root = make(0, 3)
    peers = [make(1, 3), make(2, 3)]
    for p in peers:
        p.attach_frame_buffers(root)
    rays = 0
    for b in batches:
        for r in [root] + peers:
            r.Render(cam, scene, b)
    for r in [root] + peers:
        c = r.counters()
        rays += c["closest_rays"] + c["shadow_rays"]
    assert np.array_equal(bits(root.readback(drt.DRT_ACCUM_RGB32F)), bits(ref_a))
    assert np.array_equal(bits(root.readback(drt.DRT_RGBA32F)), bits(ref_d))
    assert np.array_equal(bits(peers[0].readback(drt.DRT_ACCUM_RGB32F)), bits(ref_a)), "an attached renderer reads the shared buffer"
    assert rays == ref_c["closest_rays"] + ref_c["shadow_rays"]
    peers[0].resetAccumulationBuffer()
    assert np.array_equal(bits(root.readback(drt.DRT_ACCUM_RGB32F)), bits(ref_a)), "a peer's reset leaves the owner's buffers alone"
    peers[1].attach_frame_buffers(None)
    assert not peers[1].readback(drt.DRT_ACCUM_RGB32F).any(), "detached: back on its own (zeroed) buffers"

    # another process: CUDA IPC
    root2 = make(0, 2)
    handle = root2.export_frame_buffers()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    child = ctx.Process(target=_ipc_child, args=(handle, W, H, 1, 2, batches, q))
    child.start()
    for b in batches:
        root2.Render(cam, scene, b)
    status, val = q.get(timeout=300)
    child.join(60)
    assert status == "ok", val
    c = root2.counters()
    assert val + c["closest_rays"] + c["shadow_rays"] == ref_c["closest_rays"] + ref_c["shadow_rays"]
    assert np.array_equal(bits(root2.readback(drt.DRT_ACCUM_RGB32F)), bits(ref_a))
    assert np.array_equal(bits(root2.readback(drt.DRT_RGBA32F)), bits(ref_d))
    with pytest.raises(drt.DrtError):
        make(0, 1).import_frame_buffers(b"\0" * 160)


def test_full_size_properties(drt):
    """BASELINE config-3 size (room, 1920x1080, sun, 8 bounces): determinism, batching invariance, sample-range
    sharding = partition of the frames, ray accounting."""
    scene = gpu_scene(drt, "room")
    W, H = 1920, 1080
    cam = drt.Camera()
    st = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8, max_samples=100000)

    def run(first, stride, n, batch, flags=0, variant=4):
        r = drt.Renderer(flags=flags)
        r.set_tuning(variant, 0)
        r.m_RendererSettings = st
        r.ResizeBuffer(W, H)
        r.set_frame_schedule(first, stride)
        done = 0
        while done < n:
            r.Render(cam, scene, min(batch, n - done))
            done += min(batch, n - done)
        return r.readback(drt.DRT_ACCUM_RGB32F), r.counters()

    a, ca = run(1, 1, 4, 4)
    b, cb = run(1, 1, 4, 1)
    assert np.array_equal(bits(a), bits(b)) and ca == cb, "batched launch == one launch per frame"
    a2, _ = run(1, 1, 4, 4)
    assert np.array_equal(bits(a), bits(a2)), "deterministic"
    assert 4 * W * H <= ca["closest_rays"] <= 9 * 4 * W * H and ca["shadow_rays"] <= ca["closest_rays"]
    s0, c0 = run(1, 2, 2, 2)
    s1, c1 = run(2, 2, 2, 2)
    assert c0["closest_rays"] + c1["closest_rays"] == ca["closest_rays"]
    assert np.abs((s0 + s1) - a).max() <= 4 * np.finfo(np.float32).eps * np.abs(a).max()
    # strict arithmetic: every kernel variant agrees bit for bit at full size ...
    v4, c4 = run(1, 1, 2, 2, flags=STRICT, variant=4)
    for variant in (0, 1, 3, 9, 10, 11, 15):
        v, c = run(1, 1, 2, 1 if variant in (9, 10) else 2, flags=STRICT, variant=variant)
        assert np.array_equal(bits(v4), bits(v)) and c["closest_rays"] == c4["closest_rays"], variant
    # ... and with the host-compiled reference (2 of the 256 frames of the 1080p image, a few seconds of CPU)
    orc = O.best_oracle()
    cam_f = O.default_camera_bytes()
    st_b = O.settings_bytes(sun=True, bounces=8, max_samples=100000)
    if orc.kind == "reference":
        oh = orc.load(O.scene_path("room"))
    else:
        mats = np.frombuffer(b"".join(bytes(m) for m in scene.materials()), np.uint32).reshape(-1, 11).copy()
        oh = orc.create(scene.triangles(), mats, [])
    ref, rays, _ = orc.render(oh, cam_f, st_b, W, H, 1, 2)
    orc.free(oh)
    assert np.array_equal(bits(v4.reshape(-1)), bits(ref)), "full-size accumulation differs from the %s oracle" % orc.kind
    assert c4["closest_rays"] == rays[0] and c4["shadow_rays"] == rays[1]
    assert np.isfinite(a).all() and a.min() >= 0


def test_renderer_semantics(drt):
    scene = gpu_scene(drt, "cornell_box")
    r = drt.Renderer()
    r.ResizeBuffer(64, 48)
    cam = drt.Camera()
    r.m_RendererSettings.max_samples = 4          # Renderer.cu:82: stops when the frame index reaches max_samples
    for _ in range(10):
        r.Render(cam, scene)
    assert r.getSampleCount() == 4
    r.Render(cam, scene, 100)
    assert r.getSampleCount() == 4
    img3 = r.readback(drt.DRT_RGBA32F)
    r.ResizeBuffer(64, 48)                        # unchanged size: no reset (Renderer.cu:30)
    assert r.getSampleCount() == 4
    r.ResizeBuffer(32, 24)
    assert r.getSampleCount() == 1 and r.getBufferWidth() == 32
    r.ResizeBuffer(64, 48)
    r.m_RendererSettings.max_samples = 500
    r.Render(cam, scene, 3)
    assert np.array_equal(bits(r.readback(drt.DRT_RGBA32F)), bits(img3))
    rgba8 = r.readback(drt.DRT_RGBA8)
    assert rgba8.dtype == np.uint8 and np.all(rgba8[:, :, 3] == 255)
    assert np.abs(rgba8[:, :, :3].astype(np.float32) - np.clip(img3[:, :, :3], 0, 1) * 255).max() <= 0.5001
    # checkpoint / resume
    acc = r.readback(drt.DRT_ACCUM_RGB32F)
    r2 = drt.Renderer()
    r2.ResizeBuffer(64, 48)
    r2.upload_accum(acc, 3)
    r.Render(cam, scene, 2)
    r2.Render(cam, scene, 2)
    assert np.array_equal(bits(r.readback(drt.DRT_ACCUM_RGB32F)), bits(r2.readback(drt.DRT_ACCUM_RGB32F)))
    # debug views run through the per-sample loop
    r.m_RendererSettings.render_mode = 1
    r.m_RendererSettings.debug_mode = 1
    r.resetAccumulationBuffer()
    r.Render(cam, scene)
    n = r.readback(drt.DRT_RGBA32F)[:, :, :3]
    lens = np.linalg.norm(n, axis=2)
    assert np.all((np.abs(lens - 1) < 1e-3) | (lens > 1.0))   # unit normals on hits, sky elsewhere
    with pytest.raises(drt.DrtError):
        drt.Renderer().Render(cam, scene)     # no buffer yet


def test_gpu_builder_on_synthetic_soups(drt):
    emu = O.HostEmu()
    mat = np.zeros((1, 11), np.uint32)
    mat.view(np.float32)[0, 0:3] = 0.8
    mat.view(np.int32)[0, 6] = -1
    for n in (0, 1, 3, 4, 33, 1000, 200000):
        rng = np.random.RandomState(n + 1)
        c = rng.uniform(-10, 10, (n, 1, 3))
        pos = (c + rng.uniform(-0.4, 0.4, (n, 3, 3))).astype(np.float32)
        s = drt.Scene.from_triangles(pos)
        s.build_bvh()
        if n:
            ids = s.leaf_triangles().view(np.uint32)[:, 3]
            assert np.array_equal(np.sort(ids), np.arange(n, dtype=np.uint32)), "every triangle exactly once"
        cam = drt.Camera(pos=(0, 0, 30))
        r = drt.Renderer(flags=STRICT)
        r.ResizeBuffer(160, 90)
        hits = r.trace_primary(cam, s, 1)
        eh = emu.from_triangles(pos, mat)
        pid, t, u, v, _ = emu.trace_primary(eh, np.frombuffer(bytes(cam), np.float32).copy(), 160, 90, 1, strict=True, rtl=True)
        emu.free(eh)
        # different tree instances (atomics order), same winner: ties go to the lowest canonical id
        assert np.array_equal(hits["prim_id"], pid) and np.array_equal(bits(hits["t"]), bits(t)), n
        print("soup %d: build %.3f ms, %d wide nodes, hit fraction %.3f" % (n, s.stats()["build_ms"], s.stats()["wide_nodes"], (pid >= 0).mean()))
    # degenerate: identical centroids -> median splits, still a valid tree
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32)
    s = drt.Scene.from_triangles(np.repeat(tri, 1000, 0))
    s.build_bvh()
    assert np.array_equal(np.sort(s.leaf_triangles().view(np.uint32)[:, 3]), np.arange(1000, dtype=np.uint32))


def test_stats_counters(drt):
    scene = gpu_scene(drt, "room")
    r = drt.Renderer(flags=4)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8)
    r.ResizeBuffer(240, 135)
    r.Render(drt.Camera(), scene, 2)
    c = r.counters()
    rays = c["closest_rays"] + c["shadow_rays"]
    assert rays > 0 and c["node_visits"] >= rays and c["tri_tests"] > 0 and c["max_stack"] <= 32  # DRT_PSTACK of the production kernels
    print("room: %.2f wide-node visits/ray, %.2f triangle tests/ray" % (c["node_visits"] / rays, c["tri_tests"] / rays))


def test_cli_front_end_renders(drt, tmp_path):
    """tools/drt_render.cpp (include/drt.hpp over the C ABI): renders to a PPM, reports rays; the 8-bit image equals
    the API's DRT_RGBA8 readback of the same render."""
    import json
    import subprocess
    from dustraytracer_b200 import build
    exe = build.build_cli()
    out = str(tmp_path / "room.ppm")
    r = subprocess.run([exe, O.scene_path("room"), "--width", "160", "--height", "90", "--spp", "8", "--bounces", "8", "--sun", "--out", out],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["triangles"] == 330 and info["samples"] == 8 and info["rays"] > 160 * 90 * 8
    raw = open(out, "rb").read()
    assert raw.startswith(b"P6\n160 90\n255\n")
    img = np.frombuffer(raw[len(b"P6\n160 90\n255\n"):], np.uint8).reshape(90, 160, 3)
    rr = drt.Renderer()
    rr.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8, max_samples=1 << 30)
    rr.ResizeBuffer(160, 90)
    rr.Render(drt.Camera(), gpu_scene(drt, "room"), 8)
    api = rr.readback(drt.DRT_RGBA8)[::-1, :, :3]   # the PPM is written top-down
    assert np.array_equal(img, api)
    # --out x.png: same pixels through the built-in PNG writer (stored deflate), and --robust drops room.glb's 8 zero-area triangles
    import struct
    import zlib
    outp = str(tmp_path / "room.png")
    r = subprocess.run([exe, O.scene_path("room"), "--width", "160", "--height", "90", "--spp", "8", "--bounces", "8", "--sun", "--robust", "--out", outp],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert json.loads(r.stdout.strip().splitlines()[-1])["triangles"] == 322
    raw = open(outp, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat = 8, b""
    while pos < len(raw):
        n, ty = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(ty + body)
        if ty == b"IHDR":
            assert struct.unpack(">IIBBBBB", body) == (160, 90, 8, 2, 0, 0, 0)
        if ty == b"IDAT":
            idat += body
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(90, 1 + 160 * 3)
    assert np.all(rows[:, 0] == 0)
    png = rows[:, 1:].reshape(90, 160, 3)
    # zero-area triangles cannot be hit, so dropping them changes the tree but not the picture beyond sampling-order noise
    assert np.abs(png.astype(np.int32) - img.astype(np.int32)).mean() < 2.0


@pytest.mark.parametrize("name", ["cornell_box", "animemaid"])
def test_debug_views_bit_exact(drt, name):
    """RenderMode DEBUGMODE through the C ABI (RendererSettings.h:12-20; RayGen.cuh:137-161): albedo, normal,
    barycentric and uv views are bit-exact against the host-compiled reference; the BVH heat view runs and is finite
    (it counts the nodes of OUR tree, so it has no reference values)."""
    orc = O.best_oracle()
    if orc.kind != "reference":
        pytest.skip("needs oracle/_ref/libdrt_ref.so")
    W, H = 96, 64
    cam_f = O.default_camera_bytes()
    oh = orc.load(O.scene_path(name))
    scene = gpu_scene(drt, name)
    for mode in (0, 1, 2, 3):
        st_b = O.settings_bytes(sun=True, bounces=3, render_mode=1, debug_mode=mode)
        ref, _, _ = orc.render(oh, cam_f, st_b, W, H, 1, 2)
        r = drt.Renderer(flags=STRICT)
        r.m_RendererSettings = settings_from(drt, st_b)
        r.ResizeBuffer(W, H)
        r.Render(camera_from(drt, cam_f), scene, 2)
        got = r.readback(drt.DRT_ACCUM_RGB32F).reshape(-1)
        assert np.array_equal(bits(got), bits(ref)), "debug mode %d: max abs diff %g" % (mode, np.abs(got - ref).max())
        assert r.counters()["shadow_rays"] == 0
    r = drt.Renderer(flags=drt.DRT_FLAG_STATS)
    r.m_RendererSettings = settings_from(drt, O.settings_bytes(render_mode=1, debug_mode=4))
    r.ResizeBuffer(W, H)
    r.Render(camera_from(drt, cam_f), scene)
    heat = r.readback(drt.DRT_ACCUM_RGB32F)
    assert np.isfinite(heat).all() and heat.max() > 0.05
    orc.free(oh)


@pytest.mark.parametrize("name", ["animemaid", "mc_transparency_test"])
def test_opacity_classification_changes_nothing_but_the_work(drt, name):
    """Alpha-tested triangles proven always-opaque skip the test, never-hittable ones are collapsed (drt_host.h): the
    images, hit ids and ray counts are bit-identical with the classification on and off; only the number of alpha
    tests drops."""
    W, H, spp = 320, 180, 8
    out = []
    for on in (True, False):
        s = drt.Scene()
        s.loadGLTFmodel(O.scene_path(name))
        s.build_bvh(classify_alpha=on)
        r = drt.Renderer(flags=STRICT | drt.DRT_FLAG_STATS)
        r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=4, max_samples=1 << 20)
        r.ResizeBuffer(W, H)
        hits = r.trace_primary(drt.Camera(), s, 1)
        r.counters(reset=True)
        r.Render(drt.Camera(), s, spp)
        out.append((hits, r.readback(drt.DRT_ACCUM_RGB32F), r.counters(), s.stats()))
    (h1, a1, c1, st1), (h0, a0, c0, st0) = out
    for k in ("prim_id", "t", "u", "v"):
        assert np.array_equal(bits(h1[k]), bits(h0[k]))
    assert np.array_equal(bits(a1), bits(a0))
    assert c1["closest_rays"] == c0["closest_rays"] and c1["shadow_rays"] == c0["shadow_rays"]
    print("%s: alpha tests %d -> %d (%d of %d alpha-tested triangles proven opaque, %d hidden)"
          % (name, c0["alpha_tests"], c1["alpha_tests"], st1["alpha_opaque"], st1["alpha_tris"], st1["alpha_hidden"]))
    # (the builder's partition pass orders the primitives of a node by atomic arrival, so the any-hit traversal order --
    # and with it the number of alpha tests before the first accepted hit -- varies by a few tests from build to build)
    assert c1["alpha_tests"] <= c0["alpha_tests"] * 1.001 + 64
    if name == "animemaid":
        assert c1["alpha_tests"] < 0.9 * c0["alpha_tests"]
