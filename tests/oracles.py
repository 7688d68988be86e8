"""ctypes front-ends of the CHECKER libraries (test infrastructure only):

  RefOracle   oracle/_ref/libdrt_ref.so      the reference's own sources compiled host-side (needs the
                                             prebuilt .so; it is built in the dev container from /root/reference)
  PortOracle  oracle/_build/libdrt_oracle.so our plain-C restatement (builds anywhere gcc is)
  HostEmu     tests/hostemu/_build/...       the device headers compiled as plain C++ (logic tests, no GPU)

best_oracle() returns the strongest checker available on this machine (reference first).
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libdrt_ref.so")
PORT_LIB = os.path.join(ROOT, "oracle", "_build", "libdrt_oracle.so")
EMU_LIB = os.path.join(ROOT, "tests", "hostemu", "_build", "libdrt_hostemu.so")

_vp = ctypes.c_void_p


def P(a):
    return a.ctypes.data_as(_vp)


def scene_path(name):
    return os.path.join(SCENES, name if name.endswith(".glb") else name + ".glb")


def default_camera_bytes():
    """The reference's default Camera POD (Camera.cuh:32-46; EditorLayer.cpp:35-40), 88 bytes."""
    f = np.zeros(22, np.float32)
    f[0] = 1.0
    f[1] = np.float32(60.0) * (np.float32(3.14159265359) / np.float32(180.0))
    f[6] = 10.0
    f[7] = 10.0
    f[8:11] = (0, 2, 5)
    f[11:14] = (0, 0, -1)
    f[14:17] = (0, 1, 0)
    f[17:20] = (1, -0.0, 0)
    return f


def settings_bytes(sun=False, bounces=2, max_samples=500, gamma=True, tonemap=True, render_mode=0, debug_mode=0):
    """RendererSettings POD (RendererSettings.h:22-34), 64 bytes, padding zeroed."""
    b = np.zeros(64, np.uint8)
    b[0], b[1], b[2] = int(gamma), int(tonemap), int(sun)
    i = b.view(np.int32)
    i[1], i[2], i[3], i[4] = max_samples, bounces, render_mode, debug_mode
    f = b.view(np.float32)
    f[6:8] = (-0.803, 0.681)
    f[8:11] = (1.000, 0.944, 0.917)
    f[11] = 30
    f[12:15] = (0.25, 0.498, 0.80)
    f[15] = 20
    return b


class RefOracle:
    kind = "reference"

    def __init__(self):
        L = ctypes.CDLL(REF_LIB)
        L.ref_scene_load.restype = _vp
        L.ref_scene_load.argtypes = [ctypes.c_char_p]
        L.ref_render.restype = ctypes.c_double
        L.ref_pcg_hash.restype = ctypes.c_uint
        L.ref_random_float.restype = ctypes.c_float
        L.ref_bounds_intersect.restype = ctypes.c_float
        L.ref_camera_get_ray.restype = ctypes.c_uint
        self.L = L

    @staticmethod
    def available():
        return os.path.exists(REF_LIB)

    def max_threads(self):
        return self.L.ref_max_threads()

    def load(self, path, bins=8, leaf=20, iterative=False):
        h = _vp(self.L.ref_scene_load(path.encode()))
        sec = ctypes.c_double()
        self.L.ref_scene_build_bvh(h, bins, leaf, int(iterative), ctypes.byref(sec))
        return h

    def free(self, h):
        self.L.ref_scene_free(h)

    def triangles_load_order(self, h):
        n = self.L.ref_scene_tri_count(h)
        t = np.zeros((n, 32), np.float32)
        self.L.ref_scene_get_triangles_load_order(h, P(t))
        t[:, 3] = 0
        return t

    def materials(self, h):
        n = self.L.ref_scene_material_count(h)
        m = np.zeros((n, 11), np.uint32)
        self.L.ref_scene_get_materials(h, P(m))
        m[:, 8] &= 0xFF
        m[:, 10] &= 0xFF
        return m

    def textures(self, h):
        out = []
        for i in range(self.L.ref_scene_texture_count(h)):
            w, hh, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
            self.L.ref_scene_get_texture_info(h, i, ctypes.byref(w), ctypes.byref(hh), ctypes.byref(c))
            px = np.zeros((hh.value, w.value, c.value), np.uint8)
            self.L.ref_scene_get_texture_pixels(h, i, P(px))
            out.append(px)
        return out

    def nodes(self, h):
        n = self.L.ref_scene_node_count(h)
        a = np.zeros((n, 11), np.uint32)
        self.L.ref_scene_get_nodes(h, P(a))
        a[:, 0] &= 0xFF
        return a

    def canonical_ids(self, h):
        n = self.L.ref_scene_tri_count(h)
        a = np.zeros(n, np.int32)
        self.L.ref_scene_get_canonical_ids(h, P(a))
        return a

    def trace_primary(self, h, cam, W, H, frameidx=1, nthreads=0):
        pid = np.zeros(W * H, np.int32)
        t = np.zeros(W * H, np.float32)
        u = np.zeros_like(t)
        v = np.zeros_like(t)
        self.L.ref_trace_primary(h, P(cam), W, H, frameidx, P(pid), P(t), P(u), P(v), None, nthreads)
        return pid, t, u, v

    def render(self, h, cam, settings, W, H, frame_begin=1, frames=1, acc=None, nthreads=0):
        if acc is None:
            acc = np.zeros(W * H * 3, np.float32)
        cnt = (ctypes.c_ulonglong * 2)()
        sec = self.L.ref_render(h, P(cam), P(settings), W, H, frame_begin, frames, P(acc), nthreads, cnt)
        return acc, (cnt[0], cnt[1]), sec


class PortOracle:
    kind = "port"

    def __init__(self):
        if not os.path.exists(PORT_LIB):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"], stdout=subprocess.DEVNULL)
        L = ctypes.CDLL(PORT_LIB)
        L.orc_scene_create.restype = _vp
        L.orc_render.restype = ctypes.c_double
        L.orc_pcg_hash.restype = ctypes.c_uint
        L.orc_random_float.restype = ctypes.c_float
        L.orc_bounds_intersect.restype = ctypes.c_float
        self.L = L

    @staticmethod
    def available():
        return True

    def max_threads(self):
        return self.L.orc_max_threads()

    def create(self, tris, mats, textures, bins=8, leaf=20, rtl=True):
        """tris [N,32] f32 (reference Triangle layout, load order), mats [M,11] u32, textures list of HxWxC u8."""
        tris = np.ascontiguousarray(tris, np.float32)
        mats = np.ascontiguousarray(mats, np.uint32)
        h = _vp(self.L.orc_scene_create(P(tris), tris.shape[0], P(mats), mats.shape[0], int(rtl)))
        for px in textures:
            px = np.ascontiguousarray(px, np.uint8)
            self.L.orc_scene_add_texture(h, px.shape[1], px.shape[0], px.shape[2], P(px))
        self.L.orc_scene_build_bvh(h, bins, leaf)
        return h

    def free(self, h):
        self.L.orc_scene_free(h)

    def nodes(self, h):
        n = self.L.orc_scene_node_count(h)
        a = np.zeros((n, 11), np.uint32)
        self.L.orc_scene_get_nodes(h, P(a))
        a[:, 0] &= 0xFF
        return a

    def canonical_ids(self, h, n):
        a = np.zeros(n, np.int32)
        self.L.orc_scene_get_canonical_ids(h, P(a))
        return a

    def trace_primary(self, h, cam, W, H, frameidx=1, nthreads=0):
        pid = np.zeros(W * H, np.int32)
        t = np.zeros(W * H, np.float32)
        u = np.zeros_like(t)
        v = np.zeros_like(t)
        self.L.orc_trace_primary(h, P(cam), W, H, frameidx, P(pid), P(t), P(u), P(v), nthreads)
        return pid, t, u, v

    def camera_rays(self, cam, W, H, frameidx=1, rtl=True):
        """Primary rays as RayGen issues them: [W*H, 6] origin xyz, direction xyz (Camera.cu:82-123)."""
        out = np.zeros((W * H, 6), np.float32)
        self.L.orc_camera_rays(P(cam), W, H, frameidx, int(rtl), P(out))
        return out

    def brute_force_closest(self, tris9, rays_od6, nthreads=0):
        """Closest hit of every ray over ALL triangles with the reference's Moller-Trumbore, no tree (lowest index wins ties)."""
        tris9 = np.ascontiguousarray(tris9, np.float32).reshape(-1, 9)
        rays = np.ascontiguousarray(rays_od6, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        pid = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        u = np.zeros_like(t)
        v = np.zeros_like(t)
        self.L.orc_brute_force_closest(P(tris9), tris9.shape[0], P(rays), n, P(pid), P(t), P(u), P(v), nthreads)
        return pid, t, u, v

    def render(self, h, cam, settings, W, H, frame_begin=1, frames=1, acc=None, nthreads=0):
        if acc is None:
            acc = np.zeros(W * H * 3, np.float32)
        cnt = (ctypes.c_ulonglong * 2)()
        sec = self.L.orc_render(h, P(cam), P(settings), W, H, frame_begin, frames, P(acc), nthreads, cnt)
        return acc, (cnt[0], cnt[1]), sec


class HostEmu:
    def __init__(self):
        if not os.path.exists(EMU_LIB):
            import sys
            sys.path.insert(0, ROOT)
            from dustraytracer_b200 import build
            build.build_hostemu()
        L = ctypes.CDLL(EMU_LIB)
        L.emu_scene_load.restype = _vp
        L.emu_scene_load.argtypes = [ctypes.c_char_p]
        L.emu_scene_from_triangles.restype = _vp
        L.emu_last_error.restype = ctypes.c_char_p
        self.L = L

    def load(self, path):
        h = self.L.emu_scene_load(path.encode())
        if not h:
            raise RuntimeError(self.L.emu_last_error().decode())
        return _vp(h)

    def from_triangles(self, pos, mats44):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
        return _vp(self.L.emu_scene_from_triangles(P(pos), None, None, pos.shape[0], P(mats44), mats44.shape[0]))

    def free(self, h):
        self.L.emu_scene_free(h)

    def info(self, h):
        a = (ctypes.c_longlong * 8)()
        self.L.emu_scene_info(h, a)
        return dict(zip(("tris", "materials", "textures", "wide_nodes", "bvh2_nodes", "levels", "wide_levels", "tris_emitted"), a))

    def triangles(self, h):
        n = self.info(h)["tris"]
        t = np.zeros((n, 32), np.float32)
        self.L.emu_scene_get_triangles(h, P(t))
        return t

    def materials(self, h):
        n = self.info(h)["materials"]
        m = np.zeros((n, 11), np.uint32)
        self.L.emu_scene_get_materials(h, P(m))
        return m

    def textures(self, h):
        out = []
        for i in range(self.info(h)["textures"]):
            w, hh, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
            self.L.emu_scene_get_texture_info(h, i, ctypes.byref(w), ctypes.byref(hh), ctypes.byref(c))
            px = np.zeros((hh.value, w.value, c.value), np.uint8)
            self.L.emu_scene_get_texture_pixels(h, i, P(px))
            out.append(px)
        return out

    def wide_nodes(self, h):
        a = np.zeros((self.info(h)["wide_nodes"], 80), np.uint8)
        self.L.emu_scene_get_wide_nodes(h, P(a))
        return a

    def leaf_triangles(self, h):
        a = np.zeros((self.info(h)["tris"], 12), np.float32)
        self.L.emu_scene_get_leaf_triangles(h, P(a))
        return a

    def trace_primary(self, h, cam, W, H, frameidx=1, strict=True, rtl=True):
        hits = np.zeros((W * H, 4), np.float32)
        stats = (ctypes.c_ulonglong * 6)()
        self.L.emu_trace_primary(h, P(cam), W, H, frameidx, int(strict), int(rtl), P(hits), None, stats)
        return hits.view(np.int32)[:, 0].copy(), hits[:, 1].copy(), hits[:, 2].copy(), hits[:, 3].copy(), list(stats)

    def render(self, h, cam, settings, W, H, first=1, stride=1, n=1, strict=True, rtl=True, flat=False, acc=None, steps=None, steps5=None,
               ext=False):
        """steps: None, or the node format (0 compressed / 1 expanded) for the phased kernel's step functions.
        ext: the extended material branch (DRT_FLAG_EXTENDED_MATERIALS)."""
        if acc is None:
            acc = np.zeros(W * H * 3, np.float32)
        cnt = (ctypes.c_ulonglong * 8)()
        if ext:
            self.L.emu_render_ext(h, P(cam), P(settings), W, H, first, stride, n, int(strict), int(rtl), P(acc), cnt)
            return acc, list(cnt)
        if steps5 is not None:
            fmt, nodes_first = steps5
            self.L.emu_render_steps5(h, P(cam), P(settings), W, H, first, stride, n, int(strict), int(rtl), int(fmt), int(nodes_first), P(acc), cnt)
            return acc, list(cnt)
        if steps is not None:
            self.L.emu_render_steps(h, P(cam), P(settings), W, H, first, stride, n, int(strict), int(rtl), int(steps), P(acc), cnt)
            return acc, list(cnt)
        fn = self.L.emu_render_flat if flat else self.L.emu_render
        fn(h, P(cam), P(settings), W, H, first, stride, n, int(strict), int(rtl), P(acc), cnt)
        return acc, list(cnt)


def best_oracle():
    return RefOracle() if RefOracle.available() else PortOracle()


def pcg_hash_np(v):
    """Random.cu:6-11 on uint32 arrays."""
    v = v.astype(np.uint32)
    with np.errstate(over="ignore"):
        state = v * np.uint32(747796405) + np.uint32(2891336453)
        word = ((state >> ((state >> np.uint32(28)) + np.uint32(4))) ^ state) * np.uint32(277803737)
    return (word >> np.uint32(22)) ^ word


def soup_triangles(n, seed=0x0D057A11):
    """SURVEY.md 8d synthetic random-triangle soup (BASELINE config 5): counter-based generator on the reference's pcg_hash,
    centres uniform in [-50, 50]^3, vertices within +-0.25 of the centre. The same function as tools/run_configs.py soup()."""
    out = np.empty((n, 9), np.float32)
    chunk = 1 << 20
    for a in range(0, n, chunk):
        i = np.arange(a, min(n, a + chunk), dtype=np.uint32)
        base = pcg_hash_np(np.uint32(seed) ^ i)
        with np.errstate(over="ignore"):
            r = np.stack([pcg_hash_np(base * np.uint32(12) + np.uint32(k)) for k in range(12)], 1).astype(np.float64) / 4294967296.0
        c = r[:, 0:3] * 100.0 - 50.0
        for j in range(3):
            out[a:a + len(i), 3 * j:3 * j + 3] = (c + (r[:, 3 + 3 * j:6 + 3 * j] - 0.5) * 0.5).astype(np.float32)
    return out


def oracle_scene_for(name_or_path, orc=None):
    """(oracle, handle) of the strongest checker for a bundled scene: the reference's own loader + builder when its library
    travelled, else the C port fed by the product's HOST loader (triangle / material / texel CRCs are pinned in test_abi)."""
    orc = orc or best_oracle()
    path = name_or_path if os.path.sep in name_or_path else scene_path(name_or_path)
    if orc.kind == "reference":
        return orc, orc.load(path)
    import dustraytracer_b200 as d
    s = d.Scene()
    s.loadGLTFmodel(path)
    mats = np.frombuffer(b"".join(bytes(m) for m in s.materials()), np.uint32).reshape(-1, 11).copy()
    return orc, orc.create(s.triangles(), mats, [s.texture(i) for i in range(s.stats()["textures"])])
