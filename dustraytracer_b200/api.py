"""Host-side mirror of the reference's library surface, over the C ABI (include/drt.h).

Same class and method names, argument meaning and call order as the reference's only caller uses
(DustRayTracer/src/Editor/EditorLayer.cpp:35-55,317-318,424):

    cam = Camera(); scene = Scene(); scene.loadGLTFmodel(path)
    BVHBuilder().buildIterative(scene)            # BVHBuilder.cuh:21-22 (GPU binned-SAH build here)
    r = Renderer(); r.ResizeBuffer(w, h); r.Render(cam, scene)   # Renderer.hpp:18-30
    r.resetAccumulationBuffer(); r.getSampleCount()

The GL texture (GetRenderTargetImage_name) is replaced by readback of the device frame buffer.
Everything runs on the GPU through libdrt.so; there is no CPU path in this package.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import (DRT_ACCUM_RGB32F, DRT_FLAG_EXTENDED_MATERIALS, DRT_FLAG_HOST_ARG_ORDER, DRT_FLAG_STATS, DRT_FLAG_STRICT, DRT_RGBA8, DRT_RGBA32F,
                   DrtError, check, drt_bvh_params, drt_camera, drt_counters, drt_hit, drt_material, drt_scene_stats,
                   drt_settings)


class RendererSettings(drt_settings):
    """RendererSettings POD with the reference defaults (Core/Scene/RendererSettings.h:22-34)."""
    NORMALMODE, DEBUGMODE = 0, 1
    ALBEDO_DEBUG, NORMAL_DEBUG, BARYCENTRIC_DEBUG, UVS_DEBUG, MESHBVH_DEBUG, WORLDBVH_DEBUG = range(6)

    def __init__(self, **kw):
        super().__init__()
        _lib.lib().drt_default_settings(ctypes.byref(self))
        for k, v in kw.items():
            _assign(self, k, v)


class Camera(drt_camera):
    """Camera POD with the reference defaults (Core/Scene/Camera.cuh:17,32-46; EditorLayer.cpp:35-40)."""

    def __init__(self, pos=None, **kw):
        super().__init__()
        _lib.lib().drt_default_camera(ctypes.byref(self))
        if pos is not None:
            self.position = (ctypes.c_float * 3)(*pos)
        for k, v in kw.items():
            _assign(self, k, v)

    # reference field names
    @property
    def m_Position(self):
        return tuple(self.position)

    @m_Position.setter
    def m_Position(self, v):
        self.position = (ctypes.c_float * 3)(*v)

    @property
    def m_Forward_dir(self):
        return tuple(self.forward_dir)

    @m_Forward_dir.setter
    def m_Forward_dir(self, v):
        self.forward_dir = (ctypes.c_float * 3)(*v)

    def GetPosition(self):
        return tuple(self.position)


def _assign(struct, name, value):
    cur = getattr(struct, name)
    if isinstance(cur, ctypes.Array):
        setattr(struct, name, type(cur)(*value))
    else:
        setattr(struct, name, value)


class Scene:
    """struct Scene (Core/Scene/Scene.cuh:41-57): owns the triangle soup, materials, textures and the BVH."""

    def __init__(self):
        self._h = ctypes.c_void_p()

    def loadGLTFmodel(self, filepath, robust=False):
        """Scene::loadGLTFmodel (Scene.cu:181-317). Returns True; raises DrtError on failure
        (the reference returns a bool its caller ignores). robust=True reads the file as the glTF 2.0
        specification says instead of as the reference loader does (drt_scene_load_gltf_ex, DRT_LOAD_ROBUST)."""
        self._free()
        check(_lib.lib().drt_scene_load_gltf_ex(str(filepath).encode(), 1 if robust else 0, ctypes.byref(self._h)))
        return True

    @classmethod
    def from_triangles(cls, positions, uvs=None, material_index=None, materials=None):
        """Synthetic scenes: positions [N,3,3] f32, uvs [N,3,2] f32, material_index [N] i32."""
        s = cls()
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 9)
        n = pos.shape[0]
        uv = None if uvs is None else np.ascontiguousarray(uvs, np.float32).reshape(n, 6)
        mi = None if material_index is None else np.ascontiguousarray(material_index, np.int32).reshape(n)
        if materials is None:
            m = drt_material()
            m.albedo = (ctypes.c_float * 3)(0.8, 0.8, 0.8)
            m.albedo_texture_index = -1
            m.refractive_index = 1.45
            materials = [m]
        marr = (drt_material * len(materials))(*materials)
        check(_lib.lib().drt_scene_from_triangles(
            pos.ctypes.data, None if uv is None else uv.ctypes.data, None if mi is None else mi.ctypes.data, n,
            ctypes.cast(marr, ctypes.c_void_p), len(materials), ctypes.byref(s._h)))
        return s

    def add_texture(self, pixels):
        px = np.ascontiguousarray(pixels, np.uint8)
        h, w, c = px.shape
        check(_lib.lib().drt_scene_add_texture(self._h, w, h, c, px.ctypes.data))

    def build_bvh(self, device=0, bin_count=8, leaf_target=20, classify_alpha=True, single_launch=True):
        """single_launch=False forces the level-synchronous multi-launch builder also for small scenes (same tree; A/B)."""
        p = drt_bvh_params(bin_count=bin_count, leaf_target=leaf_target)
        p.reserved[0] = (0 if classify_alpha else 1) | (0 if single_launch else 2)
        check(_lib.lib().drt_scene_build_bvh(self._h, ctypes.byref(p), device))

    def stats(self):
        st = drt_scene_stats()
        check(_lib.lib().drt_scene_get_stats(self._h, ctypes.byref(st)))
        return {k: getattr(st, k) for k, _ in drt_scene_stats._fields_}

    def triangles(self):
        """[N,32] f32: the soup in load order, laid out like the reference's 128-byte Triangle."""
        n = self.stats()["triangles"]
        out = np.zeros((n, 32), np.float32)
        check(_lib.lib().drt_scene_get_triangles(self._h, out.ctypes.data, out.nbytes))
        return out

    def materials(self):
        n = self.stats()["materials"]
        arr = (drt_material * max(n, 1))()
        check(_lib.lib().drt_scene_get_materials(self._h, ctypes.cast(arr, ctypes.c_void_p), n))
        return list(arr)[:n]

    def texture(self, index):
        w, h, c = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        check(_lib.lib().drt_scene_get_texture(self._h, index, ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), None, 0))
        px = np.zeros((h.value, w.value, c.value), np.uint8)
        check(_lib.lib().drt_scene_get_texture(self._h, index, None, None, None, px.ctypes.data, px.nbytes))
        return px

    def wide_nodes(self):
        n = self.stats()["wide_nodes"]
        out = np.zeros((n, 80), np.uint8)
        check(_lib.lib().drt_scene_get_wide_nodes(self._h, out.ctypes.data, out.nbytes))
        return out

    def leaf_triangles(self):
        n = self.stats()["triangles"]
        out = np.zeros((n, 12), np.float32)
        check(_lib.lib().drt_scene_get_leaf_triangles(self._h, out.ctypes.data, out.nbytes))
        return out

    def _free(self):
        if self._h:
            _lib.lib().drt_scene_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self._free()
        except Exception:
            pass


def decode_image(data):
    """stbi_load_from_memory(data, len, &w, &h, &comps, 0) of the reference (Texture.cu:21-30): [H, W, C] uint8."""
    buf = bytes(data)
    w, h, c = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    check(_lib.lib().drt_decode_image(buf, len(buf), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c), None, 0))
    out = np.zeros((h.value, w.value, c.value), np.uint8)
    check(_lib.lib().drt_decode_image(buf, len(buf), None, None, None, out.ctypes.data, out.nbytes))
    return out


class BVHBuilder:
    """class BVHBuilder (Core/BVH/BVHBuilder.cuh:12-23). Both entry points run the GPU binned-SAH build and
    emit compressed 8-wide nodes; the two knobs are kept for call compatibility (see drt_bvh_params)."""

    def __init__(self, device=0):
        self.m_BinCount = 8
        self.m_TargetLeafPrimitivesCount = 6
        self.device = device

    def build(self, scene):
        scene.build_bvh(self.device, self.m_BinCount, self.m_TargetLeafPrimitivesCount)
        return scene

    buildIterative = build


DRT_FRAME_HANDLE_BYTES = 160


class Renderer:
    """class Renderer (Core/Renderer.hpp:14-47), headless: a device float4 frame buffer replaces the GL texture."""

    def __init__(self, device=0, flags=0):
        self._h = ctypes.c_void_p()
        self.device = device
        check(_lib.lib().drt_renderer_create(device, ctypes.byref(self._h)))
        self.m_RendererSettings = RendererSettings()
        if flags:
            self.set_flags(flags)

    def ResizeBuffer(self, width, height):
        check(_lib.lib().drt_renderer_resize(self._h, width, height))

    def Render(self, cam, scene, n_samples=1):
        """Renderer::Render (Renderer.cu:80-117); returns the device time in ms (the reference's *delta).
        n_samples > 1 batches that many frames into one launch."""
        ms = ctypes.c_float()
        check(_lib.lib().drt_renderer_render(self._h, ctypes.byref(cam), scene._h, ctypes.byref(self.m_RendererSettings),
                                             n_samples, ctypes.byref(ms)))
        return ms.value

    def resetAccumulationBuffer(self):
        check(_lib.lib().drt_renderer_reset(self._h))

    def getSampleCount(self):
        return _lib.lib().drt_renderer_sample_count(self._h)

    def getBufferWidth(self):
        return _lib.lib().drt_renderer_width(self._h)

    def getBufferHeight(self):
        return _lib.lib().drt_renderer_height(self._h)

    # ---- beyond the reference surface ----
    def set_flags(self, flags):
        check(_lib.lib().drt_renderer_set_flags(self._h, flags))

    def set_tuning(self, variant=-1, block_threads=0):
        check(_lib.lib().drt_renderer_set_tuning(self._h, variant, block_threads))

    def set_stream(self, cuda_stream_ptr):
        check(_lib.lib().drt_renderer_set_stream(self._h, ctypes.c_void_p(cuda_stream_ptr)))

    def set_frame_schedule(self, first_frame, stride):
        check(_lib.lib().drt_renderer_set_frame_schedule(self._h, first_frame, stride))

    def export_frame_buffers(self):
        """Bytes another process hands to import_frame_buffers to render its tiles directly into THIS renderer's buffers."""
        buf = ctypes.create_string_buffer(DRT_FRAME_HANDLE_BYTES)
        check(_lib.lib().drt_renderer_export_frame_buffers(self._h, buf))
        return bytes(buf.raw)

    def import_frame_buffers(self, handle):
        buf = ctypes.create_string_buffer(bytes(handle), DRT_FRAME_HANDLE_BYTES)
        check(_lib.lib().drt_renderer_import_frame_buffers(self._h, buf))

    def attach_frame_buffers(self, owner):
        """Same process: render into `owner`'s buffers (None detaches)."""
        check(_lib.lib().drt_renderer_attach_frame_buffers(self._h, owner._h if owner is not None else None))
        self._fb_owner = owner  # keep the buffers alive

    def set_tile_schedule(self, first_tile, stride):
        """Image-tile sharding: render only the 8x4 tiles first_tile, first_tile+stride, ... (row-major tile index)."""
        check(_lib.lib().drt_renderer_set_tile_schedule(self._h, first_tile, stride))

    def readback(self, fmt=DRT_RGBA32F, out=None):
        w, h = self.getBufferWidth(), self.getBufferHeight()
        shape, dt = {DRT_RGBA32F: ((h, w, 4), np.float32), DRT_RGBA8: ((h, w, 4), np.uint8),
                     DRT_ACCUM_RGB32F: ((h, w, 3), np.float32)}[fmt]
        if out is None:
            out = np.empty(shape, dt)
        check(_lib.lib().drt_renderer_readback(self._h, fmt, out.ctypes.data, out.nbytes))
        return out

    def upload_accum(self, rgb, samples_done):
        a = np.ascontiguousarray(rgb, np.float32)
        check(_lib.lib().drt_renderer_upload_accum(self._h, a.ctypes.data, a.nbytes, samples_done))

    def accum_device_ptr(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        check(_lib.lib().drt_renderer_accum_device_ptr(self._h, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def resolve(self, total_samples):
        check(_lib.lib().drt_renderer_resolve(self._h, total_samples))

    def counters(self, reset=False):
        c = drt_counters()
        check(_lib.lib().drt_renderer_get_counters(self._h, ctypes.byref(c), 1 if reset else 0))
        return {k: getattr(c, k) for k, _ in drt_counters._fields_}

    def kernel_launches(self):
        return _lib.lib().drt_renderer_kernel_launches(self._h)

    def last_variant(self):
        """Kernel variant the last Render ran (automatic selection resolved)."""
        return _lib.lib().drt_renderer_last_variant(self._h)

    def set_accumulate_add(self, on=True):
        """Fused render + reduce (drt_renderer_set_accumulate_add): launches add their pixel sums into the (shared) accumulation
        buffer with fp32 atomics; the owner of the buffer resolves."""
        check(_lib.lib().drt_renderer_set_accumulate_add(self._h, 1 if on else 0))

    def last_launch_ordered(self):
        """True if the last render launch ran its blocks in cost order (second and later batched launches of a view)."""
        return bool(_lib.lib().drt_renderer_last_launch_ordered(self._h))

    def trace_primary(self, cam, scene, frame_index=1):
        """Parity probe: structured array (prim_id, t, u, v) per pixel, p = x + y*W."""
        w, h = self.getBufferWidth(), self.getBufferHeight()
        out = np.zeros(w * h, dtype=[("prim_id", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32)])
        check(_lib.lib().drt_trace_primary(self._h, ctypes.byref(cam), scene._h, frame_index, out.ctypes.data, out.size))
        return out

    def trace_rays(self, scene, origins, directions, anyhit=False):
        """Ray-level parity probe (drt_trace_rays): structured array (prim_id, t, u, v) per ray."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(directions, np.float32).reshape(-1, 3)
        assert o.shape == d.shape
        out = np.zeros(o.shape[0], dtype=[("prim_id", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32)])
        check(_lib.lib().drt_trace_rays(self._h, scene._h, o.ctypes.data, d.ctypes.data, o.shape[0], 1 if anyhit else 0, out.ctypes.data))
        return out

    def trace_primary_noalloc(self, cam, scene, frame_index=1):
        """trace_primary through drt_trace_primary_noalloc: device staging in libdrt.so's static scratch (no cudaMalloc);
        width * height <= NOALLOC_MAX_PIXELS."""
        w, h = self.getBufferWidth(), self.getBufferHeight()
        out = np.zeros(w * h, dtype=[("prim_id", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32)])
        check(_lib.lib().drt_trace_primary_noalloc(self._h, ctypes.byref(cam), scene._h, frame_index, out.ctypes.data, out.size))
        return out

    def __del__(self):
        try:
            if self._h:
                _lib.lib().drt_renderer_destroy(self._h)
                self._h = ctypes.c_void_p()
        except Exception:
            pass


class MultiRenderer:
    """N GPUs of one node behind one handle (drt_multi_*): replicated scene, sample-range or image-tile sharding, one NCCL
    reduce of the accumulation buffers per render call, all inside libdrt.so (no torch.distributed)."""

    def __init__(self, devices, flags=0):
        self._h = ctypes.c_void_p()
        arr = (ctypes.c_int32 * len(devices))(*devices)
        check(_lib.lib().drt_multi_create(arr, len(devices), ctypes.byref(self._h)))
        self.devices = list(devices)
        self.m_RendererSettings = RendererSettings()
        if flags:
            self.set_flags(flags)

    def set_flags(self, flags):
        check(_lib.lib().drt_multi_set_flags(self._h, flags))

    def set_scene(self, scene):
        check(_lib.lib().drt_multi_set_scene(self._h, scene._h, None))

    def set_sharding(self, tiles=False):
        check(_lib.lib().drt_multi_set_sharding(self._h, 1 if tiles else 0))

    def ResizeBuffer(self, width, height):
        check(_lib.lib().drt_multi_resize(self._h, width, height))
        self._size = (width, height)

    def Render(self, cam, n_samples=1):
        ms = ctypes.c_float()
        check(_lib.lib().drt_multi_render(self._h, ctypes.byref(cam), ctypes.byref(self.m_RendererSettings), n_samples, ctypes.byref(ms)))
        return ms.value

    def resetAccumulationBuffer(self):
        check(_lib.lib().drt_multi_reset(self._h))

    def getSampleCount(self):
        return _lib.lib().drt_multi_sample_count(self._h)

    def nccl_reduces(self):
        return _lib.lib().drt_multi_nccl_reduces(self._h)

    def readback(self, fmt=DRT_RGBA32F):
        w, h = self._size
        shape, dt = {DRT_RGBA32F: ((h, w, 4), np.float32), DRT_RGBA8: ((h, w, 4), np.uint8), DRT_ACCUM_RGB32F: ((h, w, 3), np.float32)}[fmt]
        out = np.empty(shape, dt)
        check(_lib.lib().drt_multi_readback(self._h, fmt, out.ctypes.data, out.nbytes))
        return out

    def counters(self, reset=False):
        c = drt_counters()
        check(_lib.lib().drt_multi_get_counters(self._h, ctypes.byref(c), 1 if reset else 0))
        return {k: getattr(c, k) for k, _ in drt_counters._fields_}

    def __del__(self):
        try:
            if self._h:
                _lib.lib().drt_multi_destroy(self._h)
                self._h = ctypes.c_void_p()
        except Exception:
            pass


def tile_owner_mask(width, height, first_tile, stride):
    """Boolean (height, width) mask of the pixels a renderer with set_tile_schedule(first_tile, stride) renders: the 8x4
    tiles whose row-major index t satisfies t >= first_tile and (t - first_tile) % stride == 0 (drt_kernels.cu pixel_of_thread)."""
    import numpy as np
    tiles_x = (width + 7) // 8
    ys, xs = np.mgrid[0:height, 0:width]
    t = (ys // 4) * tiles_x + xs // 8
    return (t >= first_tile) & ((t - first_tile) % stride == 0)
