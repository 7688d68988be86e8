"""ctypes binding of libdrt.so (the C ABI in include/drt.h). No fallback: a missing library is an error."""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DRT_LIB_PATH") or os.path.join(_PKG, "libdrt.so")  # (DRT_LIB_PATH: A/B measurement builds only)


class drt_camera(ctypes.Structure):
    """Field-for-field the reference Camera POD (Core/Scene/Camera.cuh:32-46), 88 bytes."""
    _fields_ = [
        ("exposure", ctypes.c_float), ("vfov_rad", ctypes.c_float), ("zfar", ctypes.c_float), ("znear", ctypes.c_float),
        ("aspect_ratio", ctypes.c_float), ("defocus_angle", ctypes.c_float), ("focus_dist", ctypes.c_float),
        ("movement_speed", ctypes.c_float),
        ("position", ctypes.c_float * 3), ("forward_dir", ctypes.c_float * 3),
        ("up_dir", ctypes.c_float * 3), ("right_dir", ctypes.c_float * 3),
        ("viewheight", ctypes.c_uint32), ("viewwidth", ctypes.c_uint32),
    ]


class drt_settings(ctypes.Structure):
    """Field-for-field the reference RendererSettings (Core/Scene/RendererSettings.h:22-34), 64 bytes."""
    _fields_ = [
        ("gamma_correction", ctypes.c_uint8), ("tone_mapping", ctypes.c_uint8), ("enable_sunlight", ctypes.c_uint8),
        ("_pad0", ctypes.c_uint8),
        ("max_samples", ctypes.c_int32), ("ray_bounce_limit", ctypes.c_int32),
        ("render_mode", ctypes.c_int32), ("debug_mode", ctypes.c_int32), ("_pad1", ctypes.c_int32),
        ("sunlight_dir", ctypes.c_float * 2), ("sunlight_color", ctypes.c_float * 3), ("sunlight_intensity", ctypes.c_float),
        ("sky_color", ctypes.c_float * 3), ("sky_intensity", ctypes.c_float),
    ]


class drt_material(ctypes.Structure):
    _fields_ = [
        ("albedo", ctypes.c_float * 3), ("emissive_factor", ctypes.c_float * 3), ("albedo_texture_index", ctypes.c_int32),
        ("roughness", ctypes.c_float), ("transmission", ctypes.c_uint8), ("_pad0", ctypes.c_uint8 * 3),
        ("refractive_index", ctypes.c_float), ("metallic", ctypes.c_uint8), ("_pad1", ctypes.c_uint8 * 3),
    ]


class drt_bvh_params(ctypes.Structure):
    _fields_ = [("bin_count", ctypes.c_int32), ("leaf_target", ctypes.c_int32), ("reserved", ctypes.c_int32 * 6)]


class drt_scene_stats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in
                ("triangles", "materials", "textures", "meshes", "wide_nodes", "bvh2_nodes", "bvh_levels", "device_bytes")] + \
               [("build_ms", ctypes.c_double), ("load_ms", ctypes.c_double)] + \
               [(n, ctypes.c_uint64) for n in ("alpha_tris", "alpha_opaque", "alpha_hidden", "build_launches")]


class drt_hit(ctypes.Structure):
    _fields_ = [("prim_id", ctypes.c_int32), ("t", ctypes.c_float), ("u", ctypes.c_float), ("v", ctypes.c_float)]


class drt_counters(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in
                ("closest_rays", "shadow_rays", "node_visits", "tri_tests", "tie_events", "alpha_tests", "max_stack", "texel_clamps")]


assert ctypes.sizeof(drt_camera) == 88 and ctypes.sizeof(drt_settings) == 64 and ctypes.sizeof(drt_material) == 44

DRT_RGBA32F, DRT_RGBA8, DRT_ACCUM_RGB32F = 0, 1, 2
DRT_FLAG_STRICT, DRT_FLAG_HOST_ARG_ORDER, DRT_FLAG_STATS, DRT_FLAG_EXTENDED_MATERIALS, DRT_FLAG_WATERTIGHT, DRT_FLAG_NO_TILE_ORDER = 1, 2, 4, 8, 16, 32

_vp, _i, _u, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32, ctypes.c_size_t
_P = ctypes.POINTER

# name -> (restype, argtypes); every symbol include/drt.h declares
SIGNATURES = {
    "drt_scene_load_gltf": (_i, [ctypes.c_char_p, _P(_vp)]),
    "drt_scene_load_gltf_ex": (_i, [ctypes.c_char_p, ctypes.c_uint32, _P(_vp)]),
    "drt_scene_from_triangles": (_i, [_vp, _vp, _vp, _sz, _vp, _sz, _P(_vp)]),
    "drt_scene_add_texture": (_i, [_vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _vp]),
    "drt_decode_image": (_i, [_vp, _sz, _P(ctypes.c_int32), _P(ctypes.c_int32), _P(ctypes.c_int32), _vp, _sz]),
    "drt_scene_build_bvh": (_i, [_vp, _P(drt_bvh_params), _i]),
    "drt_scene_get_stats": (_i, [_vp, _P(drt_scene_stats)]),
    "drt_scene_get_triangles": (_i, [_vp, _vp, _sz]),
    "drt_scene_get_materials": (_i, [_vp, _vp, _sz]),
    "drt_scene_get_texture": (_i, [_vp, ctypes.c_int32, _P(ctypes.c_int32), _P(ctypes.c_int32), _P(ctypes.c_int32), _vp, _sz]),
    "drt_scene_get_wide_nodes": (_i, [_vp, _vp, _sz]),
    "drt_scene_get_leaf_triangles": (_i, [_vp, _vp, _sz]),
    "drt_scene_destroy": (None, [_vp]),
    "drt_renderer_create": (_i, [_i, _P(_vp)]),
    "drt_renderer_destroy": (None, [_vp]),
    "drt_renderer_set_stream": (_i, [_vp, _vp]),
    "drt_renderer_set_flags": (_i, [_vp, _u]),
    "drt_renderer_set_tuning": (_i, [_vp, ctypes.c_int32, ctypes.c_int32]),
    "drt_renderer_set_frame_schedule": (_i, [_vp, _u, _u]),
    "drt_renderer_set_tile_schedule": (_i, [_vp, _u, _u]),
    "drt_renderer_export_frame_buffers": (_i, [_vp, _vp]),
    "drt_renderer_import_frame_buffers": (_i, [_vp, _vp]),
    "drt_renderer_attach_frame_buffers": (_i, [_vp, _vp]),
    "drt_renderer_resize": (_i, [_vp, _u, _u]),
    "drt_renderer_render": (_i, [_vp, _P(drt_camera), _vp, _P(drt_settings), _u, _P(ctypes.c_float)]),
    "drt_renderer_reset": (_i, [_vp]),
    "drt_renderer_sample_count": (_u, [_vp]),
    "drt_renderer_width": (_u, [_vp]),
    "drt_renderer_height": (_u, [_vp]),
    "drt_renderer_kernel_launches": (_i, [_vp]),
    "drt_renderer_last_variant": (_i, [_vp]),
    "drt_renderer_last_launch_ordered": (_i, [_vp]),
    "drt_renderer_set_accumulate_add": (_i, [_vp, _i]),
    "drt_renderer_readback": (_i, [_vp, _i, _vp, _sz]),
    "drt_renderer_upload_accum": (_i, [_vp, _vp, _sz, _u]),
    "drt_renderer_accum_device_ptr": (_i, [_vp, _P(_vp), _P(_sz)]),
    "drt_renderer_resolve": (_i, [_vp, _u]),
    "drt_renderer_get_counters": (_i, [_vp, _P(drt_counters), _i]),
    "drt_trace_primary": (_i, [_vp, _P(drt_camera), _vp, _u, _vp, _sz]),
    "drt_multi_create": (_i, [_vp, ctypes.c_int32, _P(_vp)]),
    "drt_multi_destroy": (None, [_vp]),
    "drt_multi_device_count": (_i, [_vp]),
    "drt_multi_set_scene": (_i, [_vp, _vp, _vp]),
    "drt_multi_set_sharding": (_i, [_vp, ctypes.c_int32]),
    "drt_multi_set_flags": (_i, [_vp, _u]),
    "drt_multi_resize": (_i, [_vp, _u, _u]),
    "drt_multi_render": (_i, [_vp, _P(drt_camera), _P(drt_settings), _u, _P(ctypes.c_float)]),
    "drt_multi_reset": (_i, [_vp]),
    "drt_multi_sample_count": (_u, [_vp]),
    "drt_multi_readback": (_i, [_vp, _i, _vp, _sz]),
    "drt_multi_get_counters": (_i, [_vp, _P(drt_counters), _i]),
    "drt_multi_nccl_reduces": (_i, [_vp]),
    "drt_trace_rays": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _vp]),
    "drt_trace_primary_noalloc": (_i, [_vp, _P(drt_camera), _vp, _u, _vp, _sz]),
    "drt_default_camera": (None, [_P(drt_camera)]),
    "drt_default_settings": (None, [_P(drt_settings)]),
    "drt_last_error": (ctypes.c_char_p, []),
    "drt_version": (ctypes.c_char_p, []),
}

_lib = None


class DrtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("drt error %d: %s" % (code, message))
        self.code = code


def lib():
    """Loads libdrt.so (building it in-tree first if the sources are newer). Raises if it cannot."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from . import build
            build.build_libdrt()
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = the library does not export the ABI
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise DrtError(rc, lib().drt_last_error().decode("utf-8", "replace"))
