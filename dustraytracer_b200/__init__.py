"""dustraytracer_b200 -- B200-native path-tracing hot path behind the DustRayTracer library surface.

    from dustraytracer_b200 import Scene, BVHBuilder, Renderer, Camera, RendererSettings

The package is a thin host-side mirror (api.py) over libdrt.so (csrc/, C ABI in include/drt.h).
"""
from .api import (DRT_ACCUM_RGB32F, DRT_FLAG_EXTENDED_MATERIALS, DRT_FLAG_HOST_ARG_ORDER, DRT_FLAG_STATS, DRT_FLAG_STRICT, DRT_RGBA8, DRT_RGBA32F,
                  BVHBuilder, Camera, DrtError, MultiRenderer, Renderer, RendererSettings, Scene, decode_image, tile_owner_mask)
from ._lib import DRT_FLAG_NO_TILE_ORDER, DRT_FLAG_WATERTIGHT, drt_material  # Material POD (Material.cuh:4-23), for Scene.from_triangles(materials=[...])

__all__ = ["Scene", "BVHBuilder", "Renderer", "MultiRenderer", "decode_image", "Camera", "RendererSettings", "DrtError", "drt_material", "tile_owner_mask",
           "DRT_RGBA32F", "DRT_RGBA8", "DRT_ACCUM_RGB32F", "DRT_FLAG_STRICT", "DRT_FLAG_HOST_ARG_ORDER", "DRT_FLAG_STATS", "DRT_FLAG_EXTENDED_MATERIALS", "DRT_FLAG_WATERTIGHT", "DRT_FLAG_NO_TILE_ORDER"]
