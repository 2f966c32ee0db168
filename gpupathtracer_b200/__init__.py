"""gpupathtracer_b200 -- B200-native replacement for the OpenGL hot path of saadam1n/GPUPathTracer.

The product is gpupathtracer_b200/libb200pt.so (hand-written CUDA for sm_100a behind the C ABI in
include/b200pt.h). This package is only its loader plus a mirror of the reference's host classes.
"""
from .api import (Camera, Renderer, MultiRenderer, Scene, PtError, load_library, make_rng_table, assemble_triangles,
                  build_light_table, presubtract_edges, build_bvh_host, LIB_PATH, SYMBOLS, REFERENCE_MAX_DEPTH,
                  PT_MICROFACET_GGX, PT_MICROFACET_BECKMANN)
from . import scenes

__all__ = ["Camera", "Renderer", "MultiRenderer", "Scene", "PtError", "load_library", "make_rng_table", "scenes", "LIB_PATH",
           "SYMBOLS", "assemble_triangles", "build_light_table", "presubtract_edges", "REFERENCE_MAX_DEPTH",
           "PT_MICROFACET_GGX", "PT_MICROFACET_BECKMANN"]
