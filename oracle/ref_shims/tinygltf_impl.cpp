// oracle/ref_shims/tinygltf_impl.cpp -- TEST INFRASTRUCTURE ONLY.
// Instantiates the reference's vendored tinygltf v2.8.10 (DustRayTracer/thirdparty/tinygltf/tiny_gltf.h:28).
// stb_image / stb_image_write implementations come from the reference's own TUs
// (DustRayTracer/thirdparty/stb_image/stb_image.cpp, stb_image_write.cpp), compiled by the Makefile.
#define TINYGLTF_IMPLEMENTATION
#include "tiny_gltf.h"
