// drt_host.h -- host-side scene representation shared by the loader, the uploader and the C ABI.
#pragma once
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/drt.h"

namespace drt {

struct HostTexture {
	int width = 0, height = 0, comps = 0;
	std::vector<uint8_t> pixels; // 8-bit interleaved, `comps` channels, row 0 first (Texture.cu:21-30)
};

// Triangle soup in LOAD order (= canonical primitive numbering), as the reference's loader emits it
// (Scene.cu:272-302): per triangle 3 positions, 3 vertex normals, 3 UVs, face normal, material index.
struct HostScene {
	std::vector<float> pos;      // 9 / triangle
	std::vector<float> vnormal;  // 9 / triangle (kept only so the 128-byte reference Triangle can be reproduced)
	std::vector<float> uv;       // 6 / triangle
	std::vector<float> fnormal;  // 3 / triangle
	std::vector<int32_t> mat;    // material index as loaded (may be -1: the reference would read out of bounds)
	std::vector<drt_material> materials;
	std::vector<HostTexture> textures;
	size_t meshes = 0;
	std::vector<std::string> warnings;
	size_t tri_count() const { return mat.size(); }
};

bool decode_png(const uint8_t* data, size_t len, HostTexture& out, std::string& err);
// Loads .glb / .gltf with the reference loader's semantics (Scene.cu:22-317).
// flags = drt_load_flags: DRT_LOAD_ROBUST switches to the spec-conforming ingest (drt_gltf_robust.cpp).
int load_gltf(const char* path, HostScene& out, std::string& err, unsigned flags = 0);
// face_normal = normalize(+-cross(p1-p0, p2-p0)) oriented to the mean vertex normal (Scene.cu:275-300).
void compute_face_normal(const float* p9, const float* n9, float* out3);

// Extended material table (2 x 16 bytes per material, + the stand-in for out-of-range indices): the fields the
// reference loads but never shades (Material.cuh:10-22). Only read by the opt-in DRT_FLAG_EXTENDED_MATERIALS path.
inline void fill_materials_ext(const std::vector<drt_material>& mats, std::vector<float>& out8)
{
	out8.assign((mats.size() + 1) * 8, 0.f);
	for (size_t i = 0; i < mats.size(); i++) {
		float* o = &out8[8 * i];
		o[0] = mats[i].emissive_factor[0]; o[1] = mats[i].emissive_factor[1]; o[2] = mats[i].emissive_factor[2];
		const uint32_t kind = mats[i].transmission ? 2u : (mats[i].metallic ? 1u : 0u);
		memcpy(o + 3, &kind, 4);
		o[4] = mats[i].roughness; o[5] = mats[i].refractive_index;
	}
	out8[8 * mats.size() + 5] = 1.45f;
}

void set_error(const std::string& msg);

} // namespace drt
