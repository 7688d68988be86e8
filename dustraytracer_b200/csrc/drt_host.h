// drt_host.h -- host-side scene representation shared by the loader, the uploader and the C ABI.
#pragma once
#include <stdint.h>
#include <math.h>
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
// drt_jpeg.cpp: baseline / progressive JPEG with stb_image 2.29's arithmetic (the reference's decoder, Texture.cu:21-30)
bool looks_like_jpeg(const uint8_t* data, size_t len);
bool decode_jpeg(const uint8_t* data, size_t len, HostTexture& out, std::string& err);
// An embedded image of any supported container (what the reference hands to stbi_load_from_memory): PNG or JPEG by signature.
bool decode_image(const uint8_t* data, size_t len, HostTexture& out, std::string& err);
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

// Opacity classification of alpha-tested triangles (host side, exact). The device accepts a hit on a triangle whose
// albedo texture has 4 components only if the nearest texel at the interpolated UV has alpha 255 (AnyHit.cuh:17-25;
// drt_trace.cuh alpha_accept). The texels a triangle can ever sample lie in the bounding box of its three UVs in texel
// space (barycentrics of accepted hits are a convex combination; padded by one texel for rounding and for the
// reference's row overrun at frac() == 1). If every texel of that box has alpha 255 the test can never fail
// (class OPAQUE: skip it); if none has, the triangle can never be hit (class HIDDEN). Counted with a summed-area table
// per RGBA texture. Results are bit-identical with and without the classification.
enum { ALPHA_NONE = 0, ALPHA_TEST = 1, ALPHA_OPAQUE = 2, ALPHA_HIDDEN = 3 };
inline void classify_alpha_triangles(const HostScene& h, std::vector<uint8_t>& cls)
{
	const size_t n = h.tri_count(), nm = h.materials.size();
	cls.assign(n, (uint8_t)ALPHA_NONE);
	std::vector<std::vector<uint32_t>> sat(h.textures.size()); // (H+1) x (W+1) counts of alpha < 255
	auto table = [&](int ti) -> const std::vector<uint32_t>& {
		std::vector<uint32_t>& s = sat[ti];
		if (s.empty()) {
			const HostTexture& t = h.textures[ti];
			const size_t W = (size_t)t.width, H = (size_t)t.height;
			s.assign((H + 1) * (W + 1), 0u);
			for (size_t y = 0; y < H; y++) {
				uint32_t row = 0;
				for (size_t x = 0; x < W; x++) {
					row += t.pixels[(y * W + x) * 4 + 3] < 255 ? 1u : 0u;
					s[(y + 1) * (W + 1) + x + 1] = s[y * (W + 1) + x + 1] + row;
				}
			}
		}
		return s;
	};
	for (size_t i = 0; i < n; i++) {
		const int mi = h.mat[i];
		if (mi < 0 || (size_t)mi >= nm) continue;
		const int ti = h.materials[mi].albedo_texture_index;
		if (ti < 0 || ti >= (int)h.textures.size() || h.textures[ti].comps != 4) continue;
		cls[i] = (uint8_t)ALPHA_TEST;
		const HostTexture& t = h.textures[ti];
		const long long W = t.width, H = t.height;
		if (W <= 2 || H <= 2) continue;
		const float* uv = &h.uv[6 * i];
		float umin = uv[0], umax = uv[0], vmin = uv[1], vmax = uv[1];
		bool ok = true;
		for (int k = 0; k < 3; k++) {
			const float u = uv[2 * k], v = uv[2 * k + 1];
			if (!(u == u) || !(v == v) || u > 64.f || u < -64.f || v > 64.f || v < -64.f) ok = false; // rounding of huge UVs: do not reason
			umin = u < umin ? u : umin; umax = u > umax ? u : umax;
			vmin = v < vmin ? v : vmin; vmax = v > vmax ? v : vmax;
		}
		if (!ok) continue;
		const long long x0 = (long long)floor((double)umin * W) - 1, x1 = (long long)floor((double)umax * W) + 1;
		const long long y0 = (long long)floor((double)vmin * H) - 1, y1 = (long long)floor((double)vmax * H) + 1;
		if (x1 - x0 + 1 >= W || y1 - y0 + 1 >= H) continue; // wraps around the whole texture
		const std::vector<uint32_t>& s = table(ti);
		auto rect = [&](long long ax, long long bx, long long ay, long long by) -> uint64_t { // inclusive, inside the texture
			const size_t P = (size_t)W + 1;
			return (uint64_t)s[(size_t)(by + 1) * P + (size_t)(bx + 1)] - s[(size_t)ay * P + (size_t)(bx + 1)]
				- s[(size_t)(by + 1) * P + (size_t)ax] + s[(size_t)ay * P + (size_t)ax];
		};
		auto wrap = [](long long a, long long m) { long long r = a % m; return r < 0 ? r + m : r; };
		long long xs[2][2], ys[2][2];
		int nx = 1, ny = 1;
		{ const long long a = wrap(x0, W), b = wrap(x1, W); if (a <= b) { xs[0][0] = a; xs[0][1] = b; } else { xs[0][0] = a; xs[0][1] = W - 1; xs[1][0] = 0; xs[1][1] = b; nx = 2; } }
		{ const long long a = wrap(y0, H), b = wrap(y1, H); if (a <= b) { ys[0][0] = a; ys[0][1] = b; } else { ys[0][0] = a; ys[0][1] = H - 1; ys[1][0] = 0; ys[1][1] = b; ny = 2; } }
		uint64_t holes = 0;
		for (int a = 0; a < nx; a++) for (int b = 0; b < ny; b++) holes += rect(xs[a][0], xs[a][1], ys[b][0], ys[b][1]);
		const uint64_t area = (uint64_t)(x1 - x0 + 1) * (uint64_t)(y1 - y0 + 1);
		if (holes == 0) cls[i] = (uint8_t)ALPHA_OPAQUE;
		else if (holes == area) cls[i] = (uint8_t)ALPHA_HIDDEN;
	}
}

void set_error(const std::string& msg);

} // namespace drt
