// drt_types.h -- device-visible data layout of the hot path (shared by host and device code).
//
// HBM layout (everything 16-byte aligned so that every load on the path is one LDG.128):
//   nodes     : 80-byte compressed 8-wide BVH nodes, 5 x uint4 each            (WideNode)
//   nodes_f32 : the same nodes de-quantised to 256 bytes of fp32 planes (small scenes only; drt_phased.cuh)
//   tris      : 48 bytes per triangle in LEAF order, 3 x float4:
//                 [0] v0.xyz | canonical id (index in load order)
//                 [1] e1.xyz = v1-v0 | material index (bit 31: albedo texture carries alpha -> any-hit test)
//                 [2] e2.xyz = v2-v0 | 0
//               (the reference reads 3 x 12 B out of a 128-byte AoS Triangle: Triangle.cuh:7-19)
//   shade     : 48 bytes per triangle in leaf order, touched only when a path is shaded:
//                 [0] face_normal.xyz | uv0.x   [1] uv0.y uv1.x uv1.y uv2.x   [2] uv2.y 0 0 0
//   materials : float4 albedo.rgb | AlbedoTextureIndex (int bits)                (Material.cuh:10-13)
//   textures  : 16-byte descriptors + one RGBA8 texel pool (RGB sources padded with a=255)
#pragma once
#include <stdint.h>
#include "drt_math.cuh"

namespace drt {

struct WideNode {
	float px, py, pz;          // origin of the child-box quantisation grid
	uint8_t ex, ey, ez;        // per-axis grid scale = 2^(e-127) (biased exponent byte)
	uint8_t imask;             // bit i set: slot i holds an inner node
	uint32_t child_base;       // index of the first child wide node (children are contiguous, slot order)
	uint32_t tri_base;         // index of the first triangle referenced by this node's leaf slots
	uint8_t meta[8];           // 0 = empty; inner: 0x20 | (24+slot); leaf: unary count<<5 | offset from tri_base
	uint8_t qlox[8], qloy[8], qloz[8];
	uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(WideNode) == 80, "WideNode must be 80 bytes");

struct TexDesc {
	uint32_t offset; // first texel in the RGBA8 pool
	int32_t width, height;
	int32_t comps;   // channel count the decoder reported (Texture.cuh:15); 4 => alpha cut-out capable
};

#if defined(__CUDACC__)
typedef uint4 drt_u4;
typedef float4 drt_f4;
#else
struct alignas(16) drt_u4 { uint32_t x, y, z, w; };
struct alignas(16) drt_f4 { float x, y, z, w; };
#endif

struct SceneView {
	const drt_u4* nodes;
	const drt_f4* tris;
	const drt_f4* shade;
	const drt_f4* materials;
	const drt_f4* materials_ext; // 2 / material: emissive.rgb | kind (0 diffuse, 1 metal, 2 dielectric); roughness, ior, 0, 0
	const TexDesc* textures;
	const uint32_t* texels;
	int32_t num_tris, num_nodes, num_materials, num_textures;
	const drt_f4* nodes_f32; // expanded 256-byte nodes (drt_phased.cuh) or null: built for cache-resident scenes only
	const drt_f4* tris_v;    // 48 B per leaf-order slot: the three VERTICES (v0 | id, v1 | material word, v2 | 0) for the watertight test
};

// Camera terms that do not depend on the pixel, evaluated once on the host in the reference's
// operation order (Camera.cu:84-103) instead of once per pixel-sample on the device.
struct CameraParams {
	V3 position;
	V3 forward_focus;   // normalize(m_Forward_dir) * focus_dist
	V3 plane_h, plane_v; // world_image_plane_{horizontal,vertical}_vector
	V3 defocus_u, defocus_v;
	float defocus_angle;
	float exposure;
};

struct ShadeParams {
	V3 sunpos;       // RayGen.cuh:68-71
	V3 suncol;       // RayGen.cuh:72
	V3 sky_color;
	float sky_intensity;
	float white_scale; // 1 / uncharted2_tonemap_partial(11.2) (RayGen.cuh:39-40)
	int32_t bounce_limit;
	int32_t enable_sunlight, tone_mapping, gamma_correction;
	int32_t render_mode, debug_mode;
};

struct Counters {
	unsigned long long closest_rays;
	unsigned long long shadow_rays;
	unsigned long long node_visits;   // wide nodes fetched (stats build only)
	unsigned long long tri_tests;     // ray/triangle tests (stats build only)
	unsigned long long tie_events;    // equal-t candidates seen by closest-hit (logged for parity)
	unsigned long long alpha_tests;   // texel alpha fetches in any-hit
	unsigned long long max_stack;     // deepest traversal stack seen
	unsigned long long texel_clamps;  // texel index == width/height clamped (Texture.cu:35-36 overrun in the ref)
};

struct RenderParams {
	SceneView scene;
	CameraParams cam;
	ShadeParams shade;
	int32_t width, height;
	uint32_t frame_first;   // frame index (= RNG salt, RayGen.cuh:74-75) of the first sample of this launch
	uint32_t frame_stride;  // sample-range sharding: this renderer owns frames first, first+stride, ...
	uint32_t n_samples;     // samples rendered by this launch
	float inv_total;        // 1 / samples accumulated after this launch (display = acc * inv_total... see kernel)
	float total_samples;    // display divides by this (RenderKernel.cu:30)
	drt_f4* accum;          // W*H rgb sums (a unused)
	drt_f4* display;        // W*H acc/frameidx, a=1
	Counters* counters;
	unsigned* work_queue;   // 2 words, zero between launches: next tile, warps that have left (pixel-queue kernel)
	uint32_t tile_first, tile_stride; // image-tile sharding: this launch owns the 8x4 tiles first, first+stride, ... (row-major tile index)
	void* pool_scratch;     // sample-pool kernel (variants 22 / 23): pool colours, float4 x 32 x R per owned tile (may be null)
	// cost-ordered launch (fixed-tile kernels): block b renders the owned tiles 4 * block_order[b] + warp instead of 4 * b + warp, most expensive
	// groups first, so that the launch drains on its cheapest blocks; block_cost[g] receives the rays of group g (either may be null)
	const uint32_t* block_order;
	uint32_t* block_cost;
	// fused render + reduce (sample-range sharding without a collective): the pixel sums of this launch start at zero and are ADDED to `accum` with
	// fp32 atomics (red.global.add.f32; `accum` may be another GPU's buffer, mapped over NVLink); `display` is not written (the owner resolves)
	uint32_t accum_add;
	void* rg_overflow;      // regrouping kernel (variant 16): traversal-stack entries beyond the shared-memory ones, DRT_PSTACK x 8 bytes per resident context (may be null)
};

struct Hit {
	int32_t prim;  // leaf-order slot (device) / canonical id (API)
	float t, u, v;
};

} // namespace drt
