// drt_api.cu -- implementation of the C ABI declared in include/drt.h.
//
// Scene side : replaces Scene::loadGLTFmodel (Scene.cu:181-317) + BVHBuilder::build* (BVHBuilder.cu:11-145):
//              host triangle soup -> HBM upload -> GPU binned-SAH build -> compressed wide nodes.
// Render side: replaces Renderer::{ResizeBuffer,Render,resetAccumulationBuffer} (Renderer.cu:29-136) and
//              InvokeRenderKernel (RenderKernel.cu:37-58); the GL texture / surface is a device float4 buffer.
// No CPU fallback: every device operation reports DRT_ERR_CUDA when there is no usable GPU.
#include <cuda_runtime.h>

#include <chrono>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "drt_host.h"
#include "drt_kernels.h"
#include "drt_params.h"
#include "drt_phased.cuh"

// Expanded 256-byte nodes are kept for scenes up to this many wide nodes (4 GB of the 180 GB). Measured on the soups
// (profiles/r01c_soup_nodes.jsonl): 1.8 M wide nodes (464 MB expanded) still run 7 % faster than the 80-byte nodes,
// 174 k nodes 19 % faster: the render kernel is issue-bound there too (DRAM 62 GB/s), so fewer instructions per child
// test win over cache footprint.
#define DRT_F32_NODE_LIMIT (1 << 24)

// DRT_F32_NODE_LIMIT in the environment overrides the compiled-in limit (measurement knob: 256-byte nodes trade
// cache footprint for ~3x fewer instructions per child test).
static long long f32_node_limit()
{
	const char* e = getenv("DRT_F32_NODE_LIMIT");
	return e && *e ? atoll(e) : (long long)DRT_F32_NODE_LIMIT;
}

namespace drt {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }

static int fail(int code, const std::string& msg)
{
	set_error(msg);
	return code;
}

#define DRT_CUDA(expr)                                                                                      \
	do {                                                                                                    \
		cudaError_t e__ = (expr);                                                                           \
		if (e__ != cudaSuccess) {                                                                           \
			cudaGetLastError();                                                                             \
			return fail(DRT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));                \
		}                                                                                                   \
	} while (0)

template <class T> struct DevBuf {
	T* p = nullptr;
	size_t n = 0;
	cudaError_t alloc(size_t count)
	{
		release();
		n = count;
		if (!count) return cudaSuccess;
		return cudaMalloc((void**)&p, count * sizeof(T));
	}
	void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
	size_t bytes() const { return n * sizeof(T); }
	~DevBuf() { release(); }
};

struct CudaExec : BuildExec {
	cudaStream_t stream = nullptr;
	cudaError_t last = cudaSuccess;
	int launches = 0;
	int check(cudaError_t e) { if (e != cudaSuccess) { last = e; return -1; } return 0; }
	int launch(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n) override
	{
		launches++;
		return check(launch_build_pass(pass, A, idx_final, a, b, n, stream));
	}
	int read_state(const BuildArrays& A, BuildState& out) override
	{
		if (check(cudaMemcpyAsync(&out, A.state, sizeof(BuildState), cudaMemcpyDeviceToHost, stream))) return -1;
		return check(cudaStreamSynchronize(stream));
	}
	int write_state(const BuildArrays& A, const BuildState& in) override
	{
		return check(cudaMemcpyAsync(A.state, &in, sizeof(BuildState), cudaMemcpyHostToDevice, stream));
	}
	int write_root(const BuildArrays& A, const BuildNode& root, const BuildSplit& split) override
	{
		const int zero = 0;
		if (check(cudaMemcpyAsync(A.nodes, &root, sizeof(root), cudaMemcpyHostToDevice, stream))) return -1;
		if (check(cudaMemcpyAsync(A.splits, &split, sizeof(split), cudaMemcpyHostToDevice, stream))) return -1;
		if (check(cudaMemcpyAsync(A.wide_src, &zero, sizeof(int), cudaMemcpyHostToDevice, stream))) return -1;
		return check(cudaStreamSynchronize(stream)); // sources are stack variables
	}
};

} // namespace drt

using namespace drt;

struct drt_scene {
	HostScene host;
	int device = -1;
	bool built = false;
	DevBuf<WideNode> nodes;
	DevBuf<drt_f4> nodes_f32;
	DevBuf<drt_f4> tris, tris_v, shade, materials, materials_ext;
	DevBuf<TexDesc> texdesc;
	DevBuf<uint32_t> texels;
	SceneView view;
	drt_scene_stats stats;
	uint64_t build_serial = 0; // changes with every drt_scene_build_bvh (a renderer's cost order is tied to the tree it was measured on)
	drt_scene() { memset(&view, 0, sizeof(view)); memset(&stats, 0, sizeof(stats)); }
};

struct drt_renderer {
	int device = 0;
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	uint32_t width = 0, height = 0;
	DevBuf<drt_f4> accum, display;
	// frame buffers of ANOTHER renderer this one renders into (image-tile sharding without a collective): set by
	// drt_renderer_attach_frame_buffers (same process) or drt_renderer_import_frame_buffers (CUDA IPC, another process)
	drt_f4* ext_accum = nullptr;
	drt_f4* ext_display = nullptr;
	void* ipc_accum = nullptr;   // mappings to close
	void* ipc_display = nullptr;
	DevBuf<Counters> counters;
	DevBuf<unsigned> work_queue; // pixel-queue kernel: next tile, warps that have left (the kernel leaves both at zero)
	DevBuf<uint8_t> pool_scratch; // sample-pool kernel (variants 22 / 23): pool colours of a launch, grown on demand
	// cost-ordered launches: the first batched launch of a view records the resident time of every block's tiles (block_cost); the next
	// launches of the SAME view (scene build, camera, settings, size, tile schedule) run the blocks by descending cost (block_order),
	// so that the launch drains on its cheapest blocks instead of on whatever sits at the end of the image
	struct OrderKey {
		const drt_scene* scene; uint64_t build_serial; drt_camera cam; drt_settings settings;
		uint32_t width, height, tile_first, tile_stride, blocks, block_threads, flags, pad;
	};
	DevBuf<uint32_t> block_cost[2], block_order; // two cost arrays: one complete (it feeds the sort), one being recorded
	OrderKey order_key;
	bool order_key_valid = false;
	int cost_ready = -1;       // index of the complete cost array (-1: none)
	int cost_quality = 0;      // what it was measured on: 1 = the one-sample probe launch, 2 = a whole batched launch
	int order_quality = 0;     // quality of the costs block_order was sorted from (0: no order)
	bool last_ordered = false; // the last launch ran in cost order
	bool accum_add = false;    // drt_renderer_set_accumulate_add: launches ADD their pixel sums into the frame buffer with atomics
	DevBuf<uint8_t> rg_overflow; // regrouping kernel (variant 16): traversal-stack overflow area, allocated at its first launch
	DevBuf<Hit> hits;
	DevBuf<uint8_t> staging;
	uint32_t frame_index = 1;   // the reference's m_FrameIndex (Renderer.hpp:41): next frame to render
	uint32_t samples_done = 0;  // samples accumulated by THIS renderer
	uint32_t first_frame = 1, stride = 1;
	uint32_t tile_first = 0, tile_stride = 1; // image-tile sharding (drt_renderer_set_tile_schedule)
	uint32_t flags = 0;
	int variant = 4, block_threads = 0;
	bool auto_variant = true; // no explicit drt_renderer_set_tuning: the launch auto-tuner may pick the kernel
	int sched_bias[3] = { 0, 0, 0 };
	// Launches of at most queue_max_spp samples (the reference's interactive pattern is ONE per launch) may run the
	// persistent pixel-queue kernel (variant 9) instead of the fixed-pixel kernel (variant 4). Both give the same bits, and
	// which is faster depends on the view (room.glb 1.38x, sky-dominated dense_monkey.glb 0.82x), so the renderer MEASURES:
	// for a new (scene, size, samples, settings) it times each candidate twice with the launch's own events, keeps the
	// fastest and probes again every DRT_AUTOTUNE_PERIOD launches.
	int queue_max_spp = DRT_QUEUE_MAX_SPP_DEFAULT;
	struct AutoTune {
		const void* scene = nullptr;
		uint32_t w = 0, h = 0, n = 0, flags = 0;
		int32_t bounce = -1, sun = -1;
		int step = 0;                          // < passes * candidates: probing candidate step % candidates
		float best_ms[3] = { 0.f, 0.f, 0.f };
		int choice = 0;
		int settled_launches = 0;
		int period = 256;                      // launches between probe rounds; doubles (up to 4096) while the choice stands
		int passes = 2;                        // timings per candidate in a round: 2 for a new key (the first launch of a kernel
		                                       // in a process pays its load), 1 afterwards
	} tune;
	int sm_count = 0;
	int kernel_launches = 0;
	int last_variant = -1; // kernel variant of the last render launch (after automatic selection)
};

static void fill_stats(drt_scene* s)
{
	s->stats.triangles = s->host.tri_count();
	s->stats.materials = s->host.materials.size();
	s->stats.textures = s->host.textures.size();
	s->stats.meshes = s->host.meshes;
	std::vector<uint8_t> cls; // host-only: what the build will do with the alpha-tested triangles
	classify_alpha_triangles(s->host, cls);
	s->stats.alpha_tris = s->stats.alpha_opaque = s->stats.alpha_hidden = 0;
	for (uint8_t c : cls) {
		if (c != ALPHA_NONE) s->stats.alpha_tris++;
		if (c == ALPHA_OPAQUE) s->stats.alpha_opaque++;
		if (c == ALPHA_HIDDEN) s->stats.alpha_hidden++;
	}
}

extern "C" {

const char* drt_last_error(void) { return g_last_error.c_str(); }
const char* drt_version(void) { return "dustraytracer_b200 0.1 (sm_100a)"; }
void drt_default_camera(drt_camera* out) { if (out) default_camera(out); }
void drt_default_settings(drt_settings* out) { if (out) default_settings(out); }

int drt_scene_load_gltf(const char* path, drt_scene** out) { return drt_scene_load_gltf_ex(path, DRT_LOAD_REFERENCE_COMPAT, out); }

int drt_scene_load_gltf_ex(const char* path, uint32_t load_flags, drt_scene** out)
{
	if (!path || !out) return fail(DRT_ERR_INVALID, "drt_scene_load_gltf: null argument");
	if (load_flags & ~(uint32_t)DRT_LOAD_ROBUST) return fail(DRT_ERR_INVALID, "drt_scene_load_gltf_ex: unknown load flag");
	*out = nullptr;
	auto t0 = std::chrono::steady_clock::now();
	drt_scene* s = new drt_scene();
	std::string err;
	int rc = load_gltf(path, s->host, err, load_flags);
	if (rc != DRT_OK) { delete s; return fail(rc, err); }
	fill_stats(s);
	s->stats.load_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	*out = s;
	return DRT_OK;
}

int drt_scene_from_triangles(const float* pos, const float* uv, const int32_t* material_index, size_t n_tris,
	const drt_material* materials, size_t n_materials, drt_scene** out)
{
	if (!out || (n_tris && !pos) || (n_materials && !materials)) return fail(DRT_ERR_INVALID, "drt_scene_from_triangles: null argument");
	if (n_tris > 0x7fffffffu / 3) return fail(DRT_ERR_INVALID, "drt_scene_from_triangles: too many triangles");
	drt_scene* s = new drt_scene();
	HostScene& h = s->host;
	h.pos.assign(pos, pos + 9 * n_tris);
	if (uv) h.uv.assign(uv, uv + 6 * n_tris); else h.uv.assign(6 * n_tris, 0.f);
	h.vnormal.assign(9 * n_tris, 0.f);
	h.fnormal.resize(3 * n_tris);
	h.mat.resize(n_tris);
	const float zero9[9] = { 0, 0, 0, 0, 0, 0, 0, 0, 0 };
	for (size_t i = 0; i < n_tris; i++) {
		compute_face_normal(pos + 9 * i, zero9, h.fnormal.data() + 3 * i);
		h.mat[i] = material_index ? material_index[i] : 0;
	}
	h.materials.assign(materials, materials + n_materials);
	h.meshes = 1;
	fill_stats(s);
	*out = s;
	return DRT_OK;
}

int drt_scene_add_texture(drt_scene* scene, int32_t width, int32_t height, int32_t comps, const uint8_t* pixels)
{
	if (!scene || !pixels || width <= 0 || height <= 0 || comps < 1 || comps > 4) return fail(DRT_ERR_INVALID, "drt_scene_add_texture: bad argument");
	HostTexture t;
	t.width = width; t.height = height; t.comps = comps;
	t.pixels.assign(pixels, pixels + (size_t)width * height * comps);
	scene->host.textures.push_back(std::move(t));
	scene->built = false;
	fill_stats(scene);
	return DRT_OK;
}

void drt_scene_destroy(drt_scene* scene)
{
	if (!scene) return;
	if (scene->device >= 0) cudaSetDevice(scene->device);
	delete scene;
}

int drt_scene_get_stats(const drt_scene* scene, drt_scene_stats* out)
{
	if (!scene || !out) return fail(DRT_ERR_INVALID, "drt_scene_get_stats: null argument");
	*out = scene->stats;
	return DRT_OK;
}

int drt_scene_get_triangles(const drt_scene* scene, void* out, size_t bytes)
{
	if (!scene || !out) return fail(DRT_ERR_INVALID, "drt_scene_get_triangles: null argument");
	const HostScene& s = scene->host;
	if (bytes < s.tri_count() * 128) return fail(DRT_ERR_INVALID, "drt_scene_get_triangles: buffer too small");
	float* o = (float*)out;
	for (size_t i = 0; i < s.tri_count(); i++, o += 32) {
		const float* p = &s.pos[9 * i]; const float* n = &s.vnormal[9 * i]; const float* t = &s.uv[6 * i];
		// centroid = (v0 + v1 + v2) / 3 (Triangle.cuh:11); 4 bytes of padding follow (float2 alignment)
		o[0] = (p[0] + p[3] + p[6]) / 3; o[1] = (p[1] + p[4] + p[7]) / 3; o[2] = (p[2] + p[5] + p[8]) / 3; o[3] = 0.f;
		for (int k = 0; k < 3; k++) {
			float* v = o + 4 + 8 * k;
			v[0] = p[3 * k]; v[1] = p[3 * k + 1]; v[2] = p[3 * k + 2];
			v[3] = n[3 * k]; v[4] = n[3 * k + 1]; v[5] = n[3 * k + 2];
			v[6] = t[2 * k]; v[7] = t[2 * k + 1];
		}
		o[28] = s.fnormal[3 * i]; o[29] = s.fnormal[3 * i + 1]; o[30] = s.fnormal[3 * i + 2];
		memcpy(o + 31, &s.mat[i], 4);
	}
	return DRT_OK;
}

int drt_scene_get_materials(const drt_scene* scene, drt_material* out, size_t count)
{
	if (!scene || (!out && count)) return fail(DRT_ERR_INVALID, "drt_scene_get_materials: null argument");
	if (count < scene->host.materials.size()) return fail(DRT_ERR_INVALID, "drt_scene_get_materials: buffer too small");
	memcpy(out, scene->host.materials.data(), scene->host.materials.size() * sizeof(drt_material));
	return DRT_OK;
}

int drt_scene_get_texture(const drt_scene* scene, int32_t index, int32_t* width, int32_t* height, int32_t* comps,
	uint8_t* pixels_or_null, size_t bytes)
{
	if (!scene || index < 0 || (size_t)index >= scene->host.textures.size()) return fail(DRT_ERR_INVALID, "drt_scene_get_texture: bad index");
	const HostTexture& t = scene->host.textures[index];
	if (width) *width = t.width;
	if (height) *height = t.height;
	if (comps) *comps = t.comps;
	if (pixels_or_null) {
		if (bytes < t.pixels.size()) return fail(DRT_ERR_INVALID, "drt_scene_get_texture: buffer too small");
		memcpy(pixels_or_null, t.pixels.data(), t.pixels.size());
	}
	return DRT_OK;
}

int drt_decode_image(const uint8_t* data, size_t len, int32_t* width, int32_t* height, int32_t* comps, uint8_t* pixels_or_null, size_t bytes)
{
	if (!data || !len) return fail(DRT_ERR_INVALID, "drt_decode_image: null input");
	HostTexture t;
	std::string err;
	if (!decode_image(data, len, t, err)) return fail(DRT_ERR_FORMAT, "drt_decode_image: " + err);
	if (width) *width = t.width;
	if (height) *height = t.height;
	if (comps) *comps = t.comps;
	if (pixels_or_null) {
		if (bytes < t.pixels.size()) return fail(DRT_ERR_INVALID, "drt_decode_image: buffer too small");
		memcpy(pixels_or_null, t.pixels.data(), t.pixels.size());
	}
	return DRT_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// scene upload + GPU build
// ---------------------------------------------------------------------------------------------
extern "C" int drt_scene_build_bvh(drt_scene* scene, const drt_bvh_params* params_or_null, int device)
{
	const bool classify_alpha = !(params_or_null && (params_or_null->reserved[0] & 1));
	const bool single_launch_build = !(params_or_null && (params_or_null->reserved[0] & 2)); // bit 1: force the level-synchronous build
	if (!scene) return fail(DRT_ERR_INVALID, "drt_scene_build_bvh: null scene");
	int ndev = 0;
	DRT_CUDA(cudaGetDeviceCount(&ndev));
	if (device < 0 || device >= ndev) return fail(DRT_ERR_CUDA, "drt_scene_build_bvh: no such CUDA device (this library has no CPU path)");
	DRT_CUDA(cudaSetDevice(device));
	scene->device = device;
	scene->built = false;
	const HostScene& h = scene->host;
	const size_t n = h.tri_count();
	if (n > 0x7fffffffu / 3) return fail(DRT_ERR_INVALID, "too many triangles");

	cudaStream_t stream;
	DRT_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
	struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{ stream };

	// ---- tables: textures (RGBA8 pool), materials, per-triangle material word ----
	std::vector<TexDesc> texdesc;
	std::vector<uint32_t> texels;
	for (const HostTexture& t : h.textures) {
		TexDesc d;
		d.offset = (uint32_t)texels.size(); d.width = t.width; d.height = t.height; d.comps = t.comps;
		const size_t np = (size_t)t.width * t.height;
		texels.reserve(texels.size() + np);
		for (size_t p = 0; p < np; p++) {
			const uint8_t* px = t.pixels.data() + p * t.comps;
			uint32_t r = px[0], g = t.comps >= 3 ? px[1] : px[0], b = t.comps >= 3 ? px[2] : px[0];
			uint32_t a = t.comps == 4 ? px[3] : (t.comps == 2 ? px[1] : 255u);
			texels.push_back(r | (g << 8) | (b << 16) | (a << 24));
		}
		texdesc.push_back(d);
	}
	const size_t nm = h.materials.size();
	std::vector<drt_f4> mats(nm + 1);
	for (size_t i = 0; i <= nm; i++) {
		drt_f4 m;
		if (i < nm) {
			int tex = h.materials[i].albedo_texture_index;
			if (tex >= (int)h.textures.size()) tex = -1; // the reference would index past m_Textures
			m.x = h.materials[i].albedo[0]; m.y = h.materials[i].albedo[1]; m.z = h.materials[i].albedo[2];
			m.w = as_float((uint32_t)tex);
		}
		else { m.x = m.y = m.z = 1.f; m.w = as_float((uint32_t)-1); } // stand-in for material index -1 / out of range
		mats[i] = m;
	}
	std::vector<int32_t> mat_word(n);
	for (size_t i = 0; i < n; i++) {
		int mi = h.mat[i];
		if (mi < 0 || (size_t)mi >= nm) mi = (int)nm;
		uint32_t flag = 0;
		if ((size_t)mi < nm) {
			int tex = h.materials[mi].albedo_texture_index;
			if (tex >= 0 && tex < (int)h.textures.size() && h.textures[tex].comps == 4) flag = 0x80000000u; // AnyHit.cuh:17
		}
		mat_word[i] = (int32_t)((uint32_t)mi | flag);
	}
	// exact opacity classification (drt_host.h): triangles that can never fail the alpha test lose the flag, triangles
	// that can never pass it are collapsed to a point for the build (a degenerate triangle is rejected by Moller-Trumbore)
	std::vector<float> pos_build;
	if (classify_alpha) {
		std::vector<uint8_t> cls;
		classify_alpha_triangles(h, cls);
		for (size_t i = 0; i < n; i++) {
			if (cls[i] == ALPHA_OPAQUE) mat_word[i] &= 0x7fffffff;
			else if (cls[i] == ALPHA_HIDDEN) {
				if (pos_build.empty()) pos_build = h.pos;
				for (int k = 1; k < 3; k++) for (int c = 0; c < 3; c++) pos_build[9 * i + 3 * k + c] = pos_build[9 * i + c];
			}
		}
	}
	DRT_CUDA(scene->texdesc.alloc(texdesc.size() ? texdesc.size() : 1));
	DRT_CUDA(scene->texels.alloc(texels.size() ? texels.size() : 1));
	DRT_CUDA(scene->materials.alloc(mats.size()));
	if (!texdesc.empty()) DRT_CUDA(cudaMemcpyAsync(scene->texdesc.p, texdesc.data(), texdesc.size() * sizeof(TexDesc), cudaMemcpyHostToDevice, stream));
	if (!texels.empty()) DRT_CUDA(cudaMemcpyAsync(scene->texels.p, texels.data(), texels.size() * 4, cudaMemcpyHostToDevice, stream));
	DRT_CUDA(cudaMemcpyAsync(scene->materials.p, mats.data(), mats.size() * sizeof(drt_f4), cudaMemcpyHostToDevice, stream));
	std::vector<float> mats_ext;
	fill_materials_ext(h.materials, mats_ext);
	DRT_CUDA(scene->materials_ext.alloc(mats_ext.size() / 4));
	DRT_CUDA(cudaMemcpyAsync(scene->materials_ext.p, mats_ext.data(), mats_ext.size() * sizeof(float), cudaMemcpyHostToDevice, stream));

	// ---- build inputs + scratch ----
	DevBuf<float> d_pos, d_uv, d_fn;
	DevBuf<int32_t> d_mat;
	DevBuf<drt_f4> tri_lo, tri_hi;
	DevBuf<int> idx0, idx1, pn0, pn1, bins, wide_src;
	DevBuf<BuildNode> bnodes;
	DevBuf<BuildSplit> bsplits;
	DevBuf<BuildState> bstate;
	DevBuf<WideNode> wide_tmp;
	BuildResult res;
	double build_ms = 0.0;
	int build_launches = 0;
	if (n > 0) {
		DRT_CUDA(d_pos.alloc(9 * n)); DRT_CUDA(d_uv.alloc(6 * n)); DRT_CUDA(d_fn.alloc(3 * n)); DRT_CUDA(d_mat.alloc(n));
		DRT_CUDA(tri_lo.alloc(n)); DRT_CUDA(tri_hi.alloc(n));
		DRT_CUDA(idx0.alloc(n)); DRT_CUDA(idx1.alloc(n)); DRT_CUDA(pn0.alloc(n)); DRT_CUDA(pn1.alloc(n));
		DRT_CUDA(bnodes.alloc(2 * n + 2)); DRT_CUDA(bsplits.alloc(2 * n + 2)); DRT_CUDA(bstate.alloc(1));
		DRT_CUDA(bins.alloc(build_bin_slots_capacity((int)n) * 3 * DRT_BINS * DRT_BIN_WORDS));
		DRT_CUDA(wide_src.alloc(n + 1)); DRT_CUDA(wide_tmp.alloc(n + 1));
		DRT_CUDA(scene->tris.alloc(3 * n)); DRT_CUDA(scene->shade.alloc(3 * n));
		DRT_CUDA(cudaMemcpyAsync(d_pos.p, pos_build.empty() ? h.pos.data() : pos_build.data(), 9 * n * 4, cudaMemcpyHostToDevice, stream));
		DRT_CUDA(cudaMemcpyAsync(d_uv.p, h.uv.data(), 6 * n * 4, cudaMemcpyHostToDevice, stream));
		DRT_CUDA(cudaMemcpyAsync(d_fn.p, h.fnormal.data(), 3 * n * 4, cudaMemcpyHostToDevice, stream));
		DRT_CUDA(cudaMemcpyAsync(d_mat.p, mat_word.data(), n * 4, cudaMemcpyHostToDevice, stream));
		DRT_CUDA(cudaStreamSynchronize(stream));

		BuildArrays A;
		memset(&A, 0, sizeof(A));
		A.pos = d_pos.p; A.uv = d_uv.p; A.fnormal = d_fn.p; A.mat = d_mat.p; A.n = (int)n;
		A.tri_lo = tri_lo.p; A.tri_hi = tri_hi.p;
		A.idx_in = idx0.p; A.idx_out = idx1.p; A.pnode_in = pn0.p; A.pnode_out = pn1.p;
		A.nodes = bnodes.p; A.splits = bsplits.p; A.bins = bins.p; A.state = bstate.p;
		A.wide_src = wide_src.p; A.wide = wide_tmp.p; A.out_tris = scene->tris.p; A.out_shade = scene->shade.p;
		CudaExec ex;
		ex.stream = stream;
		DevBuf<int> d_res; // result words of the single-launch builder (allocated outside the timed region)
		DRT_CUDA(d_res.alloc(DRT_BUILD_SMALL_RESULT_INTS));
		cudaEvent_t e0, e1;
		DRT_CUDA(cudaEventCreate(&e0)); DRT_CUDA(cudaEventCreate(&e1));
		DRT_CUDA(cudaEventRecord(e0, stream));
		int rc = 0;
		if (single_launch_build && n <= (size_t)DRT_BUILD_SMALL_MAX_TRIS) {
			// small scene: the whole level loop in one launch of one block (k_build_small), one read-back at the end
			int h_res[DRT_BUILD_SMALL_RESULT_INTS] = { 0 };
			rc = ex.check(launch_build_small(A, d_res.p, stream)) ||
				ex.check(cudaMemcpyAsync(h_res, d_res.p, sizeof(h_res), cudaMemcpyDeviceToHost, stream)) || ex.check(cudaStreamSynchronize(stream));
			ex.launches = 1;
			res.bvh2_nodes = h_res[0]; res.wide_nodes = h_res[1]; res.levels = h_res[2]; res.wide_levels = h_res[3]; res.tris_emitted = h_res[4];
#ifdef DRT_BUILD_SMALL_TIMING
			fprintf(stderr, "k_build_small clocks: warp split max sah %d partition %d | init %d", h_res[6], h_res[7], h_res[8]);
			for (int l = 0; l < h_res[2] && l < 16; l++) fprintf(stderr, " | L%d bin %d split %d part %d", l, h_res[9 + 3 * l], h_res[10 + 3 * l], h_res[11 + 3 * l]);
			for (int l = 0; l < h_res[3] && l < 6; l++) fprintf(stderr, " | W%d %d", l, h_res[57 + l]);
			fprintf(stderr, "\n");
#endif
		}
		else rc = run_build(ex, A, res);
		cudaEventRecord(e1, stream);
		cudaError_t se = cudaStreamSynchronize(stream);
		float ms = 0.f;
		cudaEventElapsedTime(&ms, e0, e1);
		cudaEventDestroy(e0); cudaEventDestroy(e1);
		if (rc != 0 || se != cudaSuccess) {
			cudaGetLastError();
			return fail(DRT_ERR_CUDA, std::string("BVH build failed: ") + cudaGetErrorString(ex.last != cudaSuccess ? ex.last : se));
		}
		build_ms = ms;
		build_launches = ex.launches;
		if (res.tris_emitted != (int)n) return fail(DRT_ERR_CUDA, "BVH build emitted " + std::to_string(res.tris_emitted) + " of " + std::to_string(n) + " triangles");
		// The render kernels keep DRT_PSTACK traversal-stack entries per lane and push at most one entry (the rest of a node group)
		// per level of the wide tree: a tree deeper than that could lose subtrees, so it is refused here rather than rendered wrong.
		if (res.wide_levels > DRT_PSTACK)
			return fail(DRT_ERR_INVALID, "BVH has " + std::to_string(res.wide_levels) + " wide levels, more than the " + std::to_string(DRT_PSTACK) + "-entry traversal stack of the render kernels");
		// keep only the wide nodes that exist
		DRT_CUDA(scene->nodes.alloc((size_t)res.wide_nodes));
		DRT_CUDA(cudaMemcpyAsync(scene->nodes.p, wide_tmp.p, (size_t)res.wide_nodes * sizeof(WideNode), cudaMemcpyDeviceToDevice, stream));
		// cache-resident scenes also get the de-quantised 256-byte nodes (issue-bound regime, drt_phased.cuh)
		scene->nodes_f32.release();
		if ((long long)res.wide_nodes <= f32_node_limit()) {
			DRT_CUDA(scene->nodes_f32.alloc((size_t)res.wide_nodes * DRT_F32NODE_F4));
			DRT_CUDA(launch_expand_nodes(scene->nodes.p, scene->nodes_f32.p, res.wide_nodes, stream));
		}
		// vertex records for the watertight test (48 B per triangle, leaf order)
		DRT_CUDA(scene->tris_v.alloc(3 * n));
		DRT_CUDA(launch_vertex_records(scene->tris.p, d_pos.p, scene->tris_v.p, (int)n, stream));
		DRT_CUDA(cudaStreamSynchronize(stream));
	}
	else {
		scene->nodes.release(); scene->nodes_f32.release(); scene->tris.release(); scene->tris_v.release(); scene->shade.release();
		DRT_CUDA(cudaStreamSynchronize(stream));
	}

	SceneView& v = scene->view;
	v.nodes = (const drt_u4*)scene->nodes.p;
	v.tris = scene->tris.p; v.shade = scene->shade.p; v.materials = scene->materials.p; v.materials_ext = scene->materials_ext.p;
	v.textures = scene->texdesc.p; v.texels = scene->texels.p;
	v.nodes_f32 = scene->nodes_f32.p;
	v.tris_v = scene->tris_v.p;
	v.num_tris = (int32_t)n; v.num_nodes = res.wide_nodes; v.num_materials = (int32_t)mats.size(); v.num_textures = (int32_t)texdesc.size();
	scene->stats.wide_nodes = (uint64_t)res.wide_nodes;
	scene->stats.bvh2_nodes = (uint64_t)res.bvh2_nodes;
	scene->stats.bvh_levels = (uint64_t)res.levels;
	scene->stats.build_ms = build_ms;
	scene->stats.build_launches = (uint64_t)build_launches;
	scene->stats.device_bytes = scene->nodes.bytes() + scene->nodes_f32.bytes() + scene->tris.bytes() + scene->tris_v.bytes() + scene->shade.bytes() + scene->materials.bytes()
		+ scene->texdesc.bytes() + scene->texels.bytes();
	scene->built = true;
	static std::atomic<uint64_t> serial{ 0 };
	scene->build_serial = ++serial;
	return DRT_OK;
}

extern "C" int drt_scene_get_wide_nodes(const drt_scene* scene, void* out, size_t bytes)
{
	if (!scene || !out || !scene->built) return fail(DRT_ERR_INVALID, "drt_scene_get_wide_nodes: scene not built");
	if (bytes < scene->nodes.bytes()) return fail(DRT_ERR_INVALID, "drt_scene_get_wide_nodes: buffer too small");
	DRT_CUDA(cudaSetDevice(scene->device));
	if (scene->nodes.n) DRT_CUDA(cudaMemcpy(out, scene->nodes.p, scene->nodes.bytes(), cudaMemcpyDeviceToHost));
	return DRT_OK;
}

extern "C" int drt_scene_get_leaf_triangles(const drt_scene* scene, void* out, size_t bytes)
{
	if (!scene || !out || !scene->built) return fail(DRT_ERR_INVALID, "drt_scene_get_leaf_triangles: scene not built");
	if (bytes < scene->tris.bytes()) return fail(DRT_ERR_INVALID, "drt_scene_get_leaf_triangles: buffer too small");
	DRT_CUDA(cudaSetDevice(scene->device));
	if (scene->tris.n) DRT_CUDA(cudaMemcpy(out, scene->tris.p, scene->tris.bytes(), cudaMemcpyDeviceToHost));
	return DRT_OK;
}

// ---------------------------------------------------------------------------------------------
// renderer
// ---------------------------------------------------------------------------------------------
static drt_f4* fb_accum(const drt_renderer* r) { return r->ext_accum ? r->ext_accum : r->accum.p; }
static drt_f4* fb_display(const drt_renderer* r) { return r->ext_display ? r->ext_display : r->display.p; }

static void detach_frame_buffers(drt_renderer* r)
{
	if (r->ipc_accum) cudaIpcCloseMemHandle(r->ipc_accum);
	if (r->ipc_display) cudaIpcCloseMemHandle(r->ipc_display);
	r->ipc_accum = r->ipc_display = nullptr;
	r->ext_accum = r->ext_display = nullptr;
}

static int renderer_bind(drt_renderer* r)
{
	DRT_CUDA(cudaSetDevice(r->device));
	return DRT_OK;
}

extern "C" int drt_renderer_create(int device, drt_renderer** out)
{
	if (!out) return fail(DRT_ERR_INVALID, "drt_renderer_create: null argument");
	*out = nullptr;
	int ndev = 0;
	DRT_CUDA(cudaGetDeviceCount(&ndev));
	if (device < 0 || device >= ndev) return fail(DRT_ERR_CUDA, "drt_renderer_create: no such CUDA device (this library has no CPU path)");
	DRT_CUDA(cudaSetDevice(device));
	drt_renderer* r = new drt_renderer();
	r->device = device;
	cudaError_t e = cudaStreamCreateWithFlags(&r->own_stream, cudaStreamNonBlocking);
	if (e == cudaSuccess) e = cudaEventCreate(&r->ev0);
	if (e == cudaSuccess) e = cudaEventCreate(&r->ev1);
	if (e == cudaSuccess) e = r->counters.alloc(1);
	if (e == cudaSuccess) e = cudaMemset(r->counters.p, 0, sizeof(Counters));
	if (e == cudaSuccess) e = r->work_queue.alloc(2);
	if (e == cudaSuccess) e = cudaMemsetAsync(r->work_queue.p, 0, 2 * sizeof(unsigned), r->own_stream);
	if (e == cudaSuccess) e = cudaDeviceGetAttribute(&r->sm_count, cudaDevAttrMultiProcessorCount, device);
	if (e != cudaSuccess) { delete r; return fail(DRT_ERR_CUDA, std::string("drt_renderer_create: ") + cudaGetErrorString(e)); }
	r->stream = r->own_stream;
	if (const char* b = getenv("DRT_SCHED_BIAS")) sscanf(b, "%d,%d,%d", &r->sched_bias[0], &r->sched_bias[1], &r->sched_bias[2]); // measurement knob
	if (const char* b = getenv("DRT_QUEUE_MAX_SPP")) r->queue_max_spp = atoi(b);                                                   // measurement knob
	*out = r;
	return DRT_OK;
}

extern "C" void drt_renderer_destroy(drt_renderer* r)
{
	if (!r) return;
	cudaSetDevice(r->device);
	cudaStreamSynchronize(r->stream);
	detach_frame_buffers(r);
	if (r->ev0) cudaEventDestroy(r->ev0);
	if (r->ev1) cudaEventDestroy(r->ev1);
	if (r->own_stream) cudaStreamDestroy(r->own_stream);
	delete r;
}

extern "C" int drt_renderer_set_stream(drt_renderer* r, void* cuda_stream)
{
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_set_stream: null renderer");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	r->stream = cuda_stream ? (cudaStream_t)cuda_stream : r->own_stream;
	return DRT_OK;
}

extern "C" int drt_renderer_set_flags(drt_renderer* r, uint32_t flags)
{
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_set_flags: null renderer");
	r->flags = flags;
	return DRT_OK;
}

extern "C" int drt_renderer_set_tuning(drt_renderer* r, int32_t variant, int32_t block_threads)
{
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_set_tuning: null renderer");
	if (variant < -1 || variant > 23 || block_threads < 0 || block_threads > 256 || block_threads % 32) return fail(DRT_ERR_INVALID, "drt_renderer_set_tuning: bad value");
	r->auto_variant = variant < 0;
	r->variant = variant < 0 ? 4 : variant; r->block_threads = block_threads;
	return DRT_OK;
}

extern "C" int drt_renderer_set_frame_schedule(drt_renderer* r, uint32_t first_frame, uint32_t stride)
{
	if (!r || first_frame == 0 || stride == 0) return fail(DRT_ERR_INVALID, "drt_renderer_set_frame_schedule: first_frame and stride must be >= 1");
	r->first_frame = first_frame; r->stride = stride;
	return drt_renderer_reset(r);
}

extern "C" int drt_renderer_set_tile_schedule(drt_renderer* r, uint32_t first_tile, uint32_t stride)
{
	if (!r || stride == 0 || first_tile >= stride || stride > 65536) return fail(DRT_ERR_INVALID, "drt_renderer_set_tile_schedule: need first_tile < stride <= 65536");
	r->tile_first = first_tile; r->tile_stride = stride;
	r->tune = drt_renderer::AutoTune();
	return drt_renderer_reset(r);
}

// ---- frame-buffer sharing: image-tile sharding where every rank's kernel stores its tiles straight into the root's buffers ----
struct FrameHandle {
	cudaIpcMemHandle_t accum, display;
	uint32_t width, height;
	int32_t device;
	uint32_t magic;
};
static_assert(sizeof(FrameHandle) <= DRT_FRAME_HANDLE_BYTES, "frame handle must fit the ABI's byte array");
#define DRT_FRAME_MAGIC 0x44525446u

extern "C" int drt_renderer_export_frame_buffers(drt_renderer* r, void* handle_out)
{
	if (!r || !handle_out || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_export_frame_buffers: renderer has no buffer");
	if (r->ext_accum) return fail(DRT_ERR_INVALID, "drt_renderer_export_frame_buffers: this renderer renders into another renderer's buffers");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	FrameHandle h;
	memset(&h, 0, sizeof(h));
	DRT_CUDA(cudaIpcGetMemHandle(&h.accum, r->accum.p));
	DRT_CUDA(cudaIpcGetMemHandle(&h.display, r->display.p));
	h.width = r->width; h.height = r->height; h.device = r->device; h.magic = DRT_FRAME_MAGIC;
	memset(handle_out, 0, DRT_FRAME_HANDLE_BYTES);
	memcpy(handle_out, &h, sizeof(h));
	return DRT_OK;
}

extern "C" int drt_renderer_import_frame_buffers(drt_renderer* r, const void* handle)
{
	if (!r || !handle || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_import_frame_buffers: resize the renderer first");
	FrameHandle h;
	memcpy(&h, handle, sizeof(h));
	if (h.magic != DRT_FRAME_MAGIC) return fail(DRT_ERR_INVALID, "drt_renderer_import_frame_buffers: not a frame-buffer handle");
	if (h.width != r->width || h.height != r->height) return fail(DRT_ERR_INVALID, "drt_renderer_import_frame_buffers: size differs from the exporting renderer");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	detach_frame_buffers(r);
	void* a = nullptr;
	void* d = nullptr;
	// (maps the exporter's allocation into this process and enables peer access to its device: loads and stores go over NVLink)
	DRT_CUDA(cudaIpcOpenMemHandle(&a, h.accum, cudaIpcMemLazyEnablePeerAccess));
	cudaError_t e = cudaIpcOpenMemHandle(&d, h.display, cudaIpcMemLazyEnablePeerAccess);
	if (e != cudaSuccess) { cudaIpcCloseMemHandle(a); return fail(DRT_ERR_CUDA, std::string("drt_renderer_import_frame_buffers: ") + cudaGetErrorString(e)); }
	r->ipc_accum = a; r->ipc_display = d;
	r->ext_accum = (drt_f4*)a; r->ext_display = (drt_f4*)d;
	return drt_renderer_reset(r);
}

extern "C" int drt_renderer_attach_frame_buffers(drt_renderer* r, drt_renderer* owner)
{
	if (!r || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_attach_frame_buffers: resize the renderer first");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	detach_frame_buffers(r);
	if (!owner || owner == r) return drt_renderer_reset(r);
	if (owner->ext_accum) return fail(DRT_ERR_INVALID, "drt_renderer_attach_frame_buffers: the owner renders into another renderer's buffers itself");
	if (owner->width != r->width || owner->height != r->height) return fail(DRT_ERR_INVALID, "drt_renderer_attach_frame_buffers: size differs from the owner");
	if (owner->device != r->device) {
		int can = 0;
		DRT_CUDA(cudaDeviceCanAccessPeer(&can, r->device, owner->device));
		if (!can) return fail(DRT_ERR_CUDA, "drt_renderer_attach_frame_buffers: no peer access between the two devices");
		cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
		if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
		else if (e != cudaSuccess) return fail(DRT_ERR_CUDA, std::string("drt_renderer_attach_frame_buffers: ") + cudaGetErrorString(e));
	}
	r->ext_accum = owner->accum.p; r->ext_display = owner->display.p;
	return drt_renderer_reset(r);
}

extern "C" int drt_renderer_reset(drt_renderer* r)
{
	// resetAccumulationBuffer (Renderer.cu:132-136): zero the sums, frame index back to 1
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_reset: null renderer");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	// (a renderer attached to another renderer's frame buffers leaves zeroing them to their owner)
	if (r->accum.n && !r->ext_accum) DRT_CUDA(cudaMemsetAsync(r->accum.p, 0, r->accum.bytes(), r->stream));
	if (r->display.n && !r->ext_display) DRT_CUDA(cudaMemsetAsync(r->display.p, 0, r->display.bytes(), r->stream));
	r->frame_index = 1;
	r->samples_done = 0;
	return DRT_OK;
}

extern "C" int drt_renderer_resize(drt_renderer* r, uint32_t width, uint32_t height)
{
	// ResizeBuffer (Renderer.cu:29-78): no-op when unchanged, otherwise reallocate and reset
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_resize: null renderer");
	if (width == 0 || height == 0 || (uint64_t)width * height > 0x7fffffffull) return fail(DRT_ERR_INVALID, "drt_renderer_resize: bad size");
	if (width == r->width && height == r->height) return DRT_OK;
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	detach_frame_buffers(r); // shared buffers have the old size
	const size_t n = (size_t)width * height;
	DRT_CUDA(r->accum.alloc(n));
	DRT_CUDA(r->display.alloc(n));
	r->hits.release(); r->staging.release();
	r->width = width; r->height = height;
	return drt_renderer_reset(r);
}

extern "C" uint32_t drt_renderer_sample_count(const drt_renderer* r) { return r ? r->frame_index : 0; }
extern "C" uint32_t drt_renderer_width(const drt_renderer* r) { return r ? r->width : 0; }
extern "C" uint32_t drt_renderer_height(const drt_renderer* r) { return r ? r->height : 0; }
extern "C" int drt_renderer_kernel_launches(const drt_renderer* r) { return r ? r->kernel_launches : 0; }
extern "C" int drt_renderer_last_variant(const drt_renderer* r) { return r ? r->last_variant : -1; }

static int make_render_params(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, const drt_settings* settings, RenderParams& P)
{
	if (!r || !cam || !scene) return fail(DRT_ERR_INVALID, "render: null argument");
	if (!scene->built) return fail(DRT_ERR_INVALID, "render: call drt_scene_build_bvh first");
	if (scene->device != r->device) return fail(DRT_ERR_INVALID, "render: scene was built on another device");
	if (!r->width) return fail(DRT_ERR_INVALID, "render: call drt_renderer_resize first");
	memset(&P, 0, sizeof(P));
	P.scene = scene->view;
	P.cam = make_camera_params(*cam, r->width, r->height);
	drt_settings dflt;
	if (!settings) { default_settings(&dflt); settings = &dflt; }
	P.shade = make_shade_params(*settings);
	P.width = (int32_t)r->width; P.height = (int32_t)r->height;
	P.accum = fb_accum(r); P.display = fb_display(r); P.counters = r->counters.p;
	P.work_queue = r->work_queue.p;
	P.tile_first = r->tile_first; P.tile_stride = r->tile_stride;
	P.rg_overflow = r->rg_overflow.p;
	P.pool_scratch = r->pool_scratch.p;
	P.accum_add = r->accum_add ? 1u : 0u;
	return DRT_OK;
}

extern "C" int drt_renderer_last_launch_ordered(const drt_renderer* r) { return r && r->last_ordered ? 1 : 0; }

extern "C" int drt_renderer_set_accumulate_add(drt_renderer* r, int on)
{
	if (!r) return fail(DRT_ERR_INVALID, "drt_renderer_set_accumulate_add: null renderer");
	r->accum_add = on != 0;
	return DRT_OK;
}

#define DRT_AUTOTUNE_CANDIDATES 3
#define DRT_AUTOTUNE_PERIOD 256
#define DRT_AUTOTUNE_PERIOD_MAX 4096

// Candidate c of the launch auto-tuner: 0 = fixed-pixel kernel, 1 / 2 = pixel-queue kernel with an early / late refill.
static void autotune_apply(int c, uint32_t n, LaunchConfig& cfg)
{
	if (c == 0) return;
	cfg.variant = 9;
	cfg.sched_bias[1] = c == 1 ? 4 : (n <= 1 ? 16 : 10);
}

// End of a probe round: keep the fastest candidate; the interval to the next round doubles while the choice stands.
static void autotune_settle(drt_renderer::AutoTune& t)
{
	int choice = 0;
	for (int c = 1; c < DRT_AUTOTUNE_CANDIDATES; c++)
		if (t.best_ms[c] < 0.98f * t.best_ms[choice]) choice = c; // the fixed-pixel kernel keeps ties
	t.period = (choice == t.choice && t.period > 0) ? (2 * t.period > DRT_AUTOTUNE_PERIOD_MAX ? DRT_AUTOTUNE_PERIOD_MAX : 2 * t.period) : DRT_AUTOTUNE_PERIOD;
	t.choice = choice;
	t.settled_launches = 0;
	t.passes = 1; // later rounds find the kernels loaded and warm: one timing per candidate
}

// Returns the candidate this launch runs (and whether its time is a probe), or -1 when the launch is not eligible.
static int autotune_pick(drt_renderer* r, const drt_scene* scene, const RenderParams& P, bool& probing)
{
	probing = false;
	if (!r->auto_variant || r->variant != 4 || (r->block_threads != 0 && r->block_threads != 128) || r->sched_bias[1] != 0 || (int64_t)P.n_samples > (int64_t)r->queue_max_spp) return -1;
	if (P.shade.render_mode != 0 || (r->flags & (DRT_FLAG_EXTENDED_MATERIALS | DRT_FLAG_WATERTIGHT)) || P.width >= 65536 || P.height >= 65536) return -1;
	drt_renderer::AutoTune& t = r->tune;
	if (t.scene != scene || t.w != (uint32_t)P.width || t.h != (uint32_t)P.height || t.n != P.n_samples || t.flags != r->flags ||
		t.bounce != P.shade.bounce_limit || t.sun != P.shade.enable_sunlight) {
		t = drt_renderer::AutoTune();
		t.period = 0; // (first round of this key: the period starts at DRT_AUTOTUNE_PERIOD)
		t.scene = scene; t.w = (uint32_t)P.width; t.h = (uint32_t)P.height; t.n = P.n_samples; t.flags = r->flags;
		t.bounce = P.shade.bounce_limit; t.sun = P.shade.enable_sunlight;
	}
	if (t.step < t.passes * DRT_AUTOTUNE_CANDIDATES) { probing = true; return t.step % DRT_AUTOTUNE_CANDIDATES; }
	return t.choice;
}

static void autotune_record(drt_renderer* r, int candidate, bool probing, float ms)
{
	drt_renderer::AutoTune& t = r->tune;
	if (!probing) {
		if (++t.settled_launches >= t.period) t.step = 0; // next probe round
		return;
	}
	if (t.step < DRT_AUTOTUNE_CANDIDATES || ms < t.best_ms[candidate]) t.best_ms[candidate] = ms;
	if (++t.step == t.passes * DRT_AUTOTUNE_CANDIDATES) autotune_settle(t);
}

static LaunchConfig launch_config(const drt_renderer* r)
{
	LaunchConfig cfg;
	cfg.strict = (r->flags & DRT_FLAG_STRICT) != 0;
	cfg.rtl = (r->flags & DRT_FLAG_HOST_ARG_ORDER) != 0;
	cfg.stats = (r->flags & DRT_FLAG_STATS) != 0;
	cfg.ext_materials = (r->flags & DRT_FLAG_EXTENDED_MATERIALS) != 0;
	cfg.watertight = (r->flags & DRT_FLAG_WATERTIGHT) != 0;
	cfg.variant = r->variant; cfg.block_threads = r->block_threads;
	cfg.auto_variant = r->auto_variant;
	cfg.resident_sms = r->sm_count;
	for (int i = 0; i < 3; i++) cfg.sched_bias[i] = r->sched_bias[i];
	return cfg;
}

// Launch half of drt_renderer_render: enqueues the batch on the renderer's stream and returns without waiting.
// *launched = samples the launch renders (0: nothing was enqueued), candidate / probing = the auto-tuner's bookkeeping.
static int render_launch(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, const drt_settings* settings, uint32_t n_samples,
	uint32_t* launched, int* candidate_out, bool* probing_out)
{
	*launched = 0; *candidate_out = -1; *probing_out = false;
	RenderParams P;
	int rc = make_render_params(r, cam, scene, settings, P);
	if (rc) return rc;
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	drt_settings dflt;
	if (!settings) { default_settings(&dflt); settings = &dflt; }
	// Renderer.cu:82: nothing is rendered once the frame index equals max_samples
	const uint64_t next_frame = (uint64_t)r->first_frame + (uint64_t)r->samples_done * r->stride;
	const int64_t max_samples = settings->max_samples;
	uint32_t n = n_samples;
	if ((int64_t)next_frame >= max_samples) {
		if ((int64_t)next_frame == max_samples || r->stride > 1) n = 0; // the reference only tests equality
	}
	else {
		const uint64_t room = ((uint64_t)max_samples - next_frame + r->stride - 1) / r->stride;
		if (n > room) n = (uint32_t)room;
	}
	if (n == 0) return DRT_OK;
	*launched = n;
	P.frame_first = (uint32_t)next_frame;
	P.frame_stride = r->stride;
	P.n_samples = n;
	P.total_samples = (float)(r->samples_done + n); // = the frame index the reference divides by (RenderKernel.cu:30)
	LaunchConfig cfg = launch_config(r);
	const bool pool_auto = cfg.auto_variant && cfg.variant == 4 && n >= DRT_POOL_MIN_SPP && scene->view.num_tris >= DRT_POOL_MIN_TRIS &&
		scene->view.nodes_f32 != nullptr && !(r->flags & (DRT_FLAG_STATS | DRT_FLAG_EXTENDED_MATERIALS | DRT_FLAG_WATERTIGHT)) && P.shade.render_mode == 0;
	if (cfg.variant == 22 || cfg.variant == 23 || pool_auto) { // (grown on demand, outside any timed launch: the pool colours of one launch)
		const long long all_tiles = (long long)((r->width + 7) / 8) * ((r->height + 3) / 4);
		const long long tiles = all_tiles > r->tile_first ? (all_tiles - r->tile_first + r->tile_stride - 1) / r->tile_stride : 0;
		const size_t need = sample_pool_scratch_bytes(tiles, sample_pool_size(n, r->sched_bias[2]));
		if (need > r->pool_scratch.n && !(pool_auto && need > DRT_POOL_MAX_SCRATCH_BYTES)) {
			DRT_CUDA(cudaStreamSynchronize(r->stream));
			DRT_CUDA(r->pool_scratch.alloc(need));
		}
		P.pool_scratch = r->pool_scratch.p;
		cfg.pool_scratch_bytes = r->pool_scratch.n;
	}
	if (cfg.variant == 16 && !r->rg_overflow.p) { // (a one-off: entries beyond the shared-memory part of the regrouping kernel's stacks)
		DRT_CUDA(r->rg_overflow.alloc(regroup_overflow_bytes(r->sm_count)));
		P.rg_overflow = r->rg_overflow.p;
	}
	bool probing = false;
	const int candidate = r->accum_add ? -1 : autotune_pick(r, scene, P, probing);
	if (candidate >= 0) autotune_apply(candidate, n, cfg);
	// cost order (see drt_renderer::OrderKey): launches of DRT_ORDER_MIN_SPP samples or more over at least two waves of blocks
	int record_into = -1; // cost array this launch records into
	uint32_t probe = 0;   // samples of the probe launch
	{
		const uint32_t threads = cfg.block_threads > 0 ? (uint32_t)cfg.block_threads : 128u, wpb = threads / 32u;
		const long long all_tiles = (long long)((r->width + 7) / 8) * ((r->height + 3) / 4);
		const long long tiles = all_tiles > r->tile_first ? (all_tiles - r->tile_first + r->tile_stride - 1) / r->tile_stride : 0;
		const long long blocks = wpb ? (tiles + wpb - 1) / wpb : 0;
		const bool order_ok = !(r->flags & (DRT_FLAG_NO_TILE_ORDER | DRT_FLAG_EXTENDED_MATERIALS)) && candidate < 0 && n >= DRT_ORDER_MIN_SPP &&
			P.shade.render_mode == 0 && blocks >= 2ll * r->sm_count * 6 && blocks < 0x7fffffffll;
		if (!order_ok) { r->order_key_valid = false; r->cost_ready = -1; r->cost_quality = r->order_quality = 0; }
		else {
			drt_renderer::OrderKey k;
			memset(&k, 0, sizeof(k));
			k.scene = scene; k.build_serial = scene->build_serial; k.cam = *cam; if (settings) k.settings = *settings;
			k.width = r->width; k.height = r->height; k.tile_first = r->tile_first; k.tile_stride = r->tile_stride;
			k.blocks = (uint32_t)blocks; k.block_threads = threads; k.flags = r->flags;
			k.settings.max_samples = 0; // (how many samples may still come does not change what a tile costs)
			k.settings._pad0 = 0; k.settings._pad1 = 0; // (padding of the reference's struct: whatever the caller's memory held)
			if (!r->order_key_valid || memcmp(&k, &r->order_key, sizeof(k)) != 0) { // a new view
				r->order_key = k; r->order_key_valid = true;
				r->cost_ready = -1; r->cost_quality = r->order_quality = 0;
				if (r->block_order.n < (size_t)blocks) {
					DRT_CUDA(cudaStreamSynchronize(r->stream));
					DRT_CUDA(r->block_cost[0].alloc((size_t)blocks)); DRT_CUDA(r->block_cost[1].alloc((size_t)blocks)); DRT_CUDA(r->block_order.alloc((size_t)blocks));
				}
				// a long launch does not wait for the NEXT launch of the view to be ordered: its first sample is launched on its own
				// and records the costs (128 paths per block)
				static const uint32_t probe_spp = getenv("DRT_ORDER_PROBE_SPP") ? (uint32_t)atoi(getenv("DRT_ORDER_PROBE_SPP")) : 1u; // (0: never; measurement knob)
				if (probe_spp > 0 && n >= DRT_ORDER_PROBE_MIN_SPP && probe_spp < n && !r->accum_add) probe = probe_spp; // (two atomic adds per pixel would not be one)
			}
			if (r->order_quality < 2) record_into = r->cost_ready == 0 ? 1 : 0; // until the order rests on a whole launch's costs
		}
	}
	DRT_CUDA(cudaEventRecord(r->ev0, r->stream));
	if (probe) {
		// (the pixel sums see the same additions in the same order as in one launch)
		RenderParams A = P;
		A.n_samples = probe;
		A.total_samples = (float)(r->samples_done + probe);
		A.block_cost = r->block_cost[record_into].p;
		int probe_variant = -1;
		DRT_CUDA(cudaMemsetAsync(A.block_cost, 0, (size_t)r->order_key.blocks * sizeof(uint32_t), r->stream));
		DRT_CUDA(launch_render(A, cfg, r->stream, &r->kernel_launches, &probe_variant));
		if (variant_honours_block_order(probe_variant)) { r->cost_ready = record_into; r->cost_quality = 1; record_into = 1 - record_into; }
		P.frame_first += probe * r->stride;
		P.n_samples = n - probe;
	}
	if (r->cost_ready >= 0) {
		if (r->cost_quality > r->order_quality) {
			DRT_CUDA(launch_block_order(r->block_cost[r->cost_ready].p, r->block_order.p, (int)r->order_key.blocks, r->stream));
			r->kernel_launches++;
			r->order_quality = r->cost_quality;
		}
		P.block_order = r->block_order.p;
	}
	if (record_into >= 0) {
		P.block_cost = r->block_cost[record_into].p;
		DRT_CUDA(cudaMemsetAsync(P.block_cost, 0, (size_t)r->order_key.blocks * sizeof(uint32_t), r->stream));
	}
	DRT_CUDA(launch_render(P, cfg, r->stream, &r->kernel_launches, &r->last_variant));
	DRT_CUDA(cudaEventRecord(r->ev1, r->stream));
	const bool honoured = variant_honours_block_order(r->last_variant);
	r->last_ordered = P.block_order != nullptr && honoured;
	if (record_into >= 0 && honoured) { r->cost_ready = record_into; r->cost_quality = 2; }
	if (r->order_key_valid && !honoured) { r->order_key_valid = false; r->cost_ready = -1; r->cost_quality = r->order_quality = 0; }
	*candidate_out = candidate; *probing_out = probing;
	r->samples_done += n;
	r->frame_index += n;
	return DRT_OK;
}

extern "C" int drt_renderer_render(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, const drt_settings* settings,
	uint32_t n_samples, float* ms_or_null)
{
	if (ms_or_null) *ms_or_null = 0.f;
	uint32_t n = 0;
	int candidate = -1;
	bool probing = false;
	const int rc = render_launch(r, cam, scene, settings, n_samples, &n, &candidate, &probing);
	if (rc || n == 0) return rc;
	DRT_CUDA(cudaEventSynchronize(r->ev1)); // like Renderer.cu:105-106 the call returns with the frame done
	float ms = 0.f;
	if (ms_or_null || candidate >= 0) DRT_CUDA(cudaEventElapsedTime(&ms, r->ev0, r->ev1));
	if (ms_or_null) *ms_or_null = ms;
	if (candidate >= 0) autotune_record(r, candidate, probing, ms);
	return DRT_OK;
}

extern "C" int drt_renderer_resolve(drt_renderer* r, uint32_t total_samples)
{
	if (!r || !r->width || total_samples == 0) return fail(DRT_ERR_INVALID, "drt_renderer_resolve: bad argument");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DRT_CUDA(launch_resolve(fb_display(r), fb_accum(r), (int)((size_t)r->width * r->height), (float)total_samples, r->stream));
	r->kernel_launches++;
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}

extern "C" int drt_renderer_accum_device_ptr(drt_renderer* r, void** ptr, size_t* bytes)
{
	if (!r || !ptr || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_accum_device_ptr: renderer has no buffer");
	*ptr = fb_accum(r);
	if (bytes) *bytes = r->accum.bytes();
	return DRT_OK;
}

extern "C" int drt_renderer_readback(drt_renderer* r, drt_format format, void* host, size_t bytes)
{
	if (!r || !host || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_readback: bad argument");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	const size_t n = (size_t)r->width * r->height;
	if (format == DRT_RGBA32F) {
		if (bytes < n * 16) return fail(DRT_ERR_INVALID, "drt_renderer_readback: buffer too small");
		DRT_CUDA(cudaMemcpyAsync(host, fb_display(r), n * 16, cudaMemcpyDeviceToHost, r->stream));
	}
	else if (format == DRT_RGBA8 || format == DRT_ACCUM_RGB32F) {
		const size_t need = format == DRT_RGBA8 ? n * 4 : n * 12;
		if (bytes < need) return fail(DRT_ERR_INVALID, "drt_renderer_readback: buffer too small");
		if (r->staging.n < need) DRT_CUDA(r->staging.alloc(need));
		if (format == DRT_RGBA8) DRT_CUDA(launch_pack_rgba8((uint32_t*)r->staging.p, fb_display(r), (int)n, r->stream));
		else DRT_CUDA(launch_pack_rgb32f((float*)r->staging.p, fb_accum(r), (int)n, r->stream));
		r->kernel_launches++;
		DRT_CUDA(cudaMemcpyAsync(host, r->staging.p, need, cudaMemcpyDeviceToHost, r->stream));
	}
	else return fail(DRT_ERR_INVALID, "drt_renderer_readback: unknown format");
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}

extern "C" int drt_renderer_upload_accum(drt_renderer* r, const float* rgb32f, size_t bytes, uint32_t samples_done)
{
	// checkpoint / resume of a progressive render: restores sums + sample count
	if (!r || !rgb32f || !r->width) return fail(DRT_ERR_INVALID, "drt_renderer_upload_accum: bad argument");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	const size_t n = (size_t)r->width * r->height;
	if (bytes < n * 12) return fail(DRT_ERR_INVALID, "drt_renderer_upload_accum: buffer too small");
	std::vector<drt_f4> tmp(n);
	for (size_t i = 0; i < n; i++) { tmp[i].x = rgb32f[3 * i]; tmp[i].y = rgb32f[3 * i + 1]; tmp[i].z = rgb32f[3 * i + 2]; tmp[i].w = 0.f; }
	DRT_CUDA(cudaMemcpyAsync(fb_accum(r), tmp.data(), n * 16, cudaMemcpyHostToDevice, r->stream));
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	r->samples_done = samples_done;
	r->frame_index = samples_done + 1;
	return DRT_OK;
}

extern "C" int drt_renderer_get_counters(drt_renderer* r, drt_counters* out, int reset)
{
	if (!r || !out) return fail(DRT_ERR_INVALID, "drt_renderer_get_counters: null argument");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	static_assert(sizeof(Counters) == sizeof(drt_counters), "counter layouts must match");
	DRT_CUDA(cudaMemcpyAsync(out, r->counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, r->stream));
	if (reset) DRT_CUDA(cudaMemsetAsync(r->counters.p, 0, sizeof(Counters), r->stream));
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}

extern "C" int drt_trace_primary(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, uint32_t frame_index,
	drt_hit* out_host, size_t count)
{
	RenderParams P;
	int rc = make_render_params(r, cam, scene, nullptr, P);
	if (rc) return rc;
	if (!out_host) return fail(DRT_ERR_INVALID, "drt_trace_primary: null output");
	const size_t n = (size_t)r->width * r->height;
	if (count < n) return fail(DRT_ERR_INVALID, "drt_trace_primary: output too small");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	if (r->hits.n < n) DRT_CUDA(r->hits.alloc(n));
	static_assert(sizeof(Hit) == sizeof(drt_hit), "hit layouts must match");
	DRT_CUDA(launch_trace_primary(P, frame_index, r->hits.p, launch_config(r), r->stream));
	r->kernel_launches++;
	DRT_CUDA(cudaMemcpyAsync(out_host, r->hits.p, n * sizeof(Hit), cudaMemcpyDeviceToHost, r->stream));
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}

extern "C" int drt_trace_rays(drt_renderer* r, const drt_scene* scene, const float* origins, const float* directions, size_t n_rays,
	int anyhit, drt_hit* out_host)
{
	if (!r || !scene) return fail(DRT_ERR_INVALID, "drt_trace_rays: null argument");
	if (!scene->built) return fail(DRT_ERR_INVALID, "drt_trace_rays: call drt_scene_build_bvh first");
	if (scene->device != r->device) return fail(DRT_ERR_INVALID, "drt_trace_rays: scene was built on another device");
	if (n_rays == 0) return DRT_OK;
	if (!origins || !directions || !out_host) return fail(DRT_ERR_INVALID, "drt_trace_rays: null ray or output array");
	if (n_rays > 0x7fffffffu / 3) return fail(DRT_ERR_INVALID, "drt_trace_rays: too many rays");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	DevBuf<float> d_o, d_d;
	DevBuf<Hit> d_h;
	DRT_CUDA(d_o.alloc(3 * n_rays)); DRT_CUDA(d_d.alloc(3 * n_rays)); DRT_CUDA(d_h.alloc(n_rays));
	DRT_CUDA(cudaMemcpyAsync(d_o.p, origins, 3 * n_rays * sizeof(float), cudaMemcpyHostToDevice, r->stream));
	DRT_CUDA(cudaMemcpyAsync(d_d.p, directions, 3 * n_rays * sizeof(float), cudaMemcpyHostToDevice, r->stream));
	DRT_CUDA(launch_trace_rays(scene->view, d_o.p, d_d.p, (int)n_rays, anyhit != 0, d_h.p, launch_config(r), r->stream));
	r->kernel_launches++;
	DRT_CUDA(cudaMemcpyAsync(out_host, d_h.p, n_rays * sizeof(Hit), cudaMemcpyDeviceToHost, r->stream));
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}

// ---------------------------------------------------------------------------------------------
// multi-device renderer: N GPUs of one node behind ONE handle, NCCL inside the library
// ---------------------------------------------------------------------------------------------
// One host thread drives all devices: every device gets a replica of the scene (built there) and a renderer; a render call
// enqueues each device's share on that device's stream (sample-range sharding: device g renders frame indices g+1, g+1+N, ...;
// or image-tile sharding: device g renders the 8x4 tiles g, g+N, ... of every frame), then ONE ncclReduce(sum) of the W*H float4
// accumulation buffers into a total buffer on the root device, then the root's resolve kernel (display = total / samples).
// NCCL is bound at run time (dlopen of libnccl.so.2: the library has no link-time dependency on it and loads on hosts without it).
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
	void* handle = nullptr;
	ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
	std::string error;
};

NcclApi& nccl_api()
{
	static NcclApi api;
	static std::once_flag once;
	std::call_once(once, [] {
		for (const char* name : { "libnccl.so.2", "libnccl.so" }) {
			api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
			if (api.handle) break;
		}
		if (!api.handle) { api.error = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?"); return; }
		api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.handle, "ncclCommInitAll");
		api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
		api.GroupStart = (decltype(api.GroupStart))dlsym(api.handle, "ncclGroupStart");
		api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.handle, "ncclGroupEnd");
		api.Reduce = (decltype(api.Reduce))dlsym(api.handle, "ncclReduce");
		api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
		if (!api.CommInitAll || !api.CommDestroy || !api.GroupStart || !api.GroupEnd || !api.Reduce) api.error = "libnccl.so.2 lacks the collective entry points";
	});
	return api;
}

} // namespace

struct drt_multi {
	std::vector<int> devices;
	std::vector<drt_renderer*> r;
	std::vector<drt_scene*> scenes;    // replicas owned by this object, one per device
	std::vector<ncclComm_t> comms;
	DevBuf<drt_f4> total;              // root device: sum over the devices of the accumulation buffers
	int sharding = 0;                  // 0 = sample ranges, 1 = image tiles
	uint32_t samples_total = 0;        // samples per pixel the total buffer holds
	uint32_t width = 0, height = 0;
	int nccl_reduces = 0;
};

#define DRT_NCCL(expr)                                                                                           \
	do {                                                                                                         \
		ncclResult_t e__ = (expr);                                                                               \
		if (e__ != ncclSuccess) {                                                                                \
			NcclApi& a__ = nccl_api();                                                                           \
			return fail(DRT_ERR_CUDA, std::string(#expr) + ": " + (a__.GetErrorString ? a__.GetErrorString(e__) : "NCCL error")); \
		}                                                                                                        \
	} while (0)

extern "C" void drt_multi_destroy(drt_multi* m)
{
	if (!m) return;
	NcclApi& api = nccl_api();
	for (size_t g = 0; g < m->comms.size() && g < m->r.size(); g++)
		if (m->comms[g] && api.CommDestroy) { cudaSetDevice(m->devices[g]); api.CommDestroy(m->comms[g]); }
	const int root = m->r.empty() ? -1 : m->r[0]->device; // (a handle whose creation failed may name devices that do not exist)
	if (root >= 0) { cudaSetDevice(root); m->total.release(); }
	for (drt_renderer* r : m->r) drt_renderer_destroy(r);
	for (drt_scene* s : m->scenes) drt_scene_destroy(s);
	if (root >= 0) cudaSetDevice(root);
	delete m;
}

extern "C" int drt_multi_create(const int32_t* devices, int32_t n, drt_multi** out)
{
	if (!out || !devices || n < 1 || n > 64) return fail(DRT_ERR_INVALID, "drt_multi_create: need 1..64 device ordinals");
	*out = nullptr;
	for (int i = 0; i < n; i++) for (int j = 0; j < i; j++) if (devices[i] == devices[j]) return fail(DRT_ERR_INVALID, "drt_multi_create: a device is named twice");
	drt_multi* m = new drt_multi();
	m->devices.assign(devices, devices + n);
	for (int g = 0; g < n; g++) {
		drt_renderer* r = nullptr;
		const int rc = drt_renderer_create(devices[g], &r);
		if (rc) { drt_multi_destroy(m); return rc; }
		m->r.push_back(r);
	}
	if (n > 1) {
		NcclApi& api = nccl_api();
		if (!api.error.empty()) { const std::string e = api.error; drt_multi_destroy(m); return fail(DRT_ERR_CUDA, "drt_multi_create: " + e); }
		m->comms.assign(n, nullptr);
		const ncclResult_t e = api.CommInitAll(m->comms.data(), n, m->devices.data());
		if (e != ncclSuccess) {
			const std::string msg = api.GetErrorString ? api.GetErrorString(e) : "NCCL error";
			m->comms.clear();
			drt_multi_destroy(m);
			return fail(DRT_ERR_CUDA, "drt_multi_create: ncclCommInitAll: " + msg);
		}
	}
	drt_multi_set_sharding(m, 0);
	*out = m;
	return DRT_OK;
}

extern "C" int drt_multi_device_count(const drt_multi* m) { return m ? (int)m->devices.size() : 0; }
extern "C" int drt_multi_nccl_reduces(const drt_multi* m) { return m ? m->nccl_reduces : 0; }
extern "C" uint32_t drt_multi_sample_count(const drt_multi* m) { return m ? m->samples_total + 1 : 0; }

extern "C" int drt_multi_set_sharding(drt_multi* m, int32_t mode)
{
	if (!m || (mode != 0 && mode != 1)) return fail(DRT_ERR_INVALID, "drt_multi_set_sharding: mode is 0 (sample ranges) or 1 (image tiles)");
	const uint32_t n = (uint32_t)m->devices.size();
	m->sharding = mode;
	for (uint32_t g = 0; g < n; g++) {
		int rc = mode == 0 ? drt_renderer_set_tile_schedule(m->r[g], 0, 1) : drt_renderer_set_frame_schedule(m->r[g], 1, 1);
		if (!rc) rc = mode == 0 ? drt_renderer_set_frame_schedule(m->r[g], g + 1, n) : drt_renderer_set_tile_schedule(m->r[g], g, n);
		if (rc) return rc;
	}
	m->samples_total = 0;
	return DRT_OK;
}

extern "C" int drt_multi_set_flags(drt_multi* m, uint32_t flags)
{
	if (!m) return fail(DRT_ERR_INVALID, "drt_multi_set_flags: null handle");
	for (drt_renderer* r : m->r) drt_renderer_set_flags(r, flags);
	return DRT_OK;
}

extern "C" int drt_multi_set_scene(drt_multi* m, const drt_scene* scene, const drt_bvh_params* params_or_null)
{
	if (!m || !scene) return fail(DRT_ERR_INVALID, "drt_multi_set_scene: null argument");
	for (drt_scene* s : m->scenes) drt_scene_destroy(s);
	m->scenes.clear();
	for (size_t g = 0; g < m->devices.size(); g++) {
		drt_scene* copy = new drt_scene();
		copy->host = scene->host; // the host-side soup, materials and texels; the tree is built on the device it is rendered on
		fill_stats(copy);
		copy->stats.load_ms = scene->stats.load_ms;
		m->scenes.push_back(copy);
		const int rc = drt_scene_build_bvh(copy, params_or_null, m->devices[g]);
		if (rc) return rc;
	}
	return drt_multi_reset(m);
}

extern "C" int drt_multi_resize(drt_multi* m, uint32_t width, uint32_t height)
{
	if (!m) return fail(DRT_ERR_INVALID, "drt_multi_resize: null handle");
	for (drt_renderer* r : m->r) { const int rc = drt_renderer_resize(r, width, height); if (rc) return rc; }
	DRT_CUDA(cudaSetDevice(m->devices[0]));
	if (m->devices.size() > 1) DRT_CUDA(m->total.alloc((size_t)width * height));
	m->width = width; m->height = height;
	return drt_multi_reset(m);
}

extern "C" int drt_multi_reset(drt_multi* m)
{
	if (!m) return fail(DRT_ERR_INVALID, "drt_multi_reset: null handle");
	for (drt_renderer* r : m->r) { const int rc = drt_renderer_reset(r); if (rc) return rc; }
	m->samples_total = 0;
	return DRT_OK;
}

extern "C" int drt_multi_render(drt_multi* m, const drt_camera* cam, const drt_settings* settings, uint32_t n_samples, float* ms_or_null)
{
	if (ms_or_null) *ms_or_null = 0.f;
	if (!m || !cam) return fail(DRT_ERR_INVALID, "drt_multi_render: null argument");
	const size_t n = m->devices.size();
	if (m->scenes.size() != n) return fail(DRT_ERR_INVALID, "drt_multi_render: call drt_multi_set_scene first");
	if (!m->width) return fail(DRT_ERR_INVALID, "drt_multi_render: call drt_multi_resize first");
	if (n == 1) { // one device: exactly the single-device renderer
		const uint32_t before = m->r[0]->samples_done;
		const int rc = drt_renderer_render(m->r[0], cam, m->scenes[0], settings, n_samples, ms_or_null);
		m->samples_total += m->r[0]->samples_done - before;
		return rc;
	}
	drt_settings dflt;
	if (!settings) { default_settings(&dflt); settings = &dflt; }
	// Renderer.cu:82 for the whole frame: never more than max_samples - 1 samples in total
	uint64_t want = n_samples;
	const int64_t room = (int64_t)settings->max_samples - 1 - (int64_t)m->samples_total;
	if (room <= 0) return DRT_OK;
	if (want > (uint64_t)room) want = (uint64_t)room;
	drt_settings per = *settings;
	per.max_samples = 0x7fffffff; // the split below already honours the limit
	const auto t0 = std::chrono::steady_clock::now();
	uint32_t launched_total = 0;
	std::vector<uint32_t> launched(n, 0);
	for (size_t g = 0; g < n; g++) {
		// sample ranges: device g takes every n-th sample of the batch (the first `want % n` devices one more); tiles: every device all samples
		const uint32_t share = m->sharding == 0 ? (uint32_t)(want / n + (g < want % n ? 1 : 0)) : (uint32_t)want;
		if (share == 0) continue;
		int candidate = -1;
		bool probing = false;
		const int rc = render_launch(m->r[g], cam, m->scenes[g], &per, share, &launched[g], &candidate, &probing);
		if (rc) return rc;
		launched_total += m->sharding == 0 ? launched[g] : (g == 0 ? launched[g] : 0);
	}
	m->samples_total += launched_total;
	// ONE reduce of the accumulation buffers to the root, enqueued behind each device's kernel on that device's stream
	NcclApi& api = nccl_api();
	const size_t count = (size_t)m->width * m->height * 4;
	DRT_NCCL(api.GroupStart());
	for (size_t g = 0; g < n; g++) {
		const ncclResult_t e = api.Reduce(fb_accum(m->r[g]), g == 0 ? (void*)m->total.p : nullptr, count, ncclFloat, ncclSum, 0, m->comms[g], m->r[g]->stream);
		if (e != ncclSuccess) { api.GroupEnd(); return fail(DRT_ERR_CUDA, std::string("ncclReduce: ") + (api.GetErrorString ? api.GetErrorString(e) : "NCCL error")); }
	}
	DRT_NCCL(api.GroupEnd());
	m->nccl_reduces++;
	DRT_CUDA(cudaSetDevice(m->devices[0]));
	if (m->samples_total > 0) {
		DRT_CUDA(launch_resolve(fb_display(m->r[0]), m->total.p, (int)((size_t)m->width * m->height), (float)m->samples_total, m->r[0]->stream));
		m->r[0]->kernel_launches++;
	}
	for (size_t g = 0; g < n; g++) {
		DRT_CUDA(cudaSetDevice(m->devices[g]));
		DRT_CUDA(cudaStreamSynchronize(m->r[g]->stream));
	}
	DRT_CUDA(cudaSetDevice(m->devices[0]));
	if (ms_or_null) *ms_or_null = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
	return DRT_OK;
}

extern "C" int drt_multi_readback(drt_multi* m, drt_format format, void* host, size_t bytes)
{
	if (!m || !host || !m->width) return fail(DRT_ERR_INVALID, "drt_multi_readback: bad argument");
	drt_renderer* r0 = m->r[0];
	if (m->devices.size() == 1 || format != DRT_ACCUM_RGB32F) return drt_renderer_readback(r0, format, host, bytes); // display lives on the root
	// accumulation sums of the whole frame = the reduced total
	if (renderer_bind(r0)) return DRT_ERR_CUDA;
	const size_t np = (size_t)m->width * m->height;
	if (bytes < np * 12) return fail(DRT_ERR_INVALID, "drt_multi_readback: buffer too small");
	if (r0->staging.n < np * 12) DRT_CUDA(r0->staging.alloc(np * 12));
	DRT_CUDA(launch_pack_rgb32f((float*)r0->staging.p, m->total.p, (int)np, r0->stream));
	r0->kernel_launches++;
	DRT_CUDA(cudaMemcpyAsync(host, r0->staging.p, np * 12, cudaMemcpyDeviceToHost, r0->stream));
	DRT_CUDA(cudaStreamSynchronize(r0->stream));
	return DRT_OK;
}

extern "C" int drt_multi_get_counters(drt_multi* m, drt_counters* out, int reset)
{
	if (!m || !out) return fail(DRT_ERR_INVALID, "drt_multi_get_counters: null argument");
	memset(out, 0, sizeof(*out));
	for (drt_renderer* r : m->r) {
		drt_counters c;
		const int rc = drt_renderer_get_counters(r, &c, reset);
		if (rc) return rc;
		out->closest_rays += c.closest_rays; out->shadow_rays += c.shadow_rays; out->node_visits += c.node_visits; out->tri_tests += c.tri_tests;
		out->tie_events += c.tie_events; out->alpha_tests += c.alpha_tests; out->texel_clamps += c.texel_clamps;
		if (c.max_stack > out->max_stack) out->max_stack = c.max_stack;
	}
	return DRT_OK;
}

// ---------------------------------------------------------------------------------------------
// allocation-free primary-ray probe (no counterpart in the reference)
// ---------------------------------------------------------------------------------------------
// drt_trace_primary stages the kernel's hit records in a lazily cudaMalloc'ed device buffer (r->hits). Callers that must
// not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same kernel
// with that scratch in a file-scope device array of DRT_NOALLOC_MAX_PIXELS records (4 MiB; every GPU that loads the library
// holds it until exit). All calls share the array and it keeps the last call's contents; the kernel overwrites every record
// it later copies out, so nothing has to start at zero. Calls that overlap on one GPU race on it.
static __device__ __align__(256) unsigned char g_noalloc_hits[(size_t)DRT_NOALLOC_MAX_PIXELS * sizeof(drt::Hit)];
static_assert(sizeof(drt::Hit) == 16, "scratch size is stated in 16-byte hit records");

// Names the array from device code as well, so that the symbol is kept whatever the host side does with it (and gives
// tools a way to scrub it): sets the first n records to "miss".
static __global__ void k_noalloc_scrub(int n)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) reinterpret_cast<int4*>(g_noalloc_hits)[i] = make_int4(-1, 0, 0, 0);
}

#define DRT_NOALLOC_MAX_DEVICES 64
static drt::Hit* g_noalloc_ptr[DRT_NOALLOC_MAX_DEVICES]; // device address of g_noalloc_hits per device (0 = not resolved yet)
static std::mutex g_noalloc_mutex;

// In host code `g_noalloc_hits` is only a handle of the symbol, not a device address: resolve it once per device, with
// that device current, and hand the kernel the resolved pointer.
static int noalloc_scratch(int device, drt::Hit** out)
{
	if (device < 0 || device >= DRT_NOALLOC_MAX_DEVICES) return fail(DRT_ERR_INVALID, "drt_trace_primary_noalloc: device ordinal beyond the scratch table");
	std::lock_guard<std::mutex> lock(g_noalloc_mutex);
	if (!g_noalloc_ptr[device]) {
		int cur = -1;
		DRT_CUDA(cudaGetDevice(&cur));
		if (cur != device) return fail(DRT_ERR_CUDA, "drt_trace_primary_noalloc: the renderer's device is not current");
		void* p = nullptr;
		DRT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_hits));
		if (!p) return fail(DRT_ERR_CUDA, "drt_trace_primary_noalloc: cudaGetSymbolAddress returned a null address");
		g_noalloc_ptr[device] = (drt::Hit*)p;
	}
	*out = g_noalloc_ptr[device];
	return DRT_OK;
}

extern "C" int drt_trace_primary_noalloc(drt_renderer* r, const drt_camera* cam, const drt_scene* scene, uint32_t frame_index,
	drt_hit* out_host, size_t count)
{
	RenderParams P;
	int rc = make_render_params(r, cam, scene, nullptr, P);
	if (rc) return rc;
	if (!out_host) return fail(DRT_ERR_INVALID, "drt_trace_primary_noalloc: null output");
	const size_t n = (size_t)r->width * r->height;
	if (n > (size_t)DRT_NOALLOC_MAX_PIXELS)
		return fail(DRT_ERR_INVALID, "drt_trace_primary_noalloc: " + std::to_string(r->width) + "x" + std::to_string(r->height) + " = " + std::to_string(n) +
			" pixels exceed the static scratch of " + std::to_string((size_t)DRT_NOALLOC_MAX_PIXELS) + " hit records (DRT_NOALLOC_MAX_PIXELS); use drt_trace_primary");
	if (count < n) return fail(DRT_ERR_INVALID, "drt_trace_primary_noalloc: output too small");
	if (renderer_bind(r)) return DRT_ERR_CUDA;
	Hit* scratch = nullptr;
	rc = noalloc_scratch(r->device, &scratch);
	if (rc) return rc;
	DRT_CUDA(launch_trace_primary(P, frame_index, scratch, launch_config(r), r->stream));
	r->kernel_launches++;
	DRT_CUDA(cudaMemcpyAsync(out_host, scratch, n * sizeof(Hit), cudaMemcpyDeviceToHost, r->stream));
	DRT_CUDA(cudaStreamSynchronize(r->stream));
	return DRT_OK;
}
