// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Host-side driver around the UNMODIFIED reference sources, which are compiled where they lie
// under /root/reference by oracle/Makefile (g++ -x c++, thrust CPP device system, no CUDA runtime).
// Nothing from the reference is copied here: this file only #includes the reference headers and
// calls the reference's own functions:
//   Scene::loadGLTFmodel            DustRayTracer/src/Core/Scene/Scene.cu:181-317
//   BVHBuilder::build/buildIterative DustRayTracer/src/Core/BVH/BVHBuilder.cu:11-145
//   RayGen                          DustRayTracer/src/Core/Kernel/Shaders/RayGen.cuh:63-172
//   TraceRay / RayTest              DustRayTracer/src/Core/Kernel/TraceRay.cu:15-38
//   Camera::GetRay                  DustRayTracer/src/Core/Scene/Camera.cu:82-123
//   Bounds3f::intersect             DustRayTracer/src/Core/Bounds.cu:18-41
//   Intersection                    DustRayTracer/src/Core/Kernel/Shaders/Intersection.cu:4-36
//   pcg_hash / randomFloat / ...    DustRayTracer/src/Core/CudaMath/Random.cu:6-66
// The pixel loop + accumulation restates RenderKernel.cu:20-35 (that file needs GL headers and a
// CUDA surface, so it cannot be compiled here); frame-index semantics follow Renderer.cu:80-117.
//
// Exposed as a plain C ABI (libdrt_ref.so) for ctypes: used by tests/ as the parity checker and by
// bench.py as the CPU baseline ("kind": "reference").

#define private public   // Texture::d_data is private (Texture.cuh:18); the driver only reads it.
#include "Core/Scene/Texture.cuh"
#undef private

#include "Core/Scene/Scene.cuh"
#include "Core/Scene/Camera.cuh"
#include "Core/BVH/BVHNode.cuh"
#include "Core/BVH/BVHBuilder.cuh"
#include "Core/Ray.cuh"
#include "Core/HitPayload.cuh"
#include "Core/Kernel/TraceRay.cuh"
#include "Core/Kernel/Shaders/Intersection.cuh"
#include "Core/CudaMath/Random.cuh"

#include <atomic>
#include <chrono>
#include <cstring>
#include <map>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

// ---- ray counting: wrap the calls RayGen makes (RayGen.cuh:90,126) without touching the source ----
static thread_local unsigned long long g_closest_rays = 0, g_shadow_rays = 0;
static inline void count_closest() { ++g_closest_rays; }
static inline void count_shadow() { ++g_shadow_rays; }
#define TraceRay(...) (count_closest(), TraceRay(__VA_ARGS__))
#define RayTest(...) (count_shadow(), RayTest(__VA_ARGS__))
#include "Core/Kernel/Shaders/RayGen.cuh"
#undef TraceRay
#undef RayTest

namespace {

struct RefScene {
	Scene* scene = nullptr;
	std::vector<Triangle> load_order;   // snapshot before the builder permutes m_PrimitivesBuffer
	std::vector<int> canonical_of_slot; // current slot -> index in load order
	Camera* cam = nullptr;              // "managed" camera object the reference passes by pointer
};

// Content key of a Triangle without its 4 padding bytes (centroid ends at byte 12, vertex0 starts at
// byte 16 because float2 is 8-byte aligned; the padding is uninitialised and must not be compared).
struct TriKey {
	unsigned char b[sizeof(Triangle) - 4];
	void set(const Triangle& t)
	{
		static_assert(sizeof(Triangle) == 128, "Triangle layout");
		const unsigned char* p = (const unsigned char*)&t;
		memcpy(b, p, 12);
		memcpy(b + 12, p + 16, sizeof(Triangle) - 16);
	}
	bool operator<(const TriKey& o) const { return memcmp(b, o.b, sizeof(b)) < 0; }
};

SceneData make_scenedata(const RefScene* rs, const RendererSettings& settings)
{
	// Restates InvokeRenderKernel's packing (RenderKernel.cu:41-51).
	const Scene& scene = *rs->scene;
	SceneData sd;
	sd.DeviceBVHNodesBuffer = thrust::raw_pointer_cast(scene.m_BVHNodes.data());
	sd.DeviceTextureBufferPtr = thrust::raw_pointer_cast(scene.m_Textures.data());
	sd.DeviceMeshBufferPtr = thrust::raw_pointer_cast(scene.m_Meshes.data());
	sd.DeviceMaterialBufferPtr = thrust::raw_pointer_cast(scene.m_Material.data());
	sd.DevicePrimitivesBuffer = thrust::raw_pointer_cast(scene.m_PrimitivesBuffer.data());
	sd.DeviceMeshBufferSize = scene.m_Meshes.size();
	sd.DevicePrimitivesBufferSize = scene.m_PrimitivesBuffer.size();
	sd.DeviceBVHNodesBufferSize = scene.m_BVHNodes.size();
	sd.DeviceBVHTreeRootPtr = scene.d_BVHTreeRoot;
	sd.RenderSettings = settings;
	return sd;
}

void refresh_canonical(RefScene* rs)
{
	const auto& prims = rs->scene->m_PrimitivesBuffer;
	std::multimap<TriKey, int> by_content;
	for (size_t i = 0; i < rs->load_order.size(); i++) {
		TriKey k;
		k.set(rs->load_order[i]);
		by_content.insert({ k, (int)i });
	}
	rs->canonical_of_slot.assign(prims.size(), -1);
	for (size_t s = 0; s < prims.size(); s++) {
		Triangle t = prims[s];
		TriKey k;
		k.set(t);
		auto it = by_content.find(k); // exact duplicates: lowest unused canonical id first
		if (it != by_content.end()) {
			rs->canonical_of_slot[s] = it->second;
			by_content.erase(it);
		}
	}
}

} // namespace

extern "C" {

int ref_sizeof(int what)
{
	switch (what) {
	case 0: return (int)sizeof(Triangle);
	case 1: return (int)sizeof(BVHNode);
	case 2: return (int)sizeof(Material);
	case 3: return (int)sizeof(Camera);
	case 4: return (int)sizeof(RendererSettings);
	case 5: return (int)sizeof(Ray);
	case 6: return (int)sizeof(HitPayload);
	case 7: return (int)sizeof(SceneData);
	case 8: return (int)sizeof(Texture);
	default: return -1;
	}
}

void* ref_scene_load(const char* path)
{
	RefScene* rs = new RefScene();
	rs->scene = new Scene();
	rs->scene->loadGLTFmodel(path);
	const auto& prims = rs->scene->m_PrimitivesBuffer;
	rs->load_order.assign(prims.begin(), prims.end());
	rs->canonical_of_slot.resize(prims.size());
	for (size_t i = 0; i < prims.size(); i++) rs->canonical_of_slot[i] = (int)i;
	rs->cam = new Camera(); // Managed::operator new -> cudaMallocManaged (stubbed to malloc)
	return rs;
}

void ref_scene_free(void* h)
{
	RefScene* rs = (RefScene*)h;
	if (!rs) return;
	delete rs->cam;
	delete rs->scene;
	delete rs;
}

int ref_scene_tri_count(void* h) { return (int)((RefScene*)h)->scene->m_PrimitivesBuffer.size(); }
int ref_scene_material_count(void* h) { return (int)((RefScene*)h)->scene->m_Material.size(); }
int ref_scene_texture_count(void* h) { return (int)((RefScene*)h)->scene->m_Textures.size(); }
int ref_scene_mesh_count(void* h) { return (int)((RefScene*)h)->scene->m_Meshes.size(); }
int ref_scene_node_count(void* h) { return (int)((RefScene*)h)->scene->m_BVHNodes.size(); }

// Raw 128-byte Triangle records in their CURRENT order (load order before a build, leaf order after).
void ref_scene_get_triangles(void* h, void* out)
{
	RefScene* rs = (RefScene*)h;
	const auto& prims = rs->scene->m_PrimitivesBuffer;
	Triangle* o = (Triangle*)out;
	for (size_t i = 0; i < prims.size(); i++) o[i] = prims[i];
}

// Raw 128-byte Triangle records in LOAD order (the canonical numbering).
void ref_scene_get_triangles_load_order(void* h, void* out)
{
	RefScene* rs = (RefScene*)h;
	memcpy(out, rs->load_order.data(), rs->load_order.size() * sizeof(Triangle));
}

void ref_scene_get_canonical_ids(void* h, int* out)
{
	RefScene* rs = (RefScene*)h;
	memcpy(out, rs->canonical_of_slot.data(), rs->canonical_of_slot.size() * sizeof(int));
}

// Raw 44-byte Material records.
void ref_scene_get_materials(void* h, void* out)
{
	RefScene* rs = (RefScene*)h;
	const auto& m = rs->scene->m_Material;
	Material* o = (Material*)out;
	for (size_t i = 0; i < m.size(); i++) o[i] = m[i];
}

void ref_scene_get_texture_info(void* h, int idx, int* w, int* hgt, int* comps)
{
	RefScene* rs = (RefScene*)h;
	Texture t = rs->scene->m_Textures[idx];
	*w = t.width; *hgt = t.height; *comps = t.componentCount;
}

void ref_scene_get_texture_pixels(void* h, int idx, unsigned char* out)
{
	RefScene* rs = (RefScene*)h;
	Texture t = rs->scene->m_Textures[idx];
	memcpy(out, t.d_data, (size_t)t.width * t.height * t.componentCount);
}

// Raw 44-byte BVHNode records (root is the LAST node: BVHBuilder.cu:85,142; TraceRay.cu:20).
void ref_scene_get_nodes(void* h, void* out)
{
	RefScene* rs = (RefScene*)h;
	const auto& n = rs->scene->m_BVHNodes;
	BVHNode* o = (BVHNode*)out;
	for (size_t i = 0; i < n.size(); i++) o[i] = n[i];
}

// Runs the reference builder. iterative=1 -> buildIterative (what the editor calls,
// EditorLayer.cpp:52-55); iterative=0 -> build (recursive numbering). Returns node count.
int ref_scene_build_bvh(void* h, int bin_count, int leaf_target, int iterative, double* seconds)
{
	RefScene* rs = (RefScene*)h;
	rs->scene->m_BVHNodes.clear();
	BVHBuilder builder;
	builder.m_BinCount = bin_count;
	builder.m_TargetLeafPrimitivesCount = leaf_target;
	auto t0 = std::chrono::steady_clock::now();
	if (iterative)
		rs->scene->d_BVHTreeRoot = builder.buildIterative(rs->scene->m_PrimitivesBuffer, rs->scene->m_BVHNodes);
	else
		rs->scene->d_BVHTreeRoot = builder.build(rs->scene->m_PrimitivesBuffer, rs->scene->m_BVHNodes);
	auto t1 = std::chrono::steady_clock::now();
	if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
	refresh_canonical(rs);
	return (int)rs->scene->m_BVHNodes.size();
}

// cam_pod: 88 bytes, field-for-field the reference Camera (Camera.cuh:30-47).
static void set_camera(RefScene* rs, const void* cam_pod)
{
	static_assert(sizeof(Camera) == 88, "Camera POD size");
	memcpy((void*)rs->cam, cam_pod, sizeof(Camera));
}

void ref_default_camera(void* out88)
{
	Camera* c = new Camera();
	c->m_movement_speed = 10.0; // EditorLayer.cpp:38-40
	c->defocus_angle = 0.f;
	c->focus_dist = 10.f;
	memcpy(out88, (void*)c, sizeof(Camera));
	delete c;
}

void ref_default_settings(void* out64)
{
	static_assert(sizeof(RendererSettings) == 64, "RendererSettings POD size");
	RendererSettings s;
	memset(out64, 0, sizeof(RendererSettings));
	memcpy(out64, &s, sizeof(RendererSettings));
}

// Primary rays exactly as RayGen issues them (RayGen.cuh:65-78,90), closest hit only.
// Outputs per pixel p = x + y*W: canonical primitive id (-1 = miss), t, u, v (UVW.y, UVW.z),
// and optionally the ray (origin xyz, direction xyz) for debugging.
void ref_trace_primary(void* h, const void* cam_pod, int W, int H, unsigned frameidx,
	int* prim_id, float* t, float* u, float* v, float* ray_od6, int nthreads)
{
	RefScene* rs = (RefScene*)h;
	set_camera(rs, cam_pod);
	RendererSettings settings;
	const SceneData sd = make_scenedata(rs, settings);
	const Camera* cam = rs->cam;
#ifdef _OPENMP
	if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 16)
	for (int y = 0; y < H; y++) {
		for (int x = 0; x < W; x++) {
			float2 screen_uv = { (float(x) / (uint32_t)W), (float(y) / (uint32_t)H) };
			screen_uv = screen_uv * 2 - 1;
			uint32_t seed = x + y * W;
			seed *= frameidx;
			Ray ray = cam->GetRay(screen_uv, (uint32_t)W, (uint32_t)H, seed);
			ray.interval = Interval(-1, FLT_MAX);
			HitPayload payload = TraceRay(ray, &sd);
			size_t p = (size_t)x + (size_t)y * W;
			if (payload.primitiveptr == nullptr) {
				prim_id[p] = -1; t[p] = -1; u[p] = 0; v[p] = 0;
			}
			else {
				size_t slot = payload.primitiveptr - sd.DevicePrimitivesBuffer;
				prim_id[p] = rs->canonical_of_slot[slot];
				t[p] = payload.hit_distance;
				u[p] = payload.UVW.y;
				v[p] = payload.UVW.z;
			}
			if (ray_od6) {
				float3 o = ray.getOrigin(), d = ray.getDirection();
				float* r = ray_od6 + 6 * p;
				r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
			}
		}
	}
}

// Accumulates frames [frame_begin, frame_begin+frame_count) into acc (W*H float3, caller-zeroed for
// a fresh render). One pass over the image per frame, like one kernel launch per frame
// (Renderer.cu:100-116; RenderKernel.cu:27-29). counters[0]=TraceRay calls, [1]=RayTest calls.
// Returns seconds spent in the pixel loops.
double ref_render(void* h, const void* cam_pod, const void* settings_pod, int W, int H,
	unsigned frame_begin, unsigned frame_count, float* acc, int nthreads, unsigned long long* counters)
{
	RefScene* rs = (RefScene*)h;
	set_camera(rs, cam_pod);
	RendererSettings settings;
	memcpy(&settings, settings_pod, sizeof(RendererSettings));
	const SceneData sd = make_scenedata(rs, settings);
	const Camera* cam = rs->cam;
	float3* accum = (float3*)acc;
	unsigned long long closest = 0, shadow = 0;
#ifdef _OPENMP
	if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
	auto t0 = std::chrono::steady_clock::now();
	for (unsigned f = 0; f < frame_count; f++) {
		const uint32_t frameidx = frame_begin + f;
#pragma omp parallel reduction(+ : closest, shadow)
		{
			g_closest_rays = 0; g_shadow_rays = 0;
#pragma omp for schedule(dynamic, 16)
			for (int j = 0; j < H; j++) {
				for (int i = 0; i < W; i++) {
					float3 fcolor = RayGen(i, j, W, H, cam, frameidx, sd);
					accum[i + (size_t)j * W] += fcolor;
				}
			}
			closest += g_closest_rays; shadow += g_shadow_rays;
		}
	}
	auto t1 = std::chrono::steady_clock::now();
	if (counters) { counters[0] = closest; counters[1] = shadow; }
	return std::chrono::duration<double>(t1 - t0).count();
}

int ref_max_threads()
{
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}

// ---- known-answer probes of the leaf functions ----
unsigned ref_pcg_hash(unsigned x) { return pcg_hash(x); }

float ref_random_float(unsigned* seed)
{
	uint32_t s = *seed;
	float r = randomFloat(s);
	*seed = s;
	return r;
}

void ref_random_unit_vec3(unsigned* seed, float* out3)
{
	uint32_t s = *seed;
	float3 r = randomUnitVec3(s);
	*seed = s;
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

void ref_random_unit_sphere_vec3(unsigned* seed, float* out3)
{
	uint32_t s = *seed;
	float3 r = randomUnitSphereVec3(s);
	*seed = s;
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

// box6 = pMin xyz, pMax xyz; ray = origin xyz, direction xyz.
float ref_bounds_intersect(const float* box6, const float* o3, const float* d3)
{
	Bounds3f b(make_float3(box6[0], box6[1], box6[2]), make_float3(box6[3], box6[4], box6[5]));
	Ray ray(make_float3(o3[0], o3[1], o3[2]), make_float3(d3[0], d3[1], d3[2]));
	return b.intersect(ray);
}

// tri9 = v0 xyz, v1 xyz, v2 xyz. Returns 1 on hit and fills t, UVW (3 floats).
int ref_intersection(const float* tri9, const float* o3, const float* d3, float* t, float* uvw3)
{
	Triangle tri;
	tri.vertex0.position = make_float3(tri9[0], tri9[1], tri9[2]);
	tri.vertex1.position = make_float3(tri9[3], tri9[4], tri9[5]);
	tri.vertex2.position = make_float3(tri9[6], tri9[7], tri9[8]);
	Ray ray(make_float3(o3[0], o3[1], o3[2]), make_float3(d3[0], d3[1], d3[2]));
	ShortHitPayload p = Intersection(ray, &tri);
	if (p.primitiveptr == nullptr) return 0;
	*t = p.hit_distance;
	uvw3[0] = p.UVW.x; uvw3[1] = p.UVW.y; uvw3[2] = p.UVW.z;
	return 1;
}

// Camera::GetRay for one pixel; returns the seed after the call.
unsigned ref_camera_get_ray(const void* cam_pod, int x, int y, int W, int H, unsigned frameidx, float* od6)
{
	Camera* cam = new Camera();
	memcpy((void*)cam, cam_pod, sizeof(Camera));
	float2 screen_uv = { (float(x) / (uint32_t)W), (float(y) / (uint32_t)H) };
	screen_uv = screen_uv * 2 - 1;
	uint32_t seed = x + y * W;
	seed *= frameidx;
	Ray ray = cam->GetRay(screen_uv, (uint32_t)W, (uint32_t)H, seed);
	float3 o = ray.getOrigin(), d = ray.getDirection();
	od6[0] = o.x; od6[1] = o.y; od6[2] = o.z; od6[3] = d.x; od6[4] = d.y; od6[5] = d.z;
	delete cam;
	return seed;
}

// One RayGen call (one pixel-sample), for spot checks.
void ref_raygen(void* h, const void* cam_pod, const void* settings_pod, int x, int y, int W, int H,
	unsigned frameidx, float* rgb3)
{
	RefScene* rs = (RefScene*)h;
	set_camera(rs, cam_pod);
	RendererSettings settings;
	memcpy(&settings, settings_pod, sizeof(RendererSettings));
	const SceneData sd = make_scenedata(rs, settings);
	float3 c = RayGen(x, y, W, H, rs->cam, frameidx, sd);
	rgb3[0] = c.x; rgb3[1] = c.y; rgb3[2] = c.z;
}

} // extern "C"
