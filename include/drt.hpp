// drt.hpp -- thin header-only C++ RAII wrapper over the C ABI (drt.h), mirroring the names of the reference's
// classes so that a caller written against DustRayTracer's Scene / BVHBuilder / Renderer / Camera ports line by line:
//   Scene::loadGLTFmodel                 DustRayTracer/src/Core/Scene/Scene.cuh:51
//   BVHBuilder::build / buildIterative   DustRayTracer/src/Core/BVH/BVHBuilder.cuh:21-22
//   Renderer::ResizeBuffer / Render / resetAccumulationBuffer / getSampleCount / getBufferWidth / getBufferHeight
//                                        DustRayTracer/src/Core/Renderer.hpp:18-30
//   Camera, RendererSettings             DustRayTracer/src/Core/Scene/Camera.cuh:32-46, RendererSettings.h:22-34
// Errors become drt::Error (the reference prints and exit(99)s: Editor/Common/CudaCommon.cu:4-13).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "drt.h"

namespace drt {

struct Error : std::runtime_error {
	int status;
	Error(int s, const char* what) : std::runtime_error(what ? what : "drt error"), status(s) {}
};

inline void check(int rc)
{
	if (rc != DRT_OK) throw Error(rc, drt_last_error());
}

struct Camera : drt_camera {
	Camera() { drt_default_camera(this); }
};

struct RendererSettings : drt_settings {
	RendererSettings() { drt_default_settings(this); }
};

class Scene {
public:
	Scene() = default;
	Scene(const Scene&) = delete;
	Scene& operator=(const Scene&) = delete;
	~Scene() { drt_scene_destroy(h_); }

	bool loadGLTFmodel(const char* filepath, uint32_t load_flags = DRT_LOAD_REFERENCE_COMPAT)
	{
		drt_scene_destroy(h_);
		h_ = nullptr;
		return drt_scene_load_gltf_ex(filepath, load_flags, &h_) == DRT_OK; // the reference returns bool and its caller ignores it
	}
	void fromTriangles(const float* pos9, const float* uv6, const int32_t* material_index, size_t n, const drt_material* mats, size_t n_mats)
	{
		drt_scene_destroy(h_);
		h_ = nullptr;
		check(drt_scene_from_triangles(pos9, uv6, material_index, n, mats, n_mats, &h_));
	}
	drt_scene_stats stats() const
	{
		drt_scene_stats s;
		check(drt_scene_get_stats(h_, &s));
		return s;
	}
	drt_scene* handle() const { return h_; }

private:
	drt_scene* h_ = nullptr;
};

// Both entry points run the GPU binned-SAH build (there is one tree numbering; the reference's two differ only in node order).
class BVHBuilder {
public:
	int m_BinCount = 8;                   // BVHBuilder.cuh:16 (informational here)
	int m_TargetLeafPrimitivesCount = 6;  // BVHBuilder.cuh:17 (informational here)
	int device = 0;
	void build(Scene& scene) const
	{
		drt_bvh_params p = {};
		p.bin_count = m_BinCount;
		p.leaf_target = m_TargetLeafPrimitivesCount;
		check(drt_scene_build_bvh(scene.handle(), &p, device));
	}
	void buildIterative(Scene& scene) const { build(scene); }
};

class Renderer {
public:
	RendererSettings m_RendererSettings;  // public like Renderer.hpp:30

	explicit Renderer(int device = 0) { check(drt_renderer_create(device, &h_)); }
	Renderer(const Renderer&) = delete;
	Renderer& operator=(const Renderer&) = delete;
	~Renderer() { drt_renderer_destroy(h_); }

	void ResizeBuffer(uint32_t width, uint32_t height) { check(drt_renderer_resize(h_, width, height)); }
	// delta = device time of the launch in ms (Renderer.cu:105-108); n_samples > 1 batches frames in one launch
	void Render(const Camera* cam, const Scene& scene, float* delta, uint32_t n_samples = 1)
	{
		check(drt_renderer_render(h_, cam, scene.handle(), &m_RendererSettings, n_samples, delta));
	}
	void resetAccumulationBuffer() { check(drt_renderer_reset(h_)); }
	uint32_t getSampleCount() const { return drt_renderer_sample_count(h_); }
	uint32_t getBufferWidth() const { return drt_renderer_width(h_); }
	uint32_t getBufferHeight() const { return drt_renderer_height(h_); }

	// replaces glGetTexImage on GetRenderTargetImage_name() (EditorLayer.cpp:85-97)
	std::vector<float> readbackRGBA32F()
	{
		std::vector<float> px((size_t)getBufferWidth() * getBufferHeight() * 4);
		check(drt_renderer_readback(h_, DRT_RGBA32F, px.data(), px.size() * sizeof(float)));
		return px;
	}
	std::vector<uint8_t> readbackRGBA8()
	{
		std::vector<uint8_t> px((size_t)getBufferWidth() * getBufferHeight() * 4);
		check(drt_renderer_readback(h_, DRT_RGBA8, px.data(), px.size()));
		return px;
	}
	drt_counters counters(bool reset = false)
	{
		drt_counters c;
		check(drt_renderer_get_counters(h_, &c, reset ? 1 : 0));
		return c;
	}
	void setFrameSchedule(uint32_t first_frame, uint32_t stride) { check(drt_renderer_set_frame_schedule(h_, first_frame, stride)); }
	void setTileSchedule(uint32_t first_tile, uint32_t stride) { check(drt_renderer_set_tile_schedule(h_, first_tile, stride)); }
	// image-tile sharding without a collective: other renderers store their tiles into THIS renderer's buffers (drt.h)
	void exportFrameBuffers(void* handle_out /* DRT_FRAME_HANDLE_BYTES */) { check(drt_renderer_export_frame_buffers(h_, handle_out)); }
	void importFrameBuffers(const void* handle) { check(drt_renderer_import_frame_buffers(h_, handle)); }
	void attachFrameBuffers(Renderer* owner) { check(drt_renderer_attach_frame_buffers(h_, owner ? owner->h_ : nullptr)); }
	void setTuning(int32_t variant, int32_t block_threads = 0) { check(drt_renderer_set_tuning(h_, variant, block_threads)); }
	drt_renderer* handle() const { return h_; }

private:
	drt_renderer* h_ = nullptr;
};

// N GPUs of one node behind one object (drt_multi_*: replicated scene, sample-range or image-tile sharding, one NCCL reduce of
// the accumulation buffers per Render call, inside the library). The methods mirror Renderer's.
class MultiRenderer {
public:
	RendererSettings m_RendererSettings;

	explicit MultiRenderer(const std::vector<int32_t>& devices) { check(drt_multi_create(devices.data(), (int32_t)devices.size(), &h_)); }
	MultiRenderer(const MultiRenderer&) = delete;
	MultiRenderer& operator=(const MultiRenderer&) = delete;
	~MultiRenderer() { drt_multi_destroy(h_); }

	void setScene(const Scene& scene) { check(drt_multi_set_scene(h_, scene.handle(), nullptr)); } // replicates and builds on every device
	void setSharding(bool tiles) { check(drt_multi_set_sharding(h_, tiles ? 1 : 0)); }
	void ResizeBuffer(uint32_t width, uint32_t height) { check(drt_multi_resize(h_, width, height)); w_ = width; hgt_ = height; }
	void Render(const Camera* cam, float* delta, uint32_t n_samples = 1) { check(drt_multi_render(h_, cam, &m_RendererSettings, n_samples, delta)); }
	void resetAccumulationBuffer() { check(drt_multi_reset(h_)); }
	uint32_t getSampleCount() const { return drt_multi_sample_count(h_); }
	int deviceCount() const { return drt_multi_device_count(h_); }
	std::vector<float> readbackRGBA32F()
	{
		std::vector<float> px((size_t)w_ * hgt_ * 4);
		check(drt_multi_readback(h_, DRT_RGBA32F, px.data(), px.size() * sizeof(float)));
		return px;
	}
	std::vector<uint8_t> readbackRGBA8()
	{
		std::vector<uint8_t> px((size_t)w_ * hgt_ * 4);
		check(drt_multi_readback(h_, DRT_RGBA8, px.data(), px.size()));
		return px;
	}
	drt_counters counters(bool reset = false)
	{
		drt_counters c;
		check(drt_multi_get_counters(h_, &c, reset ? 1 : 0));
		return c;
	}

private:
	drt_multi* h_ = nullptr;
	uint32_t w_ = 0, hgt_ = 0;
};

} // namespace drt
