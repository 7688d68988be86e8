// drt_kernels.h -- launch entry points of drt_kernels.cu (host-callable, one CUDA stream each).
#pragma once
#include <cuda_runtime.h>
#include "drt_types.h"
#include "drt_build_driver.h"

namespace drt {

// Largest launch (samples per pixel) for which the renderer's auto-tuner considers the persistent pixel-queue kernel.
#ifndef DRT_QUEUE_MAX_SPP_DEFAULT
#define DRT_QUEUE_MAX_SPP_DEFAULT 8
#endif

// Automatic kernel selection: from this many triangles on, the two-triangle TRI step (variant 15) replaces variant 4
// (measured: +4 % to +10 % on the 6 k - 16 k triangle scenes, -1 % on the 34- and 330-triangle ones).
#ifndef DRT_TRI2_MIN_TRIS
#define DRT_TRI2_MIN_TRIS 1024
#endif

struct LaunchConfig {
	bool strict = false;  // DRT_FLAG_STRICT
	bool rtl = false;     // DRT_FLAG_HOST_ARG_ORDER
	bool stats = false;   // DRT_FLAG_STATS
	bool watertight = false; // DRT_FLAG_WATERTIGHT: watertight TRI step (production arithmetic only)
	bool ext_materials = false; // DRT_FLAG_EXTENDED_MATERIALS: emissive / metal / dielectric shading (reference-shaped loop kernel)
	int variant = 4;      // render kernel variant (see drt_kernels.cu); 4 = default
	bool auto_variant = false; // variant 4 was not asked for explicitly: scenes of DRT_TRI2_MIN_TRIS triangles or more run variant 15
	int sched_bias[3] = { 0, 0, 0 }; // phased kernels: [0] = sticky re-issue threshold (0 = default 6, 33 = never)
	int block_threads = 0; // 0 = variant default
	int resident_sms = 0;  // SM count of the device (grid of the persistent kernel)
};

// Bytes of the regrouping kernel's traversal-stack overflow area for a device of `sms` multiprocessors (RenderParams::rg_overflow).
size_t regroup_overflow_bytes(int sms);
cudaError_t launch_build_pass(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n, cudaStream_t s);
cudaError_t launch_vertex_records(const drt_f4* tris, const float* pos, drt_f4* out, int n, cudaStream_t s);
cudaError_t launch_expand_nodes(const WideNode* wide, drt_f4* out_f32, int n, cudaStream_t s);
cudaError_t launch_render(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s, int* launches);
cudaError_t launch_trace_primary(const RenderParams& P, uint32_t frame_index, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s);
cudaError_t launch_trace_rays(const SceneView& sc, const float* org_dev, const float* dir_dev, int n, bool anyhit, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s);
cudaError_t launch_resolve(drt_f4* display, const drt_f4* accum, int n_pixels, float total_samples, cudaStream_t s);
cudaError_t launch_pack_rgba8(uint32_t* out, const drt_f4* display, int n_pixels, cudaStream_t s);
cudaError_t launch_pack_rgb32f(float* out, const drt_f4* accum, int n_pixels, cudaStream_t s);

} // namespace drt
