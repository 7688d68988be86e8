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

// Automatic kernel selection: the sample-pool kernels (variants 22 / 23) replace variants 4 / 15 for launches of at least
// DRT_POOL_MIN_SPP samples on scenes of at least DRT_POOL_MIN_TRIS triangles (measured, profiles/r02c_sample_pool.txt: room.glb
// +3.5 % at 256 spp ... +16 % at 16 spp, dense_monkey / valorant_ghost / animemaid +5-9 % at 64 spp; the 34-triangle cornell_box,
// whose samples are 1.2 rays long, -4.5 %).
#ifndef DRT_POOL_MIN_SPP
#define DRT_POOL_MIN_SPP 8
#endif
#ifndef DRT_POOL_MIN_TRIS
#define DRT_POOL_MIN_TRIS 128
#endif
#ifndef DRT_POOL_TRI2_MIN_TRIS
#define DRT_POOL_TRI2_MIN_TRIS 256 /* pool launches: two-triangle TRI step from this many triangles (measured on the final build: room.glb, 330 triangles, 124.7 -> 123.3 ms at 256 spp, 16.11 -> 15.96 ms at 32 spp; sunshadow_test, 102 triangles, -0.3 %) */
#endif
// Upper bound of the scratch array the renderer allocates for the pool on its own (pixels x R x 16 bytes; beyond it the
// launch runs variants 4 / 15).
#ifndef DRT_POOL_MAX_SCRATCH_BYTES
#define DRT_POOL_MAX_SCRATCH_BYTES (4ull << 30)
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
	size_t pool_scratch_bytes = 0; // capacity of RenderParams::pool_scratch (sample-pool kernel)
};

// Sample-pool kernel: pool samples per pixel for a launch of n samples (0: the launch is too short for a pool) and the scratch
// bytes a launch over `tiles` 8x4 tiles needs.
int sample_pool_size(uint32_t n_samples, int override_r);
size_t sample_pool_scratch_bytes(long long tiles, int pool_r);
// Bytes of the regrouping kernel's traversal-stack overflow area for a device of `sms` multiprocessors (RenderParams::rg_overflow).
size_t regroup_overflow_bytes(int sms);
// Cost-ordered launches (RenderParams::block_order / block_cost): order[0 .. n-1] = group indices by descending cost.
cudaError_t launch_block_order(const uint32_t* cost, uint32_t* order, int n, cudaStream_t s);
// (honoured by the fixed-tile phased kernels: launch_render's variant_out in {2 .. 8, 15, 17, 18, 21, 22, 23})
inline bool variant_honours_block_order(int v) { return (v >= 2 && v <= 8) || v == 15 || v == 17 || v == 18 || v == 21 || v == 22 || v == 23; }
#define DRT_ORDER_MIN_SPP 16
#define DRT_ORDER_PROBE_MIN_SPP 64 /* launches this long pay a one-sample probe launch to be ordered themselves */
cudaError_t launch_build_pass(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n, cudaStream_t s);
// Single-launch build for small scenes (k_build_small: one block runs the whole level loop). `result` = 6 device ints:
// bvh2 nodes, wide nodes, levels, wide levels, triangles emitted, 1 if the final index array is A.idx_out.
#define DRT_BUILD_SMALL_THREADS 1024
#define DRT_BUILD_SMALL_RESULT_INTS 64 /* 6 used; the rest carries per-pass clocks in DRT_BUILD_SMALL_TIMING measurement builds */
#ifndef DRT_BUILD_SMALL_MAX_TRIS
#define DRT_BUILD_SMALL_MAX_TRIS 2048
#endif
cudaError_t launch_build_small(const BuildArrays& A, int* result, cudaStream_t s);

cudaError_t launch_vertex_records(const drt_f4* tris, const float* pos, drt_f4* out, int n, cudaStream_t s);
cudaError_t launch_expand_nodes(const WideNode* wide, drt_f4* out_f32, int n, cudaStream_t s);
// variant_out (may be null): the kernel variant the launch ran after automatic selection and fall-backs (-2 = extended materials)
cudaError_t launch_render(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s, int* launches, int* variant_out = nullptr);
cudaError_t launch_trace_primary(const RenderParams& P, uint32_t frame_index, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s);
cudaError_t launch_trace_rays(const SceneView& sc, const float* org_dev, const float* dir_dev, int n, bool anyhit, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s);
cudaError_t launch_resolve(drt_f4* display, const drt_f4* accum, int n_pixels, float total_samples, cudaStream_t s);
cudaError_t launch_pack_rgba8(uint32_t* out, const drt_f4* display, int n_pixels, cudaStream_t s);
cudaError_t launch_pack_rgb32f(float* out, const drt_f4* accum, int n_pixels, cudaStream_t s);

} // namespace drt
