// drt_kernels.cu -- sm_100a kernels of the hot path and of the GPU BVH builder.
//
// Render kernels (replace the reference's single `kernel`, DustRayTracer/src/Core/Kernel/RenderKernel.cu:20-35,
// launched once per sample with 8x8 blocks, Renderer.hpp:43-44):
//   variant 0  k_render_flat   one thread per pixel, ALL samples of the launch in one flat state machine
//                              with a single traversal call site (drt_shade.cuh render_pixel_flat):
//                              accumulator lives in registers, one 16-byte accum read + write and one
//                              16-byte display write per pixel per LAUNCH instead of 40 B per sample.
//   variant 1  k_render_loop   one thread per pixel, sample loop around the reference-shaped bounce loop
//                              (path_radiance); also serves the debug views.
//   variant 2  phased          per-warp phase voting over NODE / TRI / SHADE steps (drt_phased.cuh), expanded
//                              fp32 nodes when the scene has them, else compressed nodes  [default]
//   variant 3  phased, Q8      the same on the compressed nodes regardless of scene size
//   variant 4..6 phased        variant 2 compiled for 6..8 resident 128-thread blocks per SM (80 / 72 / 64 registers);
//                              4 is the default: issue-slot utilisation rises with occupancy up to 6 blocks (profiles/)
//   variant 7, 8 phased, Q8    variant 3 compiled for 8 / 6 blocks per SM; 7 is what scenes without expanded nodes get
//                              (soup-10M is latency-bound: +17 % from the extra resident warps)
//   variant 15 phased, TRI2   variant 4 whose TRI step tests up to two pending triangles (independent instruction streams, half the voted
//                              steps for lanes with two or more pending): what automatic selection runs on scenes of >= 1024 triangles
//   variant 11..14 cold state   variant 4 / 7 with the SHADE-only lane state in shared memory (k_render_cold: 11 / 12 / 13 = fp32 nodes at
//                              8 / 7 / 6 blocks per SM, 14 = compressed nodes); measured alternatives
//   variant 9, 10 pixel queue  persistent form of the cold-state kernel for launches of FEW samples (the reference renders one per
//                              launch): warps pull 8x4 tiles from a global counter, idle lanes take new pixels together (REFILL step);
//                              chosen against variant 4 by the renderer's launch auto-tuner (drt_api.cu)
// Warps cover 8x4 pixel tiles (coherent primary rays); blocks are rows of such tiles. A launch may own only the tiles
// first, first+stride, ... (image-tile sharding across GPUs, RenderParams::tile_first / tile_stride).
#include "drt_kernels.h"
#include "drt_shade.cuh"
#include "drt_phased.cuh"
#include "drt_regroup.cuh"

namespace drt {

namespace {

// group of tiles this block renders: its own index, or the entry of the cost order (RenderParams::block_order)
__device__ __forceinline__ uint32_t group_of_block(const RenderParams& P) { return P.block_order ? P.block_order[blockIdx.x] : blockIdx.x; }

__device__ __forceinline__ void pixel_of_thread(int width, int height, int warps_per_block, uint32_t tile_first, uint32_t tile_stride,
	int& x, int& y, bool& valid, uint32_t group = blockIdx.x)
{
	// 8x4 pixel tile per warp; tiles numbered row-major over the image; a launch owns tiles first, first+stride, ...
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int tiles_x = (width + 7) >> 3;
	const long long tile = ((long long)group * warps_per_block + warp) * tile_stride + tile_first;
	const int ty = (int)(tile / tiles_x), tx = (int)(tile - (long long)ty * tiles_x);
	x = tx * 8 + (lane & 7);
	y = ty * 4 + (lane >> 3);
	valid = x < width && y < height;
}

__device__ __forceinline__ void flush_counters(Counters* c, const PathCounters& pc, const TraceStats* st)
{
	unsigned closest = __reduce_add_sync(0xffffffffu, pc.closest);
	unsigned shadow = __reduce_add_sync(0xffffffffu, pc.shadow);
	if ((threadIdx.x & 31) == 0) {
		atomicAdd(&c->closest_rays, (unsigned long long)closest);
		atomicAdd(&c->shadow_rays, (unsigned long long)shadow);
	}
	if (st) {
		unsigned n = __reduce_add_sync(0xffffffffu, st->nodes), t = __reduce_add_sync(0xffffffffu, st->tris);
		unsigned ti = __reduce_add_sync(0xffffffffu, st->ties), a = __reduce_add_sync(0xffffffffu, st->alpha);
		unsigned cl = __reduce_add_sync(0xffffffffu, st->clamps), ms = __reduce_max_sync(0xffffffffu, st->max_stack);
		if ((threadIdx.x & 31) == 0) {
			atomicAdd(&c->node_visits, (unsigned long long)n);
			atomicAdd(&c->tri_tests, (unsigned long long)t);
			atomicAdd(&c->tie_events, (unsigned long long)ti);
			atomicAdd(&c->alpha_tests, (unsigned long long)a);
			atomicAdd(&c->texel_clamps, (unsigned long long)cl);
			atomicMax(&c->max_stack, (unsigned long long)ms);
		}
	}
}

// VARIANT 17 / 18 = variant 4 / 15 with the first DRT_SSTACK entries of the traversal stack in shared memory (SharedLaneStack).
template <bool STRICT, bool STATS, int VARIANT>
__global__ void __launch_bounds__(VARIANT >= 4 ? 128 : 256, VARIANT == 7 ? 8 : ((VARIANT == 8 || VARIANT >= 15) ? 6 : (VARIANT >= 4 ? VARIANT + 2 : 1))) k_render(const __grid_constant__ RenderParams P, const bool rtl, const int warps_per_block,
	const PhasedTuning tune)
{
	int x, y;
	bool valid;
	pixel_of_thread(P.width, P.height, warps_per_block, P.tile_first, P.tile_stride, x, y, valid, VARIANT >= 2 ? group_of_block(P) : blockIdx.x);
	const unsigned t_begin = (VARIANT >= 2 && P.block_cost) ? (unsigned)clock() : 0u; // (32 bits: a block lives far less than 2 s)
	PathCounters pc;
	pc.closest = 0; pc.shadow = 0;
	TraceStats ts;
	ts.nodes = ts.tris = ts.ties = ts.alpha = ts.max_stack = ts.clamps = 0;
	TraceStats* st = STATS ? &ts : nullptr;
	if (VARIANT >= 2) {
		// every lane of the warp takes part in the vote, valid pixel or not
		const size_t p = (size_t)x + (size_t)y * (size_t)P.width;
		V3 acc = mk3(0.f, 0.f, 0.f);
		if (valid && !P.accum_add) { const float4 a4 = P.accum[p]; acc = mk3(a4.x, a4.y, a4.z); }
		__shared__ alignas(16) uint8_t lut[DRT_PERMUTE_LUT_BYTES];
		if (VARIANT == 22 || VARIANT == 23) {
			// sample pool: the last tune.pool samples of the tile's pixels are handed out dynamically (render_warp_phased POOL)
			__shared__ unsigned pool_counter[8];
			__shared__ uint32_t pool_cold[4 * DRT_POOL_COLD_THREADS];
			const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
			if (lane == 0) pool_counter[warp] = 0u;
			__syncwarp();
			SamplePool sp;
			sp.counters = pool_counter;
			sp.cold = pool_cold;
			sp.warps_per_block = warps_per_block;
			acc = render_warp_phased<STRICT, STATS, NODE_FMT_F32, VARIANT == 23, LaneStack, false, true>(P, tune, (uint32_t)x, (uint32_t)y, valid, rtl, acc, pc, st, lut, nullptr, sp);
		}
		else
		if (VARIANT == 21) // variant 4 with the watertight TRI step (DRT_FLAG_WATERTIGHT; production arithmetic only)
			acc = render_warp_phased<false, STATS, NODE_FMT_F32, false, LaneStack, true>(P, tune, (uint32_t)x, (uint32_t)y, valid, rtl, acc, pc, st, lut);
		else
		if (VARIANT == 17 || VARIANT == 18) {
			__shared__ uint32_t stack_words[DRT_SSTACK_WORDS];
			acc = render_warp_phased<STRICT, STATS, NODE_FMT_F32, VARIANT == 18, SharedLaneStack>(P, tune, (uint32_t)x, (uint32_t)y, valid, rtl, acc, pc, st, lut, stack_words);
		}
		else
		acc = render_warp_phased<STRICT, STATS, (VARIANT == 3 || VARIANT == 7 || VARIANT == 8) ? NODE_FMT_Q8 : NODE_FMT_F32, VARIANT == 15>(P, tune, (uint32_t)x, (uint32_t)y, valid, rtl, acc, pc, st, lut);
		if (valid && P.accum_add) {
			// fused render + reduce: this launch's share of the pixel sum goes straight into the (possibly remote) frame buffer
			float* a = &P.accum[p].x;
			atomicAdd(a, acc.x); atomicAdd(a + 1, acc.y); atomicAdd(a + 2, acc.z);
		}
		else if (valid) {
			P.accum[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
			P.display[p] = make_float4(acc.x / P.total_samples, acc.y / P.total_samples, acc.z / P.total_samples, 1.f);
		}
	}
	else if (valid) {
		const size_t p = (size_t)x + (size_t)y * (size_t)P.width;
		const float4 a4 = P.accum[p];
		V3 acc = mk3(a4.x, a4.y, a4.z);
		if (VARIANT == 0 && P.shade.render_mode == DRT_RENDER_NORMAL) {
			acc = render_pixel_flat<STRICT, STATS>(P, (uint32_t)x, (uint32_t)y, rtl, acc, pc, st);
		}
		else {
			for (uint32_t k = 0; k < P.n_samples; k++) {
				const uint32_t frameidx = P.frame_first + k * P.frame_stride;
				const V3 c = path_radiance<STRICT, STATS>(P, (uint32_t)x, (uint32_t)y, frameidx, rtl, pc, st);
				acc = M<STRICT>::add3(acc, c);
			}
		}
		P.accum[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
		// display = acc / frameidx, alpha 1 (RenderKernel.cu:30-31)
		P.display[p] = make_float4(acc.x / P.total_samples, acc.y / P.total_samples, acc.z / P.total_samples, 1.f);
	}
	flush_counters(P.counters, pc, st);
	// cost of this group of tiles for the order of the NEXT launch of the same view: the clocks its slowest warp was resident
	if (VARIANT >= 2 && P.block_cost && (threadIdx.x & 31) == 0) atomicMax(&P.block_cost[group_of_block(P)], (((unsigned)clock() - t_begin) >> 4) + 1u);
}

// One block: block_order = the indices 0 .. n-1 sorted by DESCENDING cost (counting sort over DRT_ORDER_BUCKETS linear buckets of
// cost / max cost; the order inside a bucket is whatever the atomics give -- any order renders the same image).
#define DRT_ORDER_BUCKETS 2048
__global__ void __launch_bounds__(1024) k_block_order(const uint32_t* __restrict__ cost, uint32_t* __restrict__ order, int n)
{
	__shared__ unsigned hist[DRT_ORDER_BUCKETS];
	__shared__ unsigned scan[DRT_ORDER_BUCKETS];
	__shared__ unsigned smax;
	const int tid = (int)threadIdx.x, nt = (int)blockDim.x;
	if (tid == 0) smax = 1u;
	for (int b = tid; b < DRT_ORDER_BUCKETS; b += nt) hist[b] = 0u;
	__syncthreads();
	unsigned m = 1u;
	for (int i = tid; i < n; i += nt) m = max(m, cost[i]);
	m = __reduce_max_sync(0xffffffffu, m);
	if ((tid & 31) == 0) atomicMax(&smax, m);
	__syncthreads();
	const unsigned long long mx = smax;
	for (int i = tid; i < n; i += nt) atomicAdd(&hist[(DRT_ORDER_BUCKETS - 1) - (unsigned)(((unsigned long long)cost[i] * (DRT_ORDER_BUCKETS - 1)) / mx)], 1u);
	__syncthreads();
	// exclusive prefix sum of the bucket counts (Hillis-Steele on 2048 words, two per thread)
	for (int b = tid; b < DRT_ORDER_BUCKETS; b += nt) scan[b] = hist[b];
	__syncthreads();
	for (int d = 1; d < DRT_ORDER_BUCKETS; d <<= 1) {
		unsigned v[DRT_ORDER_BUCKETS / 1024];
		for (int k = 0, b = tid; b < DRT_ORDER_BUCKETS; b += nt, k++) v[k] = b >= d ? scan[b - d] : 0u;
		__syncthreads();
		for (int k = 0, b = tid; b < DRT_ORDER_BUCKETS; b += nt, k++) scan[b] += v[k];
		__syncthreads();
	}
	for (int b = tid; b < DRT_ORDER_BUCKETS; b += nt) hist[b] = scan[b] - hist[b]; // first position of the bucket
	__syncthreads();
	for (int i = tid; i < n; i += nt) {
		const unsigned b = (DRT_ORDER_BUCKETS - 1) - (unsigned)(((unsigned long long)cost[i] * (DRT_ORDER_BUCKETS - 1)) / mx);
		order[atomicAdd(&hist[b], 1u)] = (uint32_t)i;
	}
}

// Phased kernel with the SHADE-only lane state in shared memory (drt_phased.cuh render_warp_cold).
//   CV 0 / 1 / 2 : fixed pixel per lane, expanded fp32 nodes, 8 / 7 / 6 resident blocks per SM   (variants 11 / 12 / 13)
//   CV 5         : fixed pixel per lane, compressed nodes, 8 blocks per SM                        (variant 14)
//   CV 3 / 4 / 6 : persistent pixel-queue kernel, fp32 / compressed nodes / fp32 + two-triangle TRI step, 7 blocks per SM (variants 9 / 10):
//                  grid = resident blocks of the device, warps pull 8x4 tiles from P.work_queue[0]; the last warp to
//                  leave resets the queue words for the next launch.
template <bool STRICT, bool STATS, int CV>
__global__ void __launch_bounds__(128, (CV == 1 || CV == 3 || CV == 4 || CV == 6) ? 7 : (CV == 2 ? 6 : 8)) k_render_cold(const __grid_constant__ RenderParams P, const bool rtl, const PhasedTuning tune)
{
	constexpr bool QUEUE = CV == 3 || CV == 4 || CV == 6;
	constexpr bool TRI2 = CV == 6; // two-triangle TRI step (scenes of DRT_TRI2_MIN_TRIS triangles or more)
	constexpr int FMT = (CV == 4 || CV == 5) ? NODE_FMT_Q8 : NODE_FMT_F32;
	TraceStats ts;
	ts.nodes = ts.tris = ts.ties = ts.alpha = ts.max_stack = ts.clamps = 0;
	TraceStats* st = STATS ? &ts : nullptr;
	__shared__ alignas(16) uint8_t lut[DRT_PERMUTE_LUT_BYTES];
	__shared__ uint32_t cold[DRT_COLD_WORDS];
	int x = 0, y = 0;
	bool valid = false;
	if (!QUEUE) pixel_of_thread(P.width, P.height, 4, P.tile_first, P.tile_stride, x, y, valid);
	render_warp_cold<STRICT, STATS, FMT, QUEUE, TRI2>(P, tune, (uint32_t)x, (uint32_t)y, valid, rtl, st, lut, cold);
	const uint32_t* c = cold + threadIdx.x;
	if (!QUEUE && valid) {
		const size_t p = (size_t)x + (size_t)y * (size_t)P.width;
		const V3 acc = cold_get3(c, CF_ACC);
		P.accum[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
		P.display[p] = make_float4(acc.x / P.total_samples, acc.y / P.total_samples, acc.z / P.total_samples, 1.f);
	}
	PathCounters pc;
	pc.closest = c[CF_CLOSEST * DRT_COLD_THREADS]; pc.shadow = c[CF_SHADOW * DRT_COLD_THREADS];
	flush_counters(P.counters, pc, st);
	if (QUEUE && (threadIdx.x & 31) == 0) {
		__threadfence();
		const unsigned warps = gridDim.x * (blockDim.x >> 5);
		if (atomicAdd(P.work_queue + 1, 1u) == warps - 1u) {
			P.work_queue[0] = 0u;
			P.work_queue[1] = 0u;
		}
	}
}

// Variant 16: block-level regrouping of path contexts (drt_regroup.cuh). Persistent grid of DRT_RG_BLOCKS_PER_SM blocks per SM,
// DRT_RG_CTX contexts per block in (dynamic) shared memory; pixels come from P.work_queue[0], which the last block to leave resets.
template <bool STRICT, int FMT, bool TRI2>
__global__ void __launch_bounds__(DRT_RG_THREADS, DRT_RG_BLOCKS_PER_SM) k_render_regroup(const __grid_constant__ RenderParams P, const bool rtl)
{
	extern __shared__ uint4 rg_smem[];
	__shared__ unsigned tail[16];
	__shared__ alignas(16) uint8_t lut[DRT_PERMUTE_LUT_BYTES];
	PathCounters pc;
	pc.closest = 0; pc.shadow = 0;
	uint2* ovf = P.rg_overflow ? (uint2*)P.rg_overflow + (size_t)blockIdx.x * DRT_RG_CTX * DRT_PSTACK : nullptr;
	render_block_regroup<STRICT, FMT, TRI2>(P, rtl, ovf, (uint32_t*)rg_smem, tail, lut, pc);
	flush_counters(P.counters, pc, nullptr);
	if (threadIdx.x == 0) {
		__threadfence();
		if (atomicAdd(P.work_queue + 1, 1u) == gridDim.x - 1u) {
			P.work_queue[0] = 0u;
			P.work_queue[1] = 0u;
		}
	}
}

template <bool STRICT, int FMT, bool TRI2>
cudaError_t launch_regroup_k(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s)
{
	static bool configured = false;
	if (!configured) {
		const cudaError_t e = cudaFuncSetAttribute(k_render_regroup<STRICT, FMT, TRI2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DRT_RG_SMEM_BYTES);
		if (e != cudaSuccess) return e;
		configured = true;
	}
	const long long all_tiles = (long long)((P.width + 7) / 8) * ((P.height + 3) / 4);
	const long long tiles = all_tiles > P.tile_first ? (all_tiles - P.tile_first + P.tile_stride - 1) / P.tile_stride : 0;
	long long blocks = (long long)cfg.resident_sms * DRT_RG_BLOCKS_PER_SM;
	const long long need = (tiles * 32 + DRT_RG_CTX - 1) / DRT_RG_CTX;
	if (blocks > need) blocks = need;
	if (blocks <= 0) return cudaSuccess;
	k_render_regroup<STRICT, FMT, TRI2><<<(unsigned)blocks, DRT_RG_THREADS, DRT_RG_SMEM_BYTES, s>>>(P, cfg.rtl);
	return cudaGetLastError();
}

template <bool STRICT>
cudaError_t launch_regroup(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s)
{
	const bool q8 = P.scene.nodes_f32 == nullptr;
	const bool tri2 = !q8 && P.scene.num_tris >= DRT_TRI2_MIN_TRIS;
	if (q8) return launch_regroup_k<STRICT, NODE_FMT_Q8, false>(P, cfg, s);
	if (tri2) return launch_regroup_k<STRICT, NODE_FMT_F32, true>(P, cfg, s);
	return launch_regroup_k<STRICT, NODE_FMT_F32, false>(P, cfg, s);
}

template <bool STRICT>
__global__ void __launch_bounds__(256) k_trace_primary(const __grid_constant__ RenderParams P, const uint32_t frameidx, Hit* out,
	const bool rtl, const int warps_per_block)
{
	int x, y;
	bool valid;
	pixel_of_thread(P.width, P.height, warps_per_block, 0u, 1u, x, y, valid);
	TraceStats ts;
	ts.nodes = ts.tris = ts.ties = ts.alpha = ts.max_stack = ts.clamps = 0;
	PathCounters pc;
	pc.closest = valid ? 1u : 0u; pc.shadow = 0;
	if (valid) {
		uint32_t seed = (uint32_t)x + (uint32_t)y * (uint32_t)P.width;
		seed *= frameidx;
		const RayOD r = camera_get_ray<STRICT>(P.cam, (uint32_t)x, (uint32_t)y, (uint32_t)P.width, (uint32_t)P.height, seed, rtl);
		Hit hit;
		trace_ray<STRICT, false, true>(P.scene, r.o, r.d, FLT_MAX, hit, &ts);
		Hit o;
		o.prim = hit.prim >= 0 ? (int32_t)as_uint(ld_f4(P.scene.tris + 3 * (size_t)hit.prim).w) : -1;
		o.t = hit.prim >= 0 ? hit.t : -1.f;
		o.u = hit.u; o.v = hit.v;
		out[(size_t)x + (size_t)y * (size_t)P.width] = o;
	}
	flush_counters(P.counters, pc, &ts);
}

// Ray-level parity probe (drt_trace_rays): caller-given rays through the SAME resumable steps the phased render kernels run
// (node_step_f32 / node_step_q8 + tri_step / tri_step2 + trav_advance, drt_phased.cuh), one ray per thread, serial schedule.
// anyhit = RayTest semantics (TraceRay.cu:35-38: first accepted hit ends the ray), else TraceRay closest hit.
template <bool STRICT, int FMT, bool TRI2, bool WT = false>
__global__ void __launch_bounds__(128) k_trace_rays(const SceneView sc, const float* __restrict__ org, const float* __restrict__ dir, int n, bool anyhit, Hit* out)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const V3 o = mk3(org[3 * (size_t)i], org[3 * (size_t)i + 1], org[3 * (size_t)i + 2]);
	const V3 d = mk3(dir[3 * (size_t)i], dir[3 * (size_t)i + 1], dir[3 * (size_t)i + 2]);
	LaneTrav L;
	LaneStack K;
	RayBox R;
	ray_box_setup(o, d, R);
	trav_begin(L, FLT_MAX);
	int phase = sc.num_nodes > 0 ? PH_NODE : PH_SHADE;
	bool occluded = false;
	while (phase != PH_SHADE) {
		if (phase == PH_NODE) {
			if (FMT == NODE_FMT_F32) node_step_f32<false>(sc.nodes_f32, L, K, R, nullptr);
			else node_step_q8<false>(sc.nodes, L, K, R, nullptr);
			phase = trav_advance(L, K);
		}
		else {
			if (WT ? tri_step_wt<false>(sc, L, o, d, anyhit, nullptr)
			       : (TRI2 ? tri_step2<STRICT, false>(sc, sc.tris, L, o, d, anyhit, nullptr) : tri_step<STRICT, false>(sc, sc.tris, L, o, d, anyhit, nullptr))) {
				occluded = true;
				phase = PH_SHADE;
			}
			else phase = trav_advance(L, K);
		}
	}
	(void)occluded;
	Hit h;
	h.prim = -1; h.t = -1.f; h.u = 0.f; h.v = 0.f;
	if (L.hit.prim >= 0) {
		const V2 bc = hit_barycentrics<STRICT>(sc.tris, L.hit.prim, o, d); // the operations of tri_step: the same bits
		h.prim = (int32_t)as_uint(ld_f4(sc.tris + 3 * (size_t)L.hit.prim).w);
		h.t = L.hit.t; h.u = bc.x; h.v = bc.y;
	}
	out[i] = h;
}

__global__ void k_resolve(float4* display, const float4* accum, int n, float total)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 a = accum[i];
	display[i] = make_float4(a.x / total, a.y / total, a.z / total, 1.f);
}

__global__ void k_pack_rgba8(uint32_t* out, const float4* display, int n)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 c = display[i];
	// what glGetTexImage(GL_RGBA, GL_UNSIGNED_BYTE) yields from the RGBA32F target (EditorLayer.cpp:85-97):
	// clamp to [0,1], scale by 255, round to nearest
	uint32_t r = (uint32_t)__float2int_rn(__saturatef(c.x) * 255.f), g = (uint32_t)__float2int_rn(__saturatef(c.y) * 255.f);
	uint32_t b = (uint32_t)__float2int_rn(__saturatef(c.z) * 255.f), a = (uint32_t)__float2int_rn(__saturatef(c.w) * 255.f);
	out[i] = r | (g << 8) | (b << 16) | (a << 24);
}

__global__ void k_pack_rgb32f(float* out, const float4* accum, int n)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 a = accum[i];
	out[3 * (size_t)i] = a.x; out[3 * (size_t)i + 1] = a.y; out[3 * (size_t)i + 2] = a.z;
}

// Vertex records of the leaf-order triangles for the watertight test: the builder's records hold v0 and two EDGES; a vertex
// rebuilt as v0 + e would not be the same float in every triangle that shares it, so the original positions are fetched again.
__global__ void k_vertex_records(const drt_f4* tris, const float* pos, drt_f4* out, int n)
{
	const int slot = blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= n) return;
	const drt_f4 r0 = tris[3 * (size_t)slot], r1 = tris[3 * (size_t)slot + 1];
	const float* p = pos + 9 * (size_t)as_uint(r0.w);
	drt_f4 a, b, c;
	a.x = p[0]; a.y = p[1]; a.z = p[2]; a.w = r0.w;
	b.x = p[3]; b.y = p[4]; b.z = p[5]; b.w = r1.w;
	c.x = p[6]; c.y = p[7]; c.z = p[8]; c.w = 0.f;
	out[3 * (size_t)slot] = a; out[3 * (size_t)slot + 1] = b; out[3 * (size_t)slot + 2] = c;
}

__global__ void k_expand_nodes(const WideNode* wide, drt_f4* out, int n)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	expand_wide_node(wide[i], out + (size_t)DRT_F32NODE_F4 * i);
}

// DRT_FLAG_EXTENDED_MATERIALS: the reference-shaped sample loop with the extended shading branch (drt_shade.cuh).
template <bool STRICT, bool STATS>
__global__ void __launch_bounds__(256) k_render_ext(const __grid_constant__ RenderParams P, const bool rtl, const int warps_per_block)
{
	int x, y;
	bool valid;
	pixel_of_thread(P.width, P.height, warps_per_block, P.tile_first, P.tile_stride, x, y, valid);
	PathCounters pc;
	pc.closest = 0; pc.shadow = 0;
	TraceStats ts;
	ts.nodes = ts.tris = ts.ties = ts.alpha = ts.max_stack = ts.clamps = 0;
	TraceStats* st = STATS ? &ts : nullptr;
	if (valid) {
		const size_t p = (size_t)x + (size_t)y * (size_t)P.width;
		const float4 a4 = P.accum[p];
		V3 acc = mk3(a4.x, a4.y, a4.z);
		for (uint32_t k = 0; k < P.n_samples; k++) {
			const uint32_t frameidx = P.frame_first + k * P.frame_stride;
			acc = M<STRICT>::add3(acc, path_radiance<STRICT, STATS, true>(P, (uint32_t)x, (uint32_t)y, frameidx, rtl, pc, st));
		}
		P.accum[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
		P.display[p] = make_float4(acc.x / P.total_samples, acc.y / P.total_samples, acc.z / P.total_samples, 1.f);
	}
	flush_counters(P.counters, pc, st);
}

// ---- builder passes: one thread per element, bodies in drt_build.cuh ----
// The two per-primitive passes of a level (BIN, PARTITION) communicate through atomics on per-node words. Near the root
// millions of primitives hit the same few words (soup-10M: 44 ms per pass per level). The index arrays keep the
// primitives of a node contiguous, so a block's 256 primitives almost always belong to ONE node: the block aggregates
// for that node in shared memory and touches the global words once per block; primitives of other nodes in the same
// block take the plain per-primitive path of drt_build.cuh.
#define DRT_BINS_PER_NODE (3 * DRT_BINS * DRT_BIN_WORDS)

__device__ __forceinline__ void bin_prim_block(const BuildArrays& A, int i)
{
	__shared__ int sb[DRT_BINS_PER_NODE];
	__shared__ int s_node;
	for (int j = threadIdx.x; j < DRT_BINS_PER_NODE; j += blockDim.x) {
		const int k = j % DRT_BIN_WORDS;
		sb[j] = k == 0 ? 0 : (k <= 3 ? DRT_ENC_POS_INF : DRT_ENC_NEG_INF);
	}
	if (threadIdx.x == 0) {
		const int i0 = blockIdx.x * blockDim.x;
		int node = i0 < A.n ? A.pnode_in[i0] : -1;
		if (node < A.level_start || node >= A.level_end || A.nodes[node].count <= DRT_SMALL_NODE) node = -1;
		s_node = node;
	}
	__syncthreads();
	const int bn = s_node;
	if (i < A.n) {
		const int node = A.pnode_in[i];
		if (node != bn || bn < 0) bin_prim(A, i);
		else {
			const BuildNode& nd = A.nodes[node];
			const int prim = A.idx_in[i];
			const drt_f4 lo = A.tri_lo[prim], hi = A.tri_hi[prim];
			const float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
			const int elo[3] = { enc_f(lo.x), enc_f(lo.y), enc_f(lo.z) }, ehi[3] = { enc_f(hi.x), enc_f(hi.y), enc_f(hi.z) };
			for (int a = 0; a < 3; a++) {
				const int bin = bin_of(c[a], dec_f(nd.clo[a]), dec_f(nd.chi[a]), DRT_BINS);
				int* w = sb + (a * DRT_BINS + bin) * DRT_BIN_WORDS;
				atomicAdd(w, 1);
				atomicMin(w + 1, elo[0]); atomicMin(w + 2, elo[1]); atomicMin(w + 3, elo[2]);
				atomicMax(w + 4, ehi[0]); atomicMax(w + 5, ehi[1]); atomicMax(w + 6, ehi[2]);
			}
		}
	}
	__syncthreads();
	if (bn >= 0) {
		int* nb = A.bins + (size_t)(A.splits[bn].bin_slot - A.bin_slot_base) * DRT_BINS_PER_NODE;
		for (int j = threadIdx.x; j < DRT_BINS_PER_NODE; j += blockDim.x) {
			const int k = j % DRT_BIN_WORDS;
			if (sb[j - k] == 0) continue; // empty bin
			if (k == 0) atomicAdd(nb + j, sb[j]);
			else if (k <= 3) atomicMin(nb + j, sb[j]);
			else atomicMax(nb + j, sb[j]);
		}
	}
}

__device__ __forceinline__ void partition_prim_block(const BuildArrays& A, int i)
{
	__shared__ int s_node, s_base[2];
	__shared__ int s_warp[2][8];  // per-warp counts of left / right movers (blocks have 256 threads)
	__shared__ int s_box[2][12];  // per child: centroid lo xyz, centroid hi xyz, box lo xyz, box hi xyz
	if (threadIdx.x < 24) { const int k = threadIdx.x % 12; (&s_box[0][0])[threadIdx.x] = (k % 6) < 3 ? DRT_ENC_POS_INF : DRT_ENC_NEG_INF; }
	if (threadIdx.x == 0) {
		const int i0 = blockIdx.x * blockDim.x;
		int node = i0 < A.n ? A.pnode_in[i0] : -1;
		if (node < A.level_start || node >= A.level_end || A.nodes[node].left < 0 || A.nodes[node].count <= DRT_SMALL_NODE) node = -1;
		s_node = node;
	}
	__syncthreads();
	const int bn = s_node;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	bool mine = false, left = false;
	int prim = 0;
	float c[3] = { 0.f, 0.f, 0.f };
	drt_f4 lo = { 0.f, 0.f, 0.f, 0.f }, hi = { 0.f, 0.f, 0.f, 0.f };
	bool median = false;
	if (i < A.n) {
		const int node = A.pnode_in[i];
		if (node != bn || bn < 0) partition_prim(A, i);
		else {
			mine = true;
			const BuildNode& nd = A.nodes[node];
			const BuildSplit sp = A.splits[node];
			prim = A.idx_in[i];
			lo = A.tri_lo[prim]; hi = A.tri_hi[prim];
			c[0] = 0.5f * (lo.x + hi.x); c[1] = 0.5f * (lo.y + hi.y); c[2] = 0.5f * (lo.z + hi.z);
			median = sp.axis < 0;
			left = !median ? (bin_of(c[sp.axis], dec_f(nd.clo[sp.axis]), dec_f(nd.chi[sp.axis]), DRT_BINS) < sp.split_bin)
				: (i - nd.start < sp.n_left);
		}
	}
	const unsigned ml = __ballot_sync(0xffffffffu, mine && left), mr = __ballot_sync(0xffffffffu, mine && !left);
	if (lane == 0) { s_warp[0][warp] = __popc(ml); s_warp[1][warp] = __popc(mr); }
	if (mine) {
		int* bx = s_box[left ? 0 : 1];
		atomicMin(bx + 0, enc_f(c[0])); atomicMin(bx + 1, enc_f(c[1])); atomicMin(bx + 2, enc_f(c[2]));
		atomicMax(bx + 3, enc_f(c[0])); atomicMax(bx + 4, enc_f(c[1])); atomicMax(bx + 5, enc_f(c[2]));
		if (median) {
			atomicMin(bx + 6, enc_f(lo.x)); atomicMin(bx + 7, enc_f(lo.y)); atomicMin(bx + 8, enc_f(lo.z));
			atomicMax(bx + 9, enc_f(hi.x)); atomicMax(bx + 10, enc_f(hi.y)); atomicMax(bx + 11, enc_f(hi.z));
		}
	}
	__syncthreads();
	if (bn < 0) return;
	BuildNode* kids = A.nodes + A.nodes[bn].left;
	if (threadIdx.x < 2) { // one cursor bump per child per block
		int total = 0;
		for (int w = 0; w < 8; w++) total += s_warp[threadIdx.x][w];
		s_base[threadIdx.x] = total ? atomicAdd(&kids[threadIdx.x].fill, total) : 0;
	}
	else if (threadIdx.x >= 32 && threadIdx.x < 56) { // child bounds, once per block
		const int t = threadIdx.x - 32, side = t / 12, k = t % 12;
		const int v = s_box[side][k];
		BuildNode* cn = kids + side;
		if (k < 3) { if (v != DRT_ENC_POS_INF) atomicMin(&cn->clo[k], v); }
		else if (k < 6) { if (v != DRT_ENC_NEG_INF) atomicMax(&cn->chi[k - 3], v); }
		else if (k < 9) { if (v != DRT_ENC_POS_INF) atomicMin(&cn->lo[k - 6], v); }
		else { if (v != DRT_ENC_NEG_INF) atomicMax(&cn->hi[k - 9], v); }
	}
	__syncthreads();
	if (mine) {
		const int side = left ? 0 : 1;
		int rank = __popc((left ? ml : mr) & ((1u << lane) - 1u));
		for (int w = 0; w < warp; w++) rank += s_warp[side][w];
		const int np = kids[side].start + s_base[side] + rank;
		A.idx_out[np] = prim;
		A.pnode_out[np] = A.nodes[bn].left + side;
	}
}

template <int PASS>
__global__ void __launch_bounds__(256) k_build(const BuildArrays A, const int* idx_final, int a, int b, int n)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (PASS == PASS_BIN) { bin_prim_block(A, i); return; }             // (block-wide barriers inside: no early exit)
	if (PASS == PASS_PARTITION) { partition_prim_block(A, i); return; }
	if (i >= n) return;
	if (PASS == PASS_PRIM_BOUNDS) prim_bounds(A, i);
	else if (PASS == PASS_INIT_BINS) init_bin_word(A, i, a);
	else if (PASS == PASS_BIN) bin_prim(A, i);
	else if (PASS == PASS_SPLIT) split_node(A, i);
	else if (PASS == PASS_PARTITION) partition_prim(A, i);
	else if (PASS == PASS_COLLAPSE) collapse_node(A, idx_final, a, b, i);
}

template <bool STRICT, bool STATS>
cudaError_t launch_render_v(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s, int* variant_out)
{
#define DRT_RAN(V) do { if (variant_out) *variant_out = (V); } while (0)
	const int threads = cfg.block_threads > 0 ? cfg.block_threads : 128;
	const int wpb = threads / 32;
	const long long all_tiles = (long long)((P.width + 7) / 8) * ((P.height + 3) / 4);
	if (P.tile_stride == 0) return cudaErrorInvalidValue;
	const long long tiles = all_tiles > P.tile_first ? (all_tiles - P.tile_first + P.tile_stride - 1) / P.tile_stride : 0; // tiles this launch owns
	if (tiles == 0) return cudaSuccess;
	const unsigned blocks = (unsigned)((tiles + wpb - 1) / wpb);
	PhasedTuning tune;
	tune.stick = cfg.sched_bias[0] > 0 ? cfg.sched_bias[0] : 6; // measured: room best at 8, dense_monkey at 4 (33 = never)
	tune.refill = cfg.sched_bias[1] > 0 ? cfg.sched_bias[1] : (P.n_samples <= 1 ? 12 : (P.n_samples <= 4 ? 6 : 1));
	tune.pool = 0;
	// the phased kernels implement NORMALMODE; the debug views go through the reference-shaped loop
	int variant = cfg.variant;
	// fused render + reduce (RenderParams::accum_add) exists in the fixed-tile phased kernels only
#define DRT_NO_ACCUM_ADD do { if (P.accum_add) return cudaErrorNotSupported; } while (0)
	if (cfg.ext_materials && P.shade.render_mode == DRT_RENDER_NORMAL) {
		DRT_NO_ACCUM_ADD;
		k_render_ext<STRICT, STATS><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb);
		DRT_RAN(-2);
		return cudaGetLastError();
	}
	if (variant >= 2 && P.shade.render_mode != DRT_RENDER_NORMAL) variant = 1;
	// DRT_FLAG_WATERTIGHT: the watertight TRI step replaces the reference's Moller-Trumbore; not in parity (STRICT) mode, whose
	// whole point is the reference's predicate, and only where the scene has vertex records and expanded nodes
	if (cfg.watertight && !STRICT && variant >= 2 && P.scene.tris_v != nullptr && P.scene.nodes_f32 != nullptr && threads == 128) {
		k_render<false, STATS, 21><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
		DRT_RAN(21);
		return cudaGetLastError();
	}
	// variant 16: block-level regrouping (no statistics build of it: counted launches run the per-warp voting kernel)
	if (variant == 16) {
		if (!STATS && !P.accum_add && P.work_queue != nullptr && cfg.resident_sms > 0 && P.tile_stride > 0 && P.width < 65536 && P.height < 65536) { DRT_RAN(16); return launch_regroup<STRICT>(P, cfg, s); }
		variant = 4;
	}
	// variants 9 / 10: persistent warps with a pixel queue, for launches of few samples (the reference's interactive pattern is
	// one per launch); the renderer's auto-tuner (drt_api.cu) picks between them and variant 4 by measurement.
	if (variant >= 9 && variant <= 14) {
		const bool queue = variant <= 10;
		const bool ok = threads == 128 && !P.accum_add && P.width < 65536 && P.height < 65536 && (!queue || (P.work_queue != nullptr && cfg.resident_sms > 0));
		const bool q8 = variant == 10 || variant == 14 || P.scene.nodes_f32 == nullptr;
		if (!ok) variant = q8 ? 7 : 4;
		else if (queue) {
			// persistent grid = what is resident at once (asked from the runtime: a block that had to wait would find the queue empty)
			const bool tri2 = !q8 && cfg.auto_variant && P.scene.num_tris >= DRT_TRI2_MIN_TRIS;
			static int per_sm[2] = { 0, 0 };
			if (per_sm[q8 ? 1 : 0] == 0) {
				int nb = 0;
				const cudaError_t oe = q8 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_render_cold<STRICT, STATS, 4>, threads, 0)
				                          : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_render_cold<STRICT, STATS, 3>, threads, 0);
				if (oe != cudaSuccess) return oe;
				per_sm[q8 ? 1 : 0] = nb > 0 ? nb : 1;
			}
			long long qblocks = (long long)cfg.resident_sms * per_sm[q8 ? 1 : 0];
			if (qblocks * wpb > tiles) qblocks = (tiles + wpb - 1) / wpb; // at least one tile per warp
			if (q8) k_render_cold<STRICT, STATS, 4><<<(unsigned)qblocks, threads, 0, s>>>(P, cfg.rtl, tune);
			else if (tri2) k_render_cold<STRICT, STATS, 6><<<(unsigned)qblocks, threads, 0, s>>>(P, cfg.rtl, tune);
			else k_render_cold<STRICT, STATS, 3><<<(unsigned)qblocks, threads, 0, s>>>(P, cfg.rtl, tune);
			DRT_RAN(q8 ? 10 : 9);
			return cudaGetLastError();
		}
		else {
			if (q8) k_render_cold<STRICT, STATS, 5><<<blocks, threads, 0, s>>>(P, cfg.rtl, tune);
			else if (variant == 12) k_render_cold<STRICT, STATS, 1><<<blocks, threads, 0, s>>>(P, cfg.rtl, tune);
			else if (variant == 13) k_render_cold<STRICT, STATS, 2><<<blocks, threads, 0, s>>>(P, cfg.rtl, tune);
			else k_render_cold<STRICT, STATS, 0><<<blocks, threads, 0, s>>>(P, cfg.rtl, tune);
			DRT_RAN(variant);
			return cudaGetLastError();
		}
	}
	if (cfg.auto_variant && variant == 4 && P.scene.num_tris >= DRT_TRI2_MIN_TRIS) variant = 15;
	// automatic selection: launches of DRT_POOL_MIN_SPP samples or more on scenes of DRT_POOL_MIN_TRIS triangles or more run the
	// sample-pool form of the kernel (variants 22 / 23) when the renderer has provided the scratch array
	if (cfg.auto_variant && (variant == 4 || variant == 15) && P.n_samples >= DRT_POOL_MIN_SPP && P.scene.num_tris >= DRT_POOL_MIN_TRIS &&
		P.pool_scratch != nullptr) variant = (variant == 15 || P.scene.num_tris >= DRT_POOL_TRI2_MIN_TRIS) ? 23 : 22; // (with the pool the two-triangle step pays from 256 triangles: room.glb +1.1 %)
	if ((variant == 17 || variant == 18) && (P.scene.nodes_f32 == nullptr || threads != 128)) variant = variant == 17 ? 4 : 15; // (shared-memory stack: fp32 nodes, 128 threads)
	if (variant == 22 || variant == 23) {
		// sample pool: needs the scratch array (one float4 per pool sample of every owned tile), fp32 nodes, 128 threads
		tune.pool = sample_pool_size(P.n_samples, cfg.sched_bias[2]);
		const bool ok = tune.pool > 0 && P.scene.nodes_f32 != nullptr && threads == 128 && P.pool_scratch != nullptr && P.width < 65536 && P.height < 65536 &&
			sample_pool_scratch_bytes(tiles, tune.pool) <= cfg.pool_scratch_bytes;
		if (!ok) { variant = variant == 22 ? 4 : 15; tune.pool = 0; }
		else {
			// sticky re-issue threshold: 8 with the pool (re-swept on the pool kernels, profiles/r02f_stick_sweep.jsonl: room 256 spp
			// 125.8 -> 124.6 ms, animemaid +1.5 %, dense_monkey +1.9 %, valorant_ghost -0.3 % against 6)
			if (cfg.sched_bias[0] <= 0) tune.stick = 8;
			if (variant == 22) k_render<STRICT, STATS, 22><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
			else k_render<STRICT, STATS, 23><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
			DRT_RAN(variant);
			return cudaGetLastError();
		}
	}
	if (variant == 15 && (P.scene.nodes_f32 == nullptr || threads > 128)) variant = 4; // (two-triangle TRI step: fp32 nodes, 128 threads)
	if ((variant == 2 || (variant >= 4 && variant <= 6)) && P.scene.nodes_f32 == nullptr) variant = threads <= 128 ? 7 : 3;
	if (variant >= 7 && threads > 128) variant = 3;
	if (variant >= 4 && variant <= 6 && threads > 128) variant = 2;
	DRT_RAN(variant);
	if (variant == 17) k_render<STRICT, STATS, 17><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 18) k_render<STRICT, STATS, 18><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else
	if (variant == 7) k_render<STRICT, STATS, 7><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 8) k_render<STRICT, STATS, 8><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 15) k_render<STRICT, STATS, 15><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else
	if (variant == 4) k_render<STRICT, STATS, 4><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 5) k_render<STRICT, STATS, 5><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 6) k_render<STRICT, STATS, 6><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 3) k_render<STRICT, STATS, 3><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 2) k_render<STRICT, STATS, 2><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune);
	else if (variant == 1) { DRT_NO_ACCUM_ADD; k_render<STRICT, STATS, 1><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune); }
	else { DRT_NO_ACCUM_ADD; k_render<STRICT, STATS, 0><<<blocks, threads, 0, s>>>(P, cfg.rtl, wpb, tune); }
	return cudaGetLastError();
#undef DRT_NO_ACCUM_ADD
#undef DRT_RAN
}

} // namespace

// R = pool samples per pixel: the pool has to cover the spread between a tile's fastest and slowest lane, ~1.1 sqrt(n) samples
// (fit of the fixed-tile kernel's t(spp)); measured on room.glb (profiles/r02c_sample_pool.txt): 32 spp saturates at R = 8-9,
// 256 spp at R = 24-32. At least half of the samples stay with their own pixel. override_r > 0 forces R (sweeps).
int sample_pool_size(uint32_t n_samples, int override_r)
{
	if (n_samples < 2u) return 0;
	int r = override_r > 0 ? override_r : (int)ceilf(1.5f * sqrtf((float)n_samples)) + 1;
	if (r > 32 && override_r <= 0) r = 32;
	if ((uint32_t)r > n_samples / 2u && override_r <= 0) r = (int)(n_samples / 2u);
	if ((uint32_t)r > n_samples) r = (int)n_samples;
	return r;
}
size_t sample_pool_scratch_bytes(long long tiles, int pool_r) { return (size_t)(tiles > 0 ? tiles : 0) * 32u * (size_t)(pool_r > 0 ? pool_r : 0) * sizeof(float4); }

size_t regroup_overflow_bytes(int sms) { return (size_t)sms * DRT_RG_BLOCKS_PER_SM * DRT_RG_CTX * DRT_PSTACK * sizeof(uint2); }

cudaError_t launch_render(const RenderParams& P, const LaunchConfig& cfg, cudaStream_t s, int* launches, int* variant_out)
{
	if (launches) *launches += 1;
	if (cfg.block_threads % 32 != 0 || cfg.block_threads > 256) return cudaErrorInvalidValue;
	if (cfg.strict) return cfg.stats ? launch_render_v<true, true>(P, cfg, s, variant_out) : launch_render_v<true, false>(P, cfg, s, variant_out);
	return cfg.stats ? launch_render_v<false, true>(P, cfg, s, variant_out) : launch_render_v<false, false>(P, cfg, s, variant_out);
}

cudaError_t launch_trace_primary(const RenderParams& P, uint32_t frame_index, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s)
{
	const int threads = 128, wpb = threads / 32;
	const long long tiles = (long long)((P.width + 7) / 8) * ((P.height + 3) / 4);
	const unsigned blocks = (unsigned)((tiles + wpb - 1) / wpb);
	if (cfg.strict) k_trace_primary<true><<<blocks, threads, 0, s>>>(P, frame_index, out_dev, cfg.rtl, wpb);
	else k_trace_primary<false><<<blocks, threads, 0, s>>>(P, frame_index, out_dev, cfg.rtl, wpb);
	return cudaGetLastError();
}

cudaError_t launch_trace_rays(const SceneView& sc, const float* org_dev, const float* dir_dev, int n, bool anyhit, Hit* out_dev, const LaunchConfig& cfg, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	const unsigned blocks = (unsigned)((n + 127) / 128);
	const bool f32 = sc.nodes_f32 != nullptr && cfg.variant != 3 && cfg.variant != 7 && cfg.variant != 8 && cfg.variant != 10 && cfg.variant != 14;
	const bool tri2 = cfg.variant == 15 || (cfg.auto_variant && sc.num_tris >= DRT_TRI2_MIN_TRIS);
	if (cfg.watertight && !cfg.strict && f32 && sc.tris_v != nullptr) {
		k_trace_rays<false, NODE_FMT_F32, false, true><<<blocks, 128, 0, s>>>(sc, org_dev, dir_dev, n, anyhit, out_dev);
		return cudaGetLastError();
	}
#define DRT_TR(S, F, T2) k_trace_rays<S, F, T2><<<blocks, 128, 0, s>>>(sc, org_dev, dir_dev, n, anyhit, out_dev)
	if (cfg.strict) { if (f32) { if (tri2) DRT_TR(true, NODE_FMT_F32, true); else DRT_TR(true, NODE_FMT_F32, false); } else { if (tri2) DRT_TR(true, NODE_FMT_Q8, true); else DRT_TR(true, NODE_FMT_Q8, false); } }
	else { if (f32) { if (tri2) DRT_TR(false, NODE_FMT_F32, true); else DRT_TR(false, NODE_FMT_F32, false); } else { if (tri2) DRT_TR(false, NODE_FMT_Q8, true); else DRT_TR(false, NODE_FMT_Q8, false); } }
#undef DRT_TR
	return cudaGetLastError();
}

cudaError_t launch_block_order(const uint32_t* cost, uint32_t* order, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_block_order<<<1, 1024, 0, s>>>(cost, order, n);
	return cudaGetLastError();
}

cudaError_t launch_resolve(drt_f4* display, const drt_f4* accum, int n, float total, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_resolve<<<(n + 255) / 256, 256, 0, s>>>(display, accum, n, total);
	return cudaGetLastError();
}

cudaError_t launch_pack_rgba8(uint32_t* out, const drt_f4* display, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_pack_rgba8<<<(n + 255) / 256, 256, 0, s>>>(out, display, n);
	return cudaGetLastError();
}

cudaError_t launch_pack_rgb32f(float* out, const drt_f4* accum, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_pack_rgb32f<<<(n + 255) / 256, 256, 0, s>>>(out, accum, n);
	return cudaGetLastError();
}

cudaError_t launch_vertex_records(const drt_f4* tris, const float* pos, drt_f4* out, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_vertex_records<<<(n + 255) / 256, 256, 0, s>>>(tris, pos, out, n);
	return cudaGetLastError();
}

cudaError_t launch_expand_nodes(const WideNode* wide, drt_f4* out, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	k_expand_nodes<<<(n + 127) / 128, 128, 0, s>>>(wide, out, n);
	return cudaGetLastError();
}

// One WARP splits one small node (count <= DRT_SMALL_NODE = 32 primitives), in registers: lane k holds primitive k. Per axis the
// lanes add their boxes to the bins in shared memory (atomics on the order-preserving integer encoding of the bounds), lanes
// 0 .. 15 pick one bin each, a prefix and a suffix scan over those lanes (shuffles) give every candidate
// plane's two sides at once, lane s evaluates plane s, and one redux picks the cheapest. It takes the decisions of split_node's
// small path (drt_build.cuh) -- a union of boxes does not depend on the order, the costs are the same separately rounded expression
// on the same values, ties go to the first axis / first plane -- where one thread needs ~35 k clocks because its bin arrays live in
// local memory. `bins` = DRT_SMALL_BINS x 7 words of shared memory owned by this warp.
#ifdef DRT_BUILD_SMALL_TIMING
__device__ int* g_bs_debug = nullptr; // result + 6: max clocks of the SAH part / the partition part of a warp split
#endif
struct WarpBin { int n; float lo[3], hi[3]; };
__device__ __forceinline__ void warp_bin_merge(WarpBin& d, const WarpBin& o)
{
	d.n += o.n;
#pragma unroll
	for (int k = 0; k < 3; k++) { d.lo[k] = fminf(d.lo[k], o.lo[k]); d.hi[k] = fmaxf(d.hi[k], o.hi[k]); }
}
__device__ __forceinline__ WarpBin warp_bin_shfl(const WarpBin& v, int src_lane)
{
	WarpBin o;
	o.n = __shfl_sync(0xffffffffu, v.n, src_lane);
#pragma unroll
	for (int k = 0; k < 3; k++) { o.lo[k] = __shfl_sync(0xffffffffu, v.lo[k], src_lane); o.hi[k] = __shfl_sync(0xffffffffu, v.hi[k], src_lane); }
	return o;
}
// Lane b < NB holds bin b of one axis (the other lanes an empty bin). Inclusive prefix and suffix unions over the bins (shuffles), lane s
// evaluates plane s (bins 0 .. s-1 | bins s .. NB-1), one redux finds the cheapest plane; kmin = its cost bits (0xffffffff: every plane leaves a
// side empty), split = the first plane among the cheapest, bl / br = its two sides (on every lane).
template <int NB>
__device__ __forceinline__ void warp_plane_search(const WarpBin& v, int lane, unsigned& kmin, int& split, WarpBin& bl, WarpBin& br)
{
	const unsigned FULL = 0xffffffffu;
	WarpBin pre = v, suf = v;
#pragma unroll
	for (int d = 1; d < NB; d <<= 1) {
		const WarpBin up = warp_bin_shfl(pre, lane - d >= 0 ? lane - d : lane), dn = warp_bin_shfl(suf, lane + d < 32 ? lane + d : lane);
		if (lane >= d) warp_bin_merge(pre, up);
		if (lane + d < NB) warp_bin_merge(suf, dn);
	}
	const WarpBin L = warp_bin_shfl(pre, lane >= 1 ? lane - 1 : 0);
	unsigned key = 0xffffffffu; // cost bits of a valid plane (costs are >= 0: unsigned order = float order)
	if (lane >= 1 && lane < NB && L.n > 0 && suf.n > 0)
		key = __float_as_uint(sah_cost(box_area(L.lo[0], L.lo[1], L.lo[2], L.hi[0], L.hi[1], L.hi[2]), L.n,
		                               box_area(suf.lo[0], suf.lo[1], suf.lo[2], suf.hi[0], suf.hi[1], suf.hi[2]), suf.n));
	kmin = __reduce_min_sync(FULL, key);
	split = kmin != 0xffffffffu ? __ffs(__ballot_sync(FULL, key == kmin)) - 1 : 0;
	bl = warp_bin_shfl(L, split);
	br = warp_bin_shfl(suf, split);
}

// Greedy slot assignment of collapse_match (drt_build.cuh) on a warp: lane l owns the pairs 2l and 2l + 1 (pair = child * 8 + slot).
// Every round: the largest cost among the free pairs (redux max on the order-preserving encoding; -0 counts as +0, as it does for
// the serial `>`), then the first pair that has it (redux min) -- the serial loop's choice.
#define DRT_BS_SMEM_BIN_NODES 8
#define DRT_BS_WARP_SCRATCH 144 /* 16 bins x 7 words + 2 child records of 12 words (small split); 144 words (collapse) */
__device__ __forceinline__ void warp_collapse_match(int n, const float* cost, int* slot_of, int* child_in_slot)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = (int)(threadIdx.x & 31u);
	int key[2];
#pragma unroll
	for (int t = 0; t < 2; t++) {
		const float c = cost[2 * lane + t];
		key[t] = ((2 * lane + t) >> 3) < n ? enc_f(c == 0.f ? 0.f : c) : INT_MIN;
	}
	unsigned free_k = (1u << n) - 1u, free_s = 0xffu;
	if (lane < 8) { slot_of[lane] = -1; child_in_slot[lane] = -1; }
	for (int it = 0; it < n; it++) {
		int best = INT_MIN;
#pragma unroll
		for (int t = 0; t < 2; t++) {
			const int p = 2 * lane + t;
			if (((free_k >> (p >> 3)) & 1u) && ((free_s >> (p & 7)) & 1u) && key[t] > best) best = key[t];
		}
		const int wmax = __reduce_max_sync(FULL, best);
		int first = 64;
#pragma unroll
		for (int t = 1; t >= 0; t--) {
			const int p = 2 * lane + t;
			if (((free_k >> (p >> 3)) & 1u) && ((free_s >> (p & 7)) & 1u) && key[t] == wmax) first = p;
		}
		const int p = __reduce_min_sync(FULL, first);
		const int bk = p >> 3, bs = p & 7;
		free_k &= ~(1u << bk); free_s &= ~(1u << bs);
		if (lane == 0) { slot_of[bk] = bs; child_in_slot[bs] = bk; }
	}
}

// collapse_children (drt_build.cuh) on a warp: lane k < 8 holds child k (its BVH2 node, whether it is inner, its surface area); the inner
// child with the largest area (redux max on the area bits, -1 for leaves; ties to the lowest k like the serial `>`) is replaced by its two
// children until there are 8 or only leaves; then lane k writes child[k] and its eight slot costs.
__device__ __forceinline__ int warp_collapse_children(const BuildArrays& A, int w, int* child, float* cost)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = (int)(threadIdx.x & 31u);
	const int src = A.wide_src[w];
	const BuildNode& parent = A.nodes[src];
	const int pleft = parent.left;
	int n = pleft < 0 ? 1 : 2;
	int my = -1, my_left = -1, key = -1; // BVH2 node of this lane's child, its left child, area bits (inner) or -1 (leaf / none)
	if (lane < n) my = pleft < 0 ? src : pleft + lane;
	bool reload = lane < n;
	for (;;) {
		if (reload) {
			const BuildNode& c = A.nodes[my];
			my_left = c.left;
			key = my_left >= 0 ? (int)__float_as_uint(node_area(c)) : -1;
			reload = false;
		}
		if (n >= 8) break;
		const int wmax = __reduce_max_sync(FULL, lane < n ? key : -1);
		if (wmax < 0) break;
		const int best = __ffs(__ballot_sync(FULL, lane < n && key == wmax)) - 1;
		const int l = __shfl_sync(FULL, my_left, best);
		if (lane == best) { my = l; reload = true; }
		if (lane == n) { my = l + 1; reload = true; }
		n++;
	}
	if (lane < n) {
		child[lane] = my;
		const BuildNode& c = A.nodes[my];
		float d[3];
#pragma unroll
		for (int a = 0; a < 3; a++) d[a] = 0.5f * (dec_f(c.lo[a]) + dec_f(c.hi[a])) - 0.5f * (dec_f(parent.lo[a]) + dec_f(parent.hi[a]));
#pragma unroll
		for (int s = 0; s < 8; s++)
			cost[lane * 8 + s] = ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
	}
	return n;
}

// collapse_emit (drt_build.cuh) on a warp: lane s < 8 is SLOT s (quantises its child's box, knows whether it is inner and how many
// triangles it holds), ballot / shuffle prefixes give the ranks the serial loop counts, the 80-byte node is put together in shared
// memory (`image`, 20 words, 16-byte aligned) and stored by five lanes; jobs[] receives the triangle copies.
__device__ __forceinline__ void warp_collapse_emit(const BuildArrays& A, const int* idx, int w, const int* child, const int* child_in_slot, int* jobs, uint32_t* image)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = (int)(threadIdx.x & 31u);
	const BuildNode& parent = A.nodes[A.wide_src[w]];
	const float plo[3] = { dec_f(parent.lo[0]), dec_f(parent.lo[1]), dec_f(parent.lo[2]) };
	const float phi[3] = { dec_f(parent.hi[0]), dec_f(parent.hi[1]), dec_f(parent.hi[2]) };
	uint32_t e[3];
	float scale[3];
	quant_grid(plo, phi, e, scale);
	const int k = lane < 8 ? child_in_slot[lane] : -1;
	const int node = k >= 0 ? child[k] : -1;
	bool inner = false;
	int cnt = 0, cstart = 0;
	uint32_t ql[3] = { 0u, 0u, 0u }, qh[3] = { 0u, 0u, 0u };
	if (node >= 0) {
		const BuildNode& c = A.nodes[node];
		quant_child(c, plo, scale, ql, qh);
		inner = c.left >= 0;
		cnt = inner ? 0 : c.count;
		cstart = c.start;
	}
	const unsigned inner_mask = __ballot_sync(FULL, inner);
	int tri_off = cnt; // -> exclusive prefix sum of the triangle counts in slot order
#pragma unroll
	for (int d = 1; d < 8; d <<= 1) {
		const int up = __shfl_up_sync(FULL, tri_off, d);
		if (lane >= d) tri_off += up;
	}
	const int n_tris = __shfl_sync(FULL, tri_off, 7);
	tri_off -= cnt;
	const int n_inner = __popc(inner_mask);
	int child_base = 0, tri_base = 0;
	if (lane == 0) {
		child_base = n_inner ? atomic_add_i(&A.state->wide_count, n_inner) : 0;
		tri_base = n_tris ? atomic_add_i(&A.state->tri_cursor, n_tris) : 0;
	}
	child_base = __shfl_sync(FULL, child_base, 0);
	tri_base = __shfl_sync(FULL, tri_base, 0);
	uint8_t* img = (uint8_t*)image;
	if (lane == 0) {
		image[0] = __float_as_uint(plo[0]); image[1] = __float_as_uint(plo[1]); image[2] = __float_as_uint(plo[2]);
		image[3] = e[0] | (e[1] << 8) | (e[2] << 16) | (inner_mask << 24);
		image[4] = (uint32_t)child_base; image[5] = (uint32_t)tri_base;
	}
	if (lane < 8) {
		uint32_t meta = 0u;
		if (node >= 0) {
			if (inner) {
				meta = 0x20u | (24u + (uint32_t)lane);
				A.wide_src[child_base + __popc(inner_mask & ((1u << lane) - 1u))] = node;
			}
			else {
				meta = ((cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u)) << 5) | (uint32_t)tri_off;
				for (int t = 0; t < cnt; t++) { jobs[1 + 2 * (tri_off + t)] = tri_base + tri_off + t; jobs[2 + 2 * (tri_off + t)] = idx[cstart + t]; }
			}
		}
		img[24 + lane] = (uint8_t)meta;
		img[32 + lane] = (uint8_t)ql[0]; img[40 + lane] = (uint8_t)ql[1]; img[48 + lane] = (uint8_t)ql[2];
		img[56 + lane] = (uint8_t)qh[0]; img[64 + lane] = (uint8_t)qh[1]; img[72 + lane] = (uint8_t)qh[2];
	}
	if (lane == 0) jobs[0] = n_tris;
	__syncwarp();
	if (lane < 5) ((uint4*)(A.wide + w))[lane] = ((const uint4*)image)[lane];
}

// One warp splits one LARGE node (its bins were filled by the binning pass): the decisions of split_node's large path, with the
// sweep over the DRT_BINS bins of an axis done by warp_plane_search instead of one thread's loop over local-memory arrays.
__device__ __forceinline__ void split_large_node_warp(const BuildArrays& A, int node)
{
	const int lane = (int)(threadIdx.x & 31u);
	BuildNode* nd = A.nodes + node;
	const int count = nd->count, start = nd->start;
	int best_axis = -1, best_split = 0;
	float best_cost = 0.f;
	WarpBin bestL, bestR;
	bestL.n = bestR.n = 0;
	for (int k = 0; k < 3; k++) { bestL.lo[k] = bestR.lo[k] = INFINITY; bestL.hi[k] = bestR.hi[k] = -INFINITY; }
	if (A.level < DRT_MEDIAN_LEVEL) {
		const int* nb = A.bins + (size_t)(A.splits[node].bin_slot - A.bin_slot_base) * 3 * DRT_BINS * DRT_BIN_WORDS;
#pragma unroll
		for (int a = 0; a < 3; a++) {
			WarpBin v;
			v.n = 0; v.lo[0] = v.lo[1] = v.lo[2] = INFINITY; v.hi[0] = v.hi[1] = v.hi[2] = -INFINITY;
			if (lane < DRT_BINS) {
				const int* w = nb + (a * DRT_BINS + lane) * DRT_BIN_WORDS;
				v.n = w[0];
				for (int k = 0; k < 3; k++) { v.lo[k] = dec_f(w[1 + k]); v.hi[k] = dec_f(w[4 + k]); }
			}
			unsigned kmin;
			int split;
			WarpBin bl, br;
			warp_plane_search<DRT_BINS>(v, lane, kmin, split, bl, br);
			if (kmin != 0xffffffffu) {
				const float cost = __uint_as_float(kmin);
				if (best_axis < 0 || cost < best_cost) { best_axis = a; best_cost = cost; best_split = split; bestL = bl; bestR = br; }
			}
		}
	}
	if (lane == 0) {
		const int n_left = best_axis >= 0 ? bestL.n : count / 2;
		const int c0 = atomic_add_i(&A.state->node_count, 2);
		BuildNode* L = A.nodes + c0;
		BuildNode* R = L + 1;
		init_node(L, start, n_left);
		init_node(R, start + n_left, count - n_left);
		if (best_axis >= 0) {
			for (int k = 0; k < 3; k++) { L->lo[k] = enc_f(bestL.lo[k]); L->hi[k] = enc_f(bestL.hi[k]); R->lo[k] = enc_f(bestR.lo[k]); R->hi[k] = enc_f(bestR.hi[k]); }
		}
		BuildSplit sp = A.splits[node];
		sp.axis = best_axis; sp.split_bin = best_split; sp.n_left = n_left;
		A.splits[node] = sp;
		nd->left = c0;
		for (int k = 0; k < 2; k++) {
			BuildSplit cs;
			cs.axis = -1; cs.split_bin = 0; cs.n_left = 0;
			cs.bin_slot = A.nodes[c0 + k].count > DRT_SMALL_NODE ? atomic_add_i(&A.state->bin_slots, 1) : -1;
			A.splits[c0 + k] = cs;
		}
	}
}

__device__ __forceinline__ void split_small_node_warp(const BuildArrays& A, int node, int* bins)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = (int)(threadIdx.x & 31u);
	BuildNode* nd = A.nodes + node;
	const int count = nd->count, start = nd->start;
	if (count <= DRT_LEAF_MAX) { if (lane == 0) nd->left = -1; return; }
	const bool force_median = A.level >= DRT_MEDIAN_LEVEL;
	const float clo[3] = { dec_f(nd->clo[0]), dec_f(nd->clo[1]), dec_f(nd->clo[2]) };
	const float chi[3] = { dec_f(nd->chi[0]), dec_f(nd->chi[1]), dec_f(nd->chi[2]) };
	const bool have = lane < count;
	int prim = 0;
	drt_f4 lo, hi;
	lo.x = lo.y = lo.z = 0.f; hi.x = hi.y = hi.z = 0.f; lo.w = hi.w = 0.f;
	if (have) { prim = A.idx_in[start + lane]; lo = A.tri_lo[prim]; hi = A.tri_hi[prim]; }
	const float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
	const int e[6] = { enc_f(lo.x), enc_f(lo.y), enc_f(lo.z), enc_f(hi.x), enc_f(hi.y), enc_f(hi.z) };
#ifdef DRT_BUILD_SMALL_TIMING
	const long long wt0 = clock64();
#endif

	int best_axis = -1, best_split = 0, best_bin = 0;
	float best_cost = 0.f;
	if (!force_median) {
		// boxes of the bins: shared-memory atomics of the lanes that hold a primitive (a segmented redux -- match.any + redux per group
		// of lanes -- runs once per distinct group and measured 15 k clocks per axis). One bin area, axis after axis: three areas with
		// the three scan chains side by side measured the same (6.8 k clocks) and cost 2.5x the shared memory.
#pragma unroll
		for (int a = 0; a < 3; a++) {
			const int bin = have ? bin_of(c[a], clo[a], chi[a], DRT_SMALL_BINS) : 0;
			if (lane < DRT_SMALL_BINS) {
				int* w = bins + lane * 7;
				w[0] = 0; w[1] = w[2] = w[3] = DRT_ENC_POS_INF; w[4] = w[5] = w[6] = DRT_ENC_NEG_INF;
			}
			__syncwarp();
			if (have) {
				int* w = bins + bin * 7;
				atomicAdd(w, 1);
				atomicMin(w + 1, e[0]); atomicMin(w + 2, e[1]); atomicMin(w + 3, e[2]);
				atomicMax(w + 4, e[3]); atomicMax(w + 5, e[4]); atomicMax(w + 6, e[5]);
			}
			__syncwarp();
			WarpBin v; // lane b < DRT_SMALL_BINS: bin b; the other lanes: empty
			v.n = 0; v.lo[0] = v.lo[1] = v.lo[2] = INFINITY; v.hi[0] = v.hi[1] = v.hi[2] = -INFINITY;
			if (lane < DRT_SMALL_BINS) {
				const int* w = bins + lane * 7;
				v.n = w[0];
				for (int k = 0; k < 3; k++) { v.lo[k] = dec_f(w[1 + k]); v.hi[k] = dec_f(w[4 + k]); }
			}
			__syncwarp();
			unsigned kmin;
			int split;
			WarpBin bl, br;
			warp_plane_search<DRT_SMALL_BINS>(v, lane, kmin, split, bl, br);
			if (kmin != 0xffffffffu) {
				const float cost = __uint_as_float(kmin);
				if (best_axis < 0 || cost < best_cost) { best_axis = a; best_cost = cost; best_split = split; best_bin = bin; }
			}
		}
	}
#ifdef DRT_BUILD_SMALL_TIMING
	const long long wt1 = clock64();
#endif
	// partition (stable on both sides, like the serial code) and the children's exact boxes
	const bool left = have && (best_axis >= 0 ? best_bin < best_split : lane < count / 2);
	const unsigned lm = __ballot_sync(FULL, left), rm = __ballot_sync(FULL, have && !left);
	const int nl = __popc(lm), nr = __popc(rm);
	int c0 = 0;
	if (lane == 0) c0 = atomic_add_i(&A.state->node_count, 2);
	c0 = __shfl_sync(FULL, c0, 0);
	// the children's exact boxes and centroid bounds: shared-memory atomics into two records of 12 words (lo, hi, clo, chi)
	int* rec = bins + DRT_SMALL_BINS * 7;
	if (lane < 24) rec[lane] = ((lane % 12) < 3 || ((lane % 12) >= 6 && (lane % 12) < 9)) ? DRT_ENC_POS_INF : DRT_ENC_NEG_INF;
	__syncwarp();
	if (have) {
		int* r = rec + (left ? 0 : 12);
#pragma unroll
		for (int k = 0; k < 3; k++) {
			const int ec = enc_f(c[k]);
			atomicMin(r + k, e[k]); atomicMax(r + 3 + k, e[3 + k]);
			atomicMin(r + 6 + k, ec); atomicMax(r + 9 + k, ec);
		}
	}
	__syncwarp();
	if (lane < 2) {
		const int side = lane;
		const int* r = rec + 12 * side;
		BuildNode* ch = A.nodes + c0 + side;
		for (int k = 0; k < 3; k++) { ch->lo[k] = r[k]; ch->hi[k] = r[3 + k]; ch->clo[k] = r[6 + k]; ch->chi[k] = r[9 + k]; }
		ch->start = side == 0 ? start : start + nl;
		ch->count = side == 0 ? nl : nr;
		ch->left = -1;
		ch->fill = side == 0 ? nl : nr;
		BuildSplit cs;
		cs.axis = -1; cs.split_bin = 0; cs.n_left = 0; cs.bin_slot = -1; // (the children of a small node are small)
		A.splits[c0 + side] = cs;
	}
	if (have) {
		const unsigned below = (1u << lane) - 1u;
		const int pos = left ? __popc(lm & below) : nl + __popc(rm & below);
		A.idx_in[start + pos] = prim;
		A.pnode_in[start + pos] = c0 + (left ? 0 : 1);
	}
	if (lane == 0) {
		BuildSplit sp = A.splits[node];
		sp.axis = best_axis; sp.split_bin = best_split; sp.n_left = nl;
		A.splits[node] = sp;
		nd->left = c0;
	}
#ifdef DRT_BUILD_SMALL_TIMING
	if (lane == 0 && g_bs_debug) { atomicMax(&g_bs_debug[0], (int)(wt1 - wt0)); atomicMax(&g_bs_debug[1], (int)(clock64() - wt1)); }
#endif
}

// The whole build of a SMALL scene in ONE launch (one block): the level loop of run_build (drt_build_driver.h) with block
// barriers where the host loop has kernel boundaries and the BuildState words read back through shared memory instead of a
// host round trip per level. Same per-index passes, hence the same tree; the level-synchronous path spends ~60 launches and a
// dozen host read-backs on a 330-triangle scene (launch-latency bound: 0.9 ms), this one is bound by the block's own work.
// result[0..5] = bvh2 nodes, wide nodes, levels, wide levels, triangles emitted, 1 if the final index array is A.idx_out.
__global__ void __launch_bounds__(DRT_BUILD_SMALL_THREADS, 1) k_build_small(const BuildArrays A0, int* result)
{
	__shared__ int s_count[2];
	__shared__ int s_level_bins[DRT_BS_SMEM_BIN_NODES * 3 * DRT_BINS * DRT_BIN_WORDS];
	__shared__ alignas(16) int s_warp_scratch[(DRT_BUILD_SMALL_THREADS / 32) * DRT_BS_WARP_SCRATCH]; // per warp: the bins of a small-node split / the collapse's arrays
	BuildArrays A = A0;
	const int tid = (int)threadIdx.x, nt = (int)blockDim.x, n = A.n;
#ifdef DRT_BUILD_SMALL_TIMING
	if (tid == 0) { g_bs_debug = result + 6; result[6] = result[7] = 0; }
	__syncthreads();
#endif
	if (tid == 0) {
		init_node(&A.nodes[0], 0, n);
		BuildSplit rs;
		rs.axis = -1; rs.split_bin = 0; rs.n_left = 0; rs.bin_slot = n > DRT_SMALL_NODE ? 0 : -1;
		A.splits[0] = rs;
		A.wide_src[0] = 0;
		BuildState st;
		st.node_count = 1; st.wide_count = 1; st.tri_cursor = 0; st.bin_slots = n > DRT_SMALL_NODE ? 1 : 0;
		*A.state = st;
	}
	__syncthreads();
	for (int i = tid; i < n; i += nt) prim_bounds(A, i);
	__syncthreads();
#ifdef DRT_BUILD_SMALL_TIMING
	long long tc = clock64();
#define DRT_BS_TICK(SLOT) do { if (tid == 0 && (SLOT) < DRT_BUILD_SMALL_RESULT_INTS) { const long long now = clock64(); result[SLOT] = (int)(now - tc); tc = now; } } while (0)
#else
#define DRT_BS_TICK(SLOT) do { } while (0)
#endif
	DRT_BS_TICK(8);
	int level_start = 0, level_end = 1, level = 0, slot_base = 0, slot_end = n > DRT_SMALL_NODE ? 1 : 0, swapped = 0;
	while (level_end > level_start) {
		A.level_start = level_start; A.level_end = level_end; A.level = level; A.bin_slot_base = slot_base;
		// the bins of a level with few large nodes live in shared memory (the binning pass is ~7 k atomics on a handful of words:
		// 8-12 k clocks in global memory); levels with more large nodes use the global array
		A.bins = (slot_end - slot_base) <= DRT_BS_SMEM_BIN_NODES ? s_level_bins : A0.bins;
		const int bin_words = (slot_end - slot_base) * 3 * DRT_BINS * DRT_BIN_WORDS;
		if (bin_words > 0) {
			for (int j = tid; j < bin_words; j += nt) init_bin_word(A, j, bin_words);
			__syncthreads();
			for (int i = tid; i < n; i += nt) bin_prim(A, i);
			__syncthreads();
		}
		DRT_BS_TICK(9 + 3 * level);
		// one warp per node, in registers: small nodes (<= 32 primitives) bin their own primitives, large nodes read the bins the
		// pass above filled
		for (int j = tid >> 5; j < level_end - level_start; j += nt >> 5) {
			if (A.nodes[level_start + j].count <= DRT_SMALL_NODE) split_small_node_warp(A, level_start + j, s_warp_scratch + (tid >> 5) * DRT_BS_WARP_SCRATCH);
			else split_large_node_warp(A, level_start + j);
		}
		__syncthreads();
		DRT_BS_TICK(10 + 3 * level);
		for (int i = tid; i < n; i += nt) partition_prim(A, i);
		__syncthreads();
		DRT_BS_TICK(11 + 3 * level);
		if (tid == 0) { s_count[0] = A.state->node_count; s_count[1] = A.state->bin_slots; }
		__syncthreads();
		int* t = A.idx_in; A.idx_in = A.idx_out; A.idx_out = t;
		t = A.pnode_in; A.pnode_in = A.pnode_out; A.pnode_out = t;
		swapped ^= 1;
		level_start = level_end; level_end = s_count[0];
		slot_base = slot_end; slot_end = s_count[1];
		level++;
		__syncthreads(); // s_count is rewritten at the end of the next level
	}
	const int bvh2_nodes = level_end;
	int ws = 0, we = 1, wl = 0;
	while (we > ws) {
		// one warp per wide node, its scratch in shared memory: children and slot costs (lane = child), greedy slot assignment (64
		// pairs on 32 lanes), the node itself (lane = slot) and its triangle copies (lane = triangle)
		for (int j = tid >> 5; j < we - ws; j += nt >> 5) {
			int* scr = s_warp_scratch + (tid >> 5) * DRT_BS_WARP_SCRATCH;
			int* child = scr, * slot_of = scr + 8, * child_in_slot = scr + 16, * jobs = scr + 24; // 8 + 8 + 8 + 49 ints
			float* cost = (float*)(scr + 80);                                                  // 64 floats
			const int lane = (int)(tid & 31);
			const int n = warp_collapse_children(A, ws + j, child, cost);
			__syncwarp();
			warp_collapse_match(n, cost, slot_of, child_in_slot);
			__syncwarp();
			warp_collapse_emit(A, A.idx_in, ws + j, child, child_in_slot, jobs, (uint32_t*)cost); // (the costs are spent: their place takes the node image)
			__syncwarp();
			if (lane < jobs[0]) write_triangle(A, jobs[1 + 2 * lane], jobs[2 + 2 * lane]);
			__syncwarp();
		}
		__syncthreads();
		DRT_BS_TICK(57 + wl);
		if (tid == 0) s_count[0] = A.state->wide_count;
		__syncthreads();
		ws = we; we = s_count[0]; wl++;
		__syncthreads();
	}
	if (tid == 0) {
		result[0] = bvh2_nodes; result[1] = we; result[2] = level; result[3] = wl; result[4] = A.state->tri_cursor; result[5] = swapped;
	}
}

cudaError_t launch_build_small(const BuildArrays& A, int* result, cudaStream_t s)
{
	if (A.n <= 0) return cudaErrorInvalidValue;
	// one thread per primitive, and warps enough for the nodes of a level (one warp per small node): at least 16 warps
	int threads = ((A.n + 31) / 32) * 32;
	if (threads < 512) threads = 512;
	if (A.n > 128) threads = DRT_BUILD_SMALL_THREADS; // (levels of more than 16 nodes: one warp each)
	if (threads > DRT_BUILD_SMALL_THREADS) threads = DRT_BUILD_SMALL_THREADS;
	static const char* env = getenv("DRT_BUILD_SMALL_BLOCK"); // measurement override
	if (env && atoi(env) >= 32 && atoi(env) <= DRT_BUILD_SMALL_THREADS) threads = (atoi(env) / 32) * 32;
	k_build_small<<<1, threads, 0, s>>>(A, result);
	return cudaGetLastError();
}

cudaError_t launch_build_pass(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n, cudaStream_t s)
{
	if (n <= 0) return cudaSuccess;
	const unsigned blocks = (unsigned)((n + 255) / 256);
	switch (pass) {
	case PASS_PRIM_BOUNDS: k_build<PASS_PRIM_BOUNDS><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	case PASS_INIT_BINS: k_build<PASS_INIT_BINS><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	case PASS_BIN: k_build<PASS_BIN><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	case PASS_SPLIT: k_build<PASS_SPLIT><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	case PASS_PARTITION: k_build<PASS_PARTITION><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	case PASS_COLLAPSE: k_build<PASS_COLLAPSE><<<blocks, 256, 0, s>>>(A, idx_final, a, b, n); break;
	}
	return cudaGetLastError();
}

} // namespace drt
