// drt_regroup.cuh -- the path loop with BLOCK-LEVEL REGROUPING of path contexts (render variant 16).
//
// north_star subsystem (3) asks for "warp-level ballot/shuffle ray compaction, and a wavefront split (generate, extend, shade,
// connect) versus megakernel choice made on measured occupancy". The shipped megakernel (drt_phased.cuh render_warp_phased)
// binds a path to a lane: a warp votes for NODE / TRI / SHADE and the lanes waiting for another step idle -- ncu counts 17.8 of
// 32 lanes per issued instruction on the benchmarked launch (profiles/r02a_render_phased_raw.txt). This file is the other arm:
// a WAVEFRONT INSIDE A PERSISTENT BLOCK. All state of a path context lives in shared memory ([field][context] words); the
// block owns DRT_RG_CTX contexts and five queues of context indices, one per step:
//     END   sample finished (or context needs a pixel): sky on a miss, per-sample post-process, accumulate, next pixel / next
//           sample, camera ray                                                         (generate;  RayGen.cuh:63-86,164-171)
//     NODE  pop one wide node, slab-test its 8 children                               (extend;   BVHTraversal.cuh:22-47)
//     TRI   one (or two) Moller-Trumbore tests + closest / any-hit bookkeeping        (extend;   Intersection.cu:4-36)
//     HIT   closest hit found: material, face-forward normal, next origin, sun ray    (shade;    RayGen.cuh:104-125)
//     SCAT  sun shadow result + ONE try of the scatter rejection loop                 (connect;  RayGen.cuh:126-134, Random.cu:50-58)
// Execution is round-synchronous: in a round every queue is cut into batches of 32 contexts, the block's warps take the
// batches round-robin, each lane runs the step on ONE context (whatever context -- that is the regrouping) and pushes it, by
// warp ballot + popc + shuffle compaction, into the queue of its next step for the NEXT round; a remainder of fewer than 32
// entries is carried to the next round unless the block has nothing else to do. So every step is issued for 32 lanes that
// all need it, at the price of loading and storing the context around each step.
// Per context the sequence of steps and every arithmetic operation are those of shade_step / node_step / tri_step: the
// accumulated image has the same bits as every other variant's (tests/test_gpu_parity.py).
#pragma once
#include "drt_phased.cuh"

#if defined(__CUDACC__)
namespace drt {

#ifndef DRT_RG_CTX
#define DRT_RG_CTX 512      // path contexts per block
#endif
#define DRT_RG_THREADS 256  // 8 warps
#define DRT_RG_WARPS (DRT_RG_THREADS / 32)
#ifndef DRT_RG_BLOCKS_PER_SM
#define DRT_RG_BLOCKS_PER_SM 2
#endif
#ifndef DRT_RG_STACK
#define DRT_RG_STACK 4      // traversal stack entries per context in shared memory; deeper entries go to P.rg_overflow
#endif

enum { RQ_END = 0, RQ_NODE = 1, RQ_HIT = 2, RQ_TRI = 3, RQ_SCAT = 4, RQ_COUNT = 5, RQ_NONE = -1 };

// words of one context, [field][DRT_RG_CTX]
enum {
	RF_OX = 0, RF_OY, RF_OZ, RF_DX, RF_DY, RF_DZ,       // current ray
	RF_IDX, RF_IDY, RF_IDZ,                             // slab test: unscaled reciprocals of the direction ...
	RF_BNX, RF_BNY, RF_BNZ, RF_BFX, RF_BFY, RF_BFZ,     // ... and the near / far biases (ray_box_setup)
	RF_NGBASE, RF_NGBITS, RF_TRIBASE, RF_TRIBITS,       // traversal cursor (LaneTrav)
	RF_SP,                                              // stack depth | octant << 16
	RF_PRIM, RF_T,                                      // hit of the ray being traced (leaf-order slot, distance)
	RF_LX, RF_LY, RF_LZ, RF_TX, RF_TY, RF_TZ,           // path: light, throughput
	RF_SEED, RF_K, RF_BOUNCE, RF_FLAGS, RF_HPRIM, RF_XY, // RNG, sample of the launch, bounce, flags, closest-hit slot, pixel x | y << 16
	RF_STACK,                                           // 2 * DRT_RG_STACK words
	RF_COUNT = RF_STACK + 2 * DRT_RG_STACK
};
enum { RFL_SHADOW = 1, RFL_FLIP = 2, RFL_SKY = 4, RFL_NEWPIX = 8 };

// ring capacity per queue: a power of two >= 2 * DRT_RG_CTX (a context holds its old slot until the round ends)
#define DRT_RG_QCAP (DRT_RG_CTX <= 128 ? 256 : (DRT_RG_CTX <= 256 ? 512 : (DRT_RG_CTX <= 512 ? 1024 : 2048)))
#define DRT_RG_STATE_BYTES ((size_t)RF_COUNT * DRT_RG_CTX * 4)
#define DRT_RG_QUEUE_BYTES ((size_t)RQ_COUNT * DRT_RG_QCAP * 2)
#define DRT_RG_SMEM_BYTES (DRT_RG_STATE_BYTES + DRT_RG_QUEUE_BYTES)

// Traversal stack of one context: the first DRT_RG_STACK entries in shared memory, the rest (never touched on the bundled
// scenes: measured depth <= 3, soup-10M 7) in a global overflow area of DRT_PSTACK entries per resident context.
struct RegroupStack {
	uint32_t* s;     // &state[RF_STACK][ctx]
	uint2* overflow; // DRT_PSTACK entries of this context, or nullptr (then deeper entries are dropped like LaneStack's beyond DRT_PSTACK)
	__device__ __forceinline__ void put(int i, uint32_t vx, uint32_t vy)
	{
		if (i < DRT_RG_STACK) { s[(2 * i) * DRT_RG_CTX] = vx; s[(2 * i + 1) * DRT_RG_CTX] = vy; }
		else if (overflow) overflow[i] = make_uint2(vx, vy);
	}
	__device__ __forceinline__ void get(int i, uint32_t& vx, uint32_t& vy) const
	{
		if (i < DRT_RG_STACK) { vx = s[(2 * i) * DRT_RG_CTX]; vy = s[(2 * i + 1) * DRT_RG_CTX]; }
		else if (overflow) { const uint2 v = overflow[i]; vx = v.x; vy = v.y; }
		else { vx = 0u; vy = 0u; }
	}
};

struct RegroupShared {
	uint32_t* st;     // [RF_COUNT][DRT_RG_CTX]
	uint16_t* q;      // [RQ_COUNT][DRT_RG_QCAP] rings of context indices
	unsigned* tail;   // [2][8] entries ever pushed per queue in even / odd rounds (monotonic; the heads are block-uniform registers)
	const uint8_t* lut;
	unsigned frozen;  // lane k < 8: the OTHER parity's count of queue k, which does not move during this round
	int parity;       // parity of the current round: its pushes count in tail[parity][.]
};

__device__ __forceinline__ float rg_f(const uint32_t* st, int f, int c) { return __uint_as_float(st[f * DRT_RG_CTX + c]); }
__device__ __forceinline__ uint32_t rg_u(const uint32_t* st, int f, int c) { return st[f * DRT_RG_CTX + c]; }
__device__ __forceinline__ void rg_sf(uint32_t* st, int f, int c, float v) { st[f * DRT_RG_CTX + c] = __float_as_uint(v); }
__device__ __forceinline__ void rg_su(uint32_t* st, int f, int c, uint32_t v) { st[f * DRT_RG_CTX + c] = v; }
__device__ __forceinline__ V3 rg_f3(const uint32_t* st, int f, int c) { return mk3(rg_f(st, f, c), rg_f(st, f + 1, c), rg_f(st, f + 2, c)); }
__device__ __forceinline__ void rg_sf3(uint32_t* st, int f, int c, V3 v) { rg_sf(st, f, c, v.x); rg_sf(st, f + 1, c, v.y); rg_sf(st, f + 2, c, v.z); }

// Warp-level compaction into the queues: every lane names the queue its context goes to (RQ_NONE: nowhere). match.any groups
// the lanes by destination, the lowest lane of each group reserves the group's slots at the queue's tail with one
// shared-memory atomic, a shuffle hands the base to the group, popc of the lower lanes of the group is a lane's rank.
__device__ __forceinline__ void rg_push(const RegroupShared& G, int dest, int c)
{
	const unsigned lane = threadIdx.x & 31u;
	const unsigned same = __match_any_sync(0xffffffffu, dest);
	const int leader = __ffs((int)same) - 1;
	// ring position = (pushes of the other parity's rounds, frozen during this round) + (pushes of this parity's rounds)
	unsigned base = __shfl_sync(0xffffffffu, G.frozen, dest & 7);
	if (dest >= 0 && (int)lane == leader) base += atomicAdd(G.tail + G.parity * 8 + dest, (unsigned)__popc(same));
	base = __shfl_sync(0xffffffffu, base, leader);
	if (dest >= 0) G.q[dest * DRT_RG_QCAP + ((base + (unsigned)__popc(same & ((1u << lane) - 1u))) & (DRT_RG_QCAP - 1))] = (uint16_t)c;
}

// A new ray (o, d) of context c: slab-test constants and a fresh traversal cursor. Returns the queue of its first step.
__device__ __forceinline__ int rg_new_ray(const RenderParams& P, uint32_t* st, int c, V3 o, V3 d, bool store_o, bool shadow, uint32_t& flags)
{
	if (store_o) rg_sf3(st, RF_OX, c, o);
	rg_sf3(st, RF_DX, c, d);
	rg_su(st, RF_PRIM, c, 0xffffffffu);
	rg_sf(st, RF_T, c, FLT_MAX);
	if (P.scene.num_nodes <= 0) { // nothing to traverse: the ray is finished at once
		if (shadow) return RQ_SCAT;
		flags |= RFL_SKY;
		return RQ_END;
	}
	RayBox R;
	float id3[3];
	ray_box_setup(o, d, R, id3);
	rg_sf(st, RF_IDX, c, id3[0]); rg_sf(st, RF_IDY, c, id3[1]); rg_sf(st, RF_IDZ, c, id3[2]);
	rg_sf(st, RF_BNX, c, R.bnx); rg_sf(st, RF_BNY, c, R.bny); rg_sf(st, RF_BNZ, c, R.bnz);
	rg_sf(st, RF_BFX, c, R.bfx); rg_sf(st, RF_BFY, c, R.bfy); rg_sf(st, RF_BFZ, c, R.bfz);
	rg_su(st, RF_NGBASE, c, 0u); rg_su(st, RF_NGBITS, c, 0x80000000u); // root (trav_begin)
	rg_su(st, RF_TRIBITS, c, 0u);
	rg_su(st, RF_SP, c, R.oct_inv << 16);
	return RQ_NODE;
}

// Queue of a context whose traversal just ran out of work.
__device__ __forceinline__ int rg_finished(uint32_t* st, int c, int prim)
{
	const uint32_t flags = rg_u(st, RF_FLAGS, c);
	if (flags & RFL_SHADOW) return RQ_SCAT;
	if (prim >= 0) return RQ_HIT;
	rg_su(st, RF_FLAGS, c, flags | RFL_SKY);
	return RQ_END;
}

template <int FMT>
__device__ __forceinline__ int rg_step_node(const RenderParams& P, const RegroupShared& G, int c, uint2* overflow)
{
	uint32_t* st = G.st;
	RayBox R;
	const float idx = rg_f(st, RF_IDX, c), idy = rg_f(st, RF_IDY, c), idz = rg_f(st, RF_IDZ, c);
	R.idnx = idx * DRT_BOX_KN; R.idny = idy * DRT_BOX_KN; R.idnz = idz * DRT_BOX_KN; // the two multiplies of ray_box_setup
	R.idfx = idx * DRT_BOX_KF; R.idfy = idy * DRT_BOX_KF; R.idfz = idz * DRT_BOX_KF;
	R.bnx = rg_f(st, RF_BNX, c); R.bny = rg_f(st, RF_BNY, c); R.bnz = rg_f(st, RF_BNZ, c);
	R.bfx = rg_f(st, RF_BFX, c); R.bfy = rg_f(st, RF_BFY, c); R.bfz = rg_f(st, RF_BFZ, c);
	const uint32_t spw = rg_u(st, RF_SP, c);
	R.oct_inv = spw >> 16;
	LaneTrav L;
	L.ng_base = rg_u(st, RF_NGBASE, c); L.ng_bits = rg_u(st, RF_NGBITS, c);
	L.tri_base = 0u; L.tri_bits = 0u;
	L.sp = (int)(spw & 0xffffu);
	L.hit.prim = -1; L.hit.t = rg_f(st, RF_T, c); L.hit.u = L.hit.v = 0.f;
	RegroupStack K;
	K.s = st + RF_STACK * DRT_RG_CTX + c; K.overflow = overflow;
	if (FMT == NODE_FMT_F32) node_step_f32<false>(P.scene.nodes_f32, L, K, R, nullptr, G.lut);
	else node_step_q8<false>(P.scene.nodes, L, K, R, nullptr, G.lut);
	const int phase = trav_advance(L, K);
	rg_su(st, RF_NGBASE, c, L.ng_base); rg_su(st, RF_NGBITS, c, L.ng_bits);
	rg_su(st, RF_TRIBASE, c, L.tri_base); rg_su(st, RF_TRIBITS, c, L.tri_bits);
	rg_su(st, RF_SP, c, (uint32_t)L.sp | (R.oct_inv << 16));
	if (phase == PH_NODE) return RQ_NODE;
	if (phase == PH_TRI) return RQ_TRI;
	return rg_finished(st, c, (int)rg_u(st, RF_PRIM, c));
}

template <bool STRICT, bool TRI2>
__device__ __forceinline__ int rg_step_tri(const RenderParams& P, const RegroupShared& G, int c, uint2* overflow)
{
	uint32_t* st = G.st;
	const V3 o = rg_f3(st, RF_OX, c), d = rg_f3(st, RF_DX, c);
	const uint32_t spw = rg_u(st, RF_SP, c);
	LaneTrav L;
	L.ng_base = rg_u(st, RF_NGBASE, c); L.ng_bits = rg_u(st, RF_NGBITS, c);
	L.tri_base = rg_u(st, RF_TRIBASE, c); L.tri_bits = rg_u(st, RF_TRIBITS, c);
	L.sp = (int)(spw & 0xffffu);
	L.hit.prim = (int)rg_u(st, RF_PRIM, c); L.hit.t = rg_f(st, RF_T, c); L.hit.u = L.hit.v = 0.f;
	const bool anyhit = (rg_u(st, RF_FLAGS, c) & RFL_SHADOW) != 0u;
	const bool fin = TRI2 ? tri_step2<STRICT, false>(P.scene, P.scene.tris, L, o, d, anyhit, nullptr)
	                      : tri_step<STRICT, false>(P.scene, P.scene.tris, L, o, d, anyhit, nullptr);
	rg_su(st, RF_PRIM, c, (uint32_t)L.hit.prim); rg_sf(st, RF_T, c, L.hit.t);
	if (fin) return RQ_SCAT; // an any-hit ray found its occluder (RayTest ends at the first accepted triangle)
	RegroupStack K;
	K.s = st + RF_STACK * DRT_RG_CTX + c; K.overflow = overflow;
	const int phase = trav_advance(L, K);
	rg_su(st, RF_TRIBITS, c, L.tri_bits);
	rg_su(st, RF_NGBASE, c, L.ng_base); rg_su(st, RF_NGBITS, c, L.ng_bits);
	rg_su(st, RF_SP, c, (uint32_t)L.sp | (spw & 0xffff0000u));
	if (phase == PH_NODE) return RQ_NODE;
	if (phase == PH_TRI) return RQ_TRI;
	return rg_finished(st, c, L.hit.prim);
}

// HIT: the `found` branch of shade_step for a closest-hit ray (RayGen.cuh:104-125).
template <bool STRICT>
__device__ __forceinline__ int rg_step_hit(const RenderParams& P, const RegroupShared& G, int c, bool rtl, PathCounters& pc)
{
	typedef M<STRICT> m;
	uint32_t* st = G.st;
	const SceneView& sc = P.scene;
	const ShadeParams& SP = P.shade;
	const V3 o = rg_f3(st, RF_OX, c), d = rg_f3(st, RF_DX, c);
	const int prim = (int)rg_u(st, RF_PRIM, c);
	const float t = rg_f(st, RF_T, c);
	uint32_t seed = rg_u(st, RF_SEED, c);
	V3 thr = rg_f3(st, RF_TX, c);
	uint32_t flags = rg_u(st, RF_FLAGS, c);
	pc.closest++;
	seed += rg_u(st, RF_BOUNCE, c); // RayGen.cuh:91
	const drt_f4 r1 = ld_f4(sc.tris + 3 * (size_t)prim + 1);
	const uint32_t mat = as_uint(r1.w) & 0x7fffffffu;
	const drt_f4 mrec = ld_f4(sc.materials + mat);
	const int tex = (int)as_uint(mrec.w);
	if (tex < 0) thr = m::mul3(thr, mk3(mrec.x, mrec.y, mrec.z));
	else {
		const V2 bc = hit_barycentrics<STRICT>(sc.tris, prim, o, d);
		const V2 tuv = interp_uv<STRICT>(sc, prim, bc.x, bc.y);
		thr = m::mul3(thr, texture_get_pixel<STRICT>(sc, tex, tuv, nullptr));
	}
	rg_sf3(st, RF_TX, c, thr);
	const drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)prim);
	V3 fn = mk3(s0.x, s0.y, s0.z);
	const V3 pos = m::add3(o, m::scale(d, t));
	flags &= ~(uint32_t)RFL_FLIP;
	if ((STRICT ? m::dot(fn, m::normalize(d)) : m::dot(fn, d)) > 0.f) { fn = m::scale(fn, -1.f); flags |= RFL_FLIP; }
	rg_su(st, RF_HPRIM, c, (uint32_t)prim);
	const V3 no = m::add3(pos, m::scale(fn, 0.001f)); // RayGen.cuh:121
	int dest;
	if (SP.enable_sunlight) {
		const V3 nd = m::add3(SP.sunpos, m::scale(random_unit_vec3<STRICT>(seed, rtl), 1.5f)); // RayGen.cuh:124-125
		flags |= RFL_SHADOW;
		dest = rg_new_ray(P, st, c, no, nd, true, true, flags);
	}
	else {
		rg_sf3(st, RF_OX, c, no);
		dest = RQ_SCAT;
	}
	rg_su(st, RF_SEED, c, seed);
	rg_su(st, RF_FLAGS, c, flags);
	return dest;
}

// SCAT: the sun shadow result (RayGen.cuh:126-127) and ONE try of randomUnitSphereVec3's rejection loop (Random.cu:50-58);
// a rejected try leaves the context in the SCAT queue with its RNG state advanced, exactly where the loop would continue.
template <bool STRICT>
__device__ __forceinline__ int rg_step_scat(const RenderParams& P, const RegroupShared& G, int c, bool rtl, PathCounters& pc)
{
	typedef M<STRICT> m;
	uint32_t* st = G.st;
	const ShadeParams& SP = P.shade;
	uint32_t flags = rg_u(st, RF_FLAGS, c);
	uint32_t seed = rg_u(st, RF_SEED, c);
	if (flags & RFL_SHADOW) {
		pc.shadow++;
		if ((int)rg_u(st, RF_PRIM, c) < 0) rg_sf3(st, RF_LX, c, m::add3(rg_f3(st, RF_LX, c), m::mul3(SP.suncol, rg_f3(st, RF_TX, c))));
		flags &= ~(uint32_t)RFL_SHADOW;
		rg_su(st, RF_FLAGS, c, flags);
	}
	const V3 p = random_unit_vec3<STRICT>(seed, rtl);
	rg_su(st, RF_SEED, c, seed);
	bool accept;
	if (STRICT) { const float len = m::length(p); accept = m::mul(len, len) < 1.f; }
	else accept = m::dot(p, p) < 1.f;
	if (!accept) return RQ_SCAT;
	const drt_f4 s0 = ld_f4(P.scene.shade + 3 * (size_t)rg_u(st, RF_HPRIM, c));
	V3 fn = mk3(s0.x, s0.y, s0.z);
	if (flags & RFL_FLIP) fn = m::scale(fn, -1.f);
	const V3 nd = m::add3(fn, p); // RayGen.cuh:133-134
	const uint32_t bounce = rg_u(st, RF_BOUNCE, c) + 1u;
	rg_su(st, RF_BOUNCE, c, bounce);
	if ((int)bounce > SP.bounce_limit) return RQ_END; // path over without a sky term
	const int dest = rg_new_ray(P, st, c, rg_f3(st, RF_OX, c), nd, false, false, flags);
	if (flags & RFL_SKY) rg_su(st, RF_FLAGS, c, flags); // (empty scene: the new ray is a miss at once)
	return dest;
}

// END: sky term of a missed closest-hit ray, per-sample post-process + accumulation (RenderKernel.cu:29), then the next sample of
// the pixel or the next pixel of the launch, and its camera ray. Pixels come from ONE global counter, 8x4 tile order.
template <bool STRICT>
__device__ __forceinline__ int rg_step_end(const RenderParams& P, const RegroupShared& G, int c, bool valid, bool rtl, PathCounters& pc)
{
	typedef M<STRICT> m;
	uint32_t* st = G.st;
	const ShadeParams& SP = P.shade;
	const unsigned lane = threadIdx.x & 31u;
	uint32_t flags = valid ? rg_u(st, RF_FLAGS, c) : 0u;
	uint32_t k = valid ? rg_u(st, RF_K, c) : 0u;
	uint32_t xy = valid ? rg_u(st, RF_XY, c) : 0u;
	bool need_pixel = valid && (flags & RFL_NEWPIX) != 0u;
	if (valid && !need_pixel) {
		V3 light = rg_f3(st, RF_LX, c);
		if (flags & RFL_SKY) {
			pc.closest++;
			const V3 sky = sky_model<STRICT>(rg_f3(st, RF_DX, c), SP.sky_color);
			light = m::add3(light, m::scale(m::mul3(sky, rg_f3(st, RF_TX, c)), SP.sky_intensity));
		}
		const size_t p = (size_t)(xy & 0xffffu) + (size_t)(xy >> 16) * (size_t)P.width;
		const float4 a4 = P.accum[p];
		const V3 acc = m::add3(mk3(a4.x, a4.y, a4.z), post_process<STRICT>(P, light));
		P.accum[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
		k++;
		if (k >= P.n_samples) {
			P.display[p] = make_float4(acc.x / P.total_samples, acc.y / P.total_samples, acc.z / P.total_samples, 1.f);
			need_pixel = true;
		}
	}
	// next pixel of the launch for the lanes that need one (warp-uniform: one global atomic per batch)
	const unsigned need = __ballot_sync(0xffffffffu, need_pixel);
	if (need != 0u) {
		unsigned base = 0u;
		if (lane == 0u) base = atomicAdd(P.work_queue, (unsigned)__popc(need));
		base = __shfl_sync(0xffffffffu, base, 0);
		if (need_pixel) {
			const unsigned long long i = (unsigned long long)base + (unsigned)__popc(need & ((1u << lane) - 1u));
			const unsigned tiles_x = (unsigned)(P.width + 7) >> 3, tiles_y = (unsigned)(P.height + 3) >> 2;
			const unsigned long long tile = (i >> 5) * P.tile_stride + P.tile_first; // k-th 8x4 tile this launch owns
			if (tile >= (unsigned long long)tiles_x * tiles_y || P.n_samples == 0u) return RQ_NONE; // the launch has no more pixels: the context retires
			const unsigned ty = (unsigned)(tile / tiles_x), tx = (unsigned)(tile - (unsigned long long)ty * tiles_x);
			const unsigned x = tx * 8u + ((unsigned)i & 7u), y = ty * 4u + (((unsigned)i >> 3) & 3u);
			if (x >= (unsigned)P.width || y >= (unsigned)P.height) { // ragged edge: ask again next round
				rg_su(st, RF_FLAGS, c, RFL_NEWPIX);
				return RQ_END;
			}
			xy = x | (y << 16);
			k = 0u;
			rg_su(st, RF_XY, c, xy);
		}
	}
	if (!valid) return RQ_NONE;
	// path_begin_sample (RayGen.cuh:68-78)
	const uint32_t x = xy & 0xffffu, y = xy >> 16;
	uint32_t seed = (x + y * (uint32_t)P.width) * (P.frame_first + k * P.frame_stride);
	const RayOD r = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, seed, rtl);
	rg_su(st, RF_K, c, k);
	rg_su(st, RF_SEED, c, seed);
	rg_su(st, RF_BOUNCE, c, 0u);
	rg_sf3(st, RF_LX, c, mk3(0.f, 0.f, 0.f));
	rg_sf3(st, RF_TX, c, mk3(1.f, 1.f, 1.f));
	flags = 0u;
	const int dest = rg_new_ray(P, st, c, r.o, r.d, true, false, flags);
	rg_su(st, RF_FLAGS, c, flags);
	return dest;
}

// The block: DRT_RG_CTX contexts, round-synchronous batches. In a round the warps take the FULL batches (32 contexts) of every
// queue round-robin; what a queue holds beyond a multiple of 32 simply stays in its ring for the next round. Only when no queue
// of the block has a full batch (start and end of the launch) do the remainders run as partial batches.
template <bool STRICT, int FMT, bool TRI2>
__device__ __forceinline__ void render_block_regroup(const RenderParams& P, bool rtl, uint2* overflow_block, uint32_t* smem_words, unsigned* tail, uint8_t* lut,
	PathCounters& pc)
{
	RegroupShared G;
	G.st = smem_words;
	G.q = (uint16_t*)(smem_words + RF_COUNT * DRT_RG_CTX);
	G.tail = tail;
	G.lut = lut;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	for (int i = tid; i < DRT_PERMUTE_LUT_BYTES / 16; i += DRT_RG_THREADS) ((uint4*)lut)[i] = __ldg((const uint4*)g_permute_lut.v + i);
	if (tid < 16) tail[tid] = tid == 8 + RQ_END ? (unsigned)DRT_RG_CTX : 0u; // the initial entries count as pushed in an odd round
	for (int c = tid; c < DRT_RG_CTX; c += DRT_RG_THREADS) { // every context starts by asking for a pixel
		G.st[RF_FLAGS * DRT_RG_CTX + c] = RFL_NEWPIX;
		G.st[RF_K * DRT_RG_CTX + c] = 0u;
		G.st[RF_XY * DRT_RG_CTX + c] = 0u;
		G.q[RQ_END * DRT_RG_QCAP + c] = (uint16_t)c;
	}
	unsigned head[RQ_COUNT];
#pragma unroll
	for (int k = 0; k < RQ_COUNT; k++) head[k] = 0u;
	// One barrier per round: the pushes of a round count in tail[parity of the round], so at the start of a round the other
	// parity's counters are final (last written in the previous round, which the barrier closed) and this parity's counters
	// were final when every thread read them one round ago (cached in `seen`).
	unsigned seen[RQ_COUNT];
#pragma unroll
	for (int k = 0; k < RQ_COUNT; k++) seen[k] = 0u;
	for (unsigned round = 0;; round++) {
		__syncthreads(); // every push of the previous round is in its ring
		const int par = (int)(round & 1u);
		G.parity = par;
		G.frozen = tail[(par ^ 1) * 8 + (lane & 7)];
		int take[RQ_COUNT], items[RQ_COUNT];
		int total = 0, nfull = 0;
#pragma unroll
		for (int k = 0; k < RQ_COUNT; k++) {
			const unsigned other = tail[(par ^ 1) * 8 + k];
			take[k] = (int)(other + seen[k] - head[k]); // seen[k] = this parity's counter as of one round ago = as of now
			seen[k] = other;
			total += take[k]; nfull += take[k] >> 5;
		}
		if (total == 0) break;
		int nitems = 0;
#pragma unroll
		for (int k = 0; k < RQ_COUNT; k++) {
			if (nfull > 0) take[k] &= ~31;           // full batches only; the remainder waits in the ring
			items[k] = (take[k] + 31) >> 5;
			nitems += items[k];
		}
		for (int it = warp; it < nitems; it += DRT_RG_WARPS) {
			int k = 0, b = it;
#pragma unroll
			for (int kk = 0; kk < RQ_COUNT - 1; kk++) if (k == kk && b >= items[kk]) { b -= items[kk]; k = kk + 1; }
			const int idx = b * 32 + lane;
			const int tk = k == 0 ? take[0] : (k == 1 ? take[1] : (k == 2 ? take[2] : (k == 3 ? take[3] : take[4])));
			const unsigned hk = k == 0 ? head[0] : (k == 1 ? head[1] : (k == 2 ? head[2] : (k == 3 ? head[3] : head[4])));
			const bool valid = idx < tk;
			const int c = valid ? (int)G.q[k * DRT_RG_QCAP + ((hk + (unsigned)idx) & (DRT_RG_QCAP - 1))] : 0;
			int dest = RQ_NONE;
			uint2* ovf = overflow_block ? overflow_block + (size_t)c * DRT_PSTACK : nullptr;
			if (k == RQ_NODE) { if (valid) dest = rg_step_node<FMT>(P, G, c, ovf); }
			else if (k == RQ_TRI) { if (valid) dest = rg_step_tri<STRICT, TRI2>(P, G, c, ovf); }
			else if (k == RQ_HIT) { if (valid) dest = rg_step_hit<STRICT>(P, G, c, rtl, pc); }
			else if (k == RQ_SCAT) { if (valid) dest = rg_step_scat<STRICT>(P, G, c, rtl, pc); }
			else dest = rg_step_end<STRICT>(P, G, c, valid, rtl, pc); // (warp-collective pixel fetch inside)
			rg_push(G, dest, c);
		}
#pragma unroll
		for (int k = 0; k < RQ_COUNT; k++) head[k] += (unsigned)take[k];
	}
}

} // namespace drt
#endif
