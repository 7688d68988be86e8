// drt_phased.cuh -- the path loop as per-lane PHASE STEPS, scheduled per warp by vote.
//
// Same per-lane arithmetic as drt_trace.cuh / drt_shade.cuh (and therefore the same results as the
// reference's RayGen -> TraceRay/RayTest -> traverseBVH* -> Intersection chain, RayGen.cuh:63-172,
// BVHTraversal.cuh:14-134, Intersection.cu:4-36), cut into three resumable steps per lane:
//   NODE  : pop one wide node, slab-test its 8 children                (node_step_*)
//   TRI   : one Moller-Trumbore test + closest / any-hit bookkeeping   (tri_step)
//   SHADE : consume a finished ray (sky / material / sun / scatter / per-sample post) and emit the
//           lane's next ray, or the next sample of its pixel           (shade_step)
// A warp repeatedly votes (ballot + popc) for the phase most of its lanes are waiting in and executes
// only that step, so each step's instructions are issued for many lanes at once instead of for the
// few that happen to be in that branch (ncu on the monolithic loop: 10.5 of 32 lanes active,
// profiles/r01a_*). A lane's own sequence of steps -- and its RNG stream -- does not depend on the
// schedule, so results are identical to render_pixel_flat()/path_radiance() bit for bit.
//
// Two node formats feed NODE (chosen per scene, never per node):
//   Q8  : the builder's 80-byte compressed wide node (drt_types.h WideNode)        -- large scenes
//   F32 : the same boxes de-quantised once into 256-byte nodes of fp32 planes, laid out
//         [axis][lo|hi][child] so that "near plane for this ray's sign" is an address offset --
//         6 FFMA per child instead of 18 convert+select+FFMA (cache-resident scenes are issue-bound)
#pragma once
#include "drt_types.h"
#include "drt_rng.cuh"
#include "drt_trace.cuh"
#include "drt_shade.cuh"

namespace drt {

enum { PH_NODE = 0, PH_TRI = 1, PH_SHADE = 2, PH_DONE = 3 };
enum { NODE_FMT_Q8 = 0, NODE_FMT_F32 = 1 };

#define DRT_F32NODE_F4 16 // float4 per expanded node (256 bytes)

// Expanded node, 16 x float4:
//   [0]      u4 : child_base, tri_base, imask, 0
//   [1..2]   u4 : hit contribution of child 0..7 for octant 0: inner slot s -> 1 << (24+s); leaf -> unary << offset; empty -> 0
//   [3+4a+2s+h] : plane of axis a, side s (0 = lo, 1 = hi), children 4h..4h+3
//   [15]        : unused
DRT_HD void expand_wide_node(const WideNode& n, drt_f4* out)
{
	drt_u4 h;
	h.x = n.child_base; h.y = n.tri_base; h.z = n.imask; h.w = 0u;
	uint32_t contrib[8];
	float plane[3][2][8];
	const float p[3] = { n.px, n.py, n.pz };
	const float sc[3] = { as_float((uint32_t)n.ex << 23), as_float((uint32_t)n.ey << 23), as_float((uint32_t)n.ez << 23) };
	const uint8_t* q[3][2] = { { n.qlox, n.qhix }, { n.qloy, n.qhiy }, { n.qloz, n.qhiz } };
	for (int i = 0; i < 8; i++) {
		const uint32_t mt = n.meta[i];
		const bool inner = ((mt & (mt << 1)) & 0x10u) != 0u;
		contrib[i] = mt == 0u ? 0u : (inner ? (1u << (mt & 31u)) : ((mt >> 5) << (mt & 31u)));
		for (int a = 0; a < 3; a++)
			for (int s = 0; s < 2; s++)
				plane[a][s][i] = mt == 0u ? 0.f : p[a] + (float)q[a][s][i] * sc[a]; // exact: q * 2^e, then one rounding (outward-checked below)
	}
	// the builder verified base + q*scale brackets the child box in fp32 with this very expression (drt_build.cuh collapse_node)
	drt_u4* o = (drt_u4*)out;
	o[0] = h;
	drt_u4 c0, c1;
	c0.x = contrib[0]; c0.y = contrib[1]; c0.z = contrib[2]; c0.w = contrib[3];
	c1.x = contrib[4]; c1.y = contrib[5]; c1.z = contrib[6]; c1.w = contrib[7];
	o[1] = c0; o[2] = c1;
	for (int a = 0; a < 3; a++)
		for (int s = 0; s < 2; s++)
			for (int hh = 0; hh < 2; hh++) {
				drt_f4 v;
				v.x = plane[a][s][4 * hh]; v.y = plane[a][s][4 * hh + 1]; v.z = plane[a][s][4 * hh + 2]; v.w = plane[a][s][4 * hh + 3];
				out[3 + 4 * a + 2 * s + hh] = v;
			}
	drt_f4 z;
	z.x = z.y = z.z = z.w = 0.f;
	out[15] = z;
}

DRT_HD float fast_rcp(float x)
{
#if defined(__CUDA_ARCH__)
	float r;
	asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); // box tests carry 1e-6 relative slack; MUFU.RCP is within 1 ulp
	return r;
#else
	return 1.0f / x;
#endif
}

// Per-ray constants of the conservative slab test. t_near = plane*idn + bn, t_far = plane*idf + bf with
// idn = id*(1-1e-6), idf = id*(1+1e-6) and the biases pushed outward by a few ulp of |o*id| (the
// rounding error of the fused form is absolute in that magnitude, not relative to t).
struct RayBox {
	float idnx, idny, idnz, bnx, bny, bnz;
	float idfx, idfy, idfz, bfx, bfy, bfz;
	uint32_t oct_inv;
};

// `id3`, when given, receives the unscaled reciprocals (R.idn* = id * (1 - 1e-6), R.idf* = id * (1 + 1e-6)): kernels that park
// a ray outside registers keep these 3 words instead of the 6 scaled ones and re-derive both with the same two multiplies.
#define DRT_BOX_KN 0.999999f
#define DRT_BOX_KF 1.000001f
DRT_HD void ray_box_setup(V3 o, V3 d, RayBox& R, float* id3 = nullptr)
{
	const float tiny = 1e-20f;
	const float dx = fabsf(d.x) > tiny ? d.x : (d.x < 0.f ? -tiny : tiny);
	const float dy = fabsf(d.y) > tiny ? d.y : (d.y < 0.f ? -tiny : tiny);
	const float dz = fabsf(d.z) > tiny ? d.z : (d.z < 0.f ? -tiny : tiny);
	const float idx = fast_rcp(dx), idy = fast_rcp(dy), idz = fast_rcp(dz);
	R.oct_inv = (d.x >= 0.f ? 4u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 1u : 0u);
	const float kn = DRT_BOX_KN, kf = DRT_BOX_KF, ka = 4.8e-7f;
	if (id3) { id3[0] = idx; id3[1] = idy; id3[2] = idz; }
	R.idnx = idx * kn; R.idny = idy * kn; R.idnz = idz * kn;
	R.idfx = idx * kf; R.idfy = idy * kf; R.idfz = idz * kf;
	const float ex = fabsf(o.x * idx) * ka, ey = fabsf(o.y * idy) * ka, ez = fabsf(o.z * idz) * ka;
	R.bnx = -(o.x * R.idnx) - ex; R.bny = -(o.y * R.idny) - ey; R.bnz = -(o.z * R.idnz) - ez;
	R.bfx = -(o.x * R.idfx) + ex; R.bfy = -(o.y * R.idfy) + ey; R.bfz = -(o.z * R.idfz) + ez;
}

// Traversal state of one lane between steps.
struct LaneTrav {
	uint32_t ng_base, ng_bits;   // current node group: child_base | (hit inner slots << 24 | imask)
	uint32_t tri_base, tri_bits; // pending triangles of the last node
	int sp;
	Hit hit; // only prim and t are carried between steps: u, v (and the id of the current winner, for exact-t ties)
	         // are recomputed from the triangle where they are needed (same operations -> same bits)
};

#ifndef DRT_PSTACK
#define DRT_PSTACK 32
#endif

// Traversal stack of one lane. The step functions below are templates over the stack type so that the same code runs with
// the stack in per-thread local memory (LaneStack, the default), in shared memory (drt_regroup.cuh) or wherever a kernel
// keeps it; a stack type provides put(i, x, y) / get(i, x, y) for 0 <= i < DRT_PSTACK.
struct LaneStack {
	uint32_t x[DRT_PSTACK], y[DRT_PSTACK];
	DRT_HD void put(int i, uint32_t vx, uint32_t vy) { x[i] = vx; y[i] = vy; }
	DRT_HD void get(int i, uint32_t& vx, uint32_t& vy) const { vx = x[i]; vy = y[i]; }
};

// The same stack with its first DRT_SSTACK entries in SHARED memory ([entry][thread] words, conflict-free) and only the rest in
// local memory: measured depths are <= 3 (room.glb) ... 7 (soup-10M), so the local part is never touched on the bundled scenes.
// ncu on the all-local stack: 22-30 % of the kernel's L1 sector traffic was local loads / stores (profiles/r02a_*).
#ifndef DRT_SSTACK
#define DRT_SSTACK 8
#endif
#define DRT_SSTACK_THREADS 128
#define DRT_SSTACK_WORDS (2 * DRT_SSTACK * DRT_SSTACK_THREADS)
struct SharedLaneStack {
	uint32_t* s; // shared-memory words of this thread: s[(2 * i) * DRT_SSTACK_THREADS], s[(2 * i + 1) * DRT_SSTACK_THREADS]
	uint32_t lx[DRT_PSTACK - DRT_SSTACK], ly[DRT_PSTACK - DRT_SSTACK];
	DRT_HD void put(int i, uint32_t vx, uint32_t vy)
	{
		if (i < DRT_SSTACK) { s[(2 * i) * DRT_SSTACK_THREADS] = vx; s[(2 * i + 1) * DRT_SSTACK_THREADS] = vy; }
		else { lx[i - DRT_SSTACK] = vx; ly[i - DRT_SSTACK] = vy; }
	}
	DRT_HD void get(int i, uint32_t& vx, uint32_t& vy) const
	{
		if (i < DRT_SSTACK) { vx = s[(2 * i) * DRT_SSTACK_THREADS]; vy = s[(2 * i + 1) * DRT_SSTACK_THREADS]; }
		else { vx = lx[i - DRT_SSTACK]; vy = ly[i - DRT_SSTACK]; }
	}
};

DRT_HD void trav_begin(LaneTrav& L, float tmax)
{
	L.ng_base = 0u; L.ng_bits = 0x80000000u; // root: one pending inner slot, imask 0 -> relative index 0
	L.tri_base = 0u; L.tri_bits = 0u;
	L.sp = 0;
	L.hit.prim = -1; L.hit.t = tmax; L.hit.u = 0.f; L.hit.v = 0.f;
}

// Phase of a lane after a NODE or TRI step (pops the stack when the current group is exhausted).
template <class STACK>
DRT_HD int trav_advance(LaneTrav& L, STACK& S)
{
	const bool has_tri = L.tri_bits != 0u, has_node = L.ng_bits > 0x00ffffffu;
	int phase = has_tri ? PH_TRI : PH_NODE;
	if (!has_tri && !has_node) {
		if (L.sp == 0) phase = PH_SHADE;
		else {
			// entries beyond the stack were dropped when pushed (counted via max_stack): skip them in one step
			const int sp = L.sp - 1;
			L.sp = sp < DRT_PSTACK - 1 ? sp : DRT_PSTACK - 1;
			S.get(L.sp, L.ng_base, L.ng_bits);
		}
	}
	return phase;
}

// Picks the nearest pending inner child of the current group (octant order) and returns its node index.
template <bool STATS, class STACK>
DRT_HD uint32_t trav_pop_child(LaneTrav& L, STACK& S, uint32_t oct_inv, TraceStats* st)
{
	const int bit = hi_bit(L.ng_bits);
	L.ng_bits &= ~(1u << bit);
	const uint32_t slot = (uint32_t)(bit - 24) ^ oct_inv;
	const uint32_t rel = (uint32_t)pop_count(L.ng_bits & 0xffu & ((1u << slot) - 1u));
	const uint32_t node_index = L.ng_base + rel;
	if (L.ng_bits > 0x00ffffffu) {
		if (L.sp < DRT_PSTACK) S.put(L.sp, L.ng_base, L.ng_bits);
		++L.sp;
		if (STATS && st && (uint32_t)L.sp > st->max_stack) st->max_stack = (uint32_t)L.sp;
	}
	if (STATS && st) st->nodes++;
	return node_index;
}

// bits 24..31 of `m` hold child-slot hits in slot order; move bit s to bit s ^ oct_inv
DRT_HD uint32_t permute_slots(uint32_t m, uint32_t oct_inv)
{
	uint32_t b = m >> 24;
	if (oct_inv & 1u) b = ((b & 0x55u) << 1) | ((b >> 1) & 0x55u);
	if (oct_inv & 2u) b = ((b & 0x33u) << 2) | ((b >> 2) & 0x33u);
	if (oct_inv & 4u) b = ((b & 0x0fu) << 4) | (b >> 4);
	return (m & 0x00ffffffu) | (b << 24);
}

// hitmask |= C when tn <= tf. The contribution words of one node have disjoint bits, so the OR is an ADD; on the
// device it is issued as a predicated integer multiply-add (FMA pipe) because the NODE step is bound by the half-rate
// ALU pipe (compares, min/max, selects): ncu r01c, ALU 65 % vs FMA 26 %.
#if defined(__CUDA_ARCH__)
#define DRT_ACC_HIT(H, TN, TF, C) \
	asm("{ .reg .pred p; setp.le.f32 p, %1, %2; @p mad.lo.u32 %0, %3, 1, %0; }" : "+r"(H) : "f"(TN), "f"(TF), "r"(C));
#else
#define DRT_ACC_HIT(H, TN, TF, C) if ((TN) <= (TF)) (H) |= (C);
#endif
// Table form of permute_slots for the device: lut[(oct_inv << 8) | b] = b with bit s moved to bit s ^ oct_inv
// (2 KB of shared memory filled per block by fill_permute_lut). One PRMT + one LDS.U8 instead of ~12 ALU instructions.
#define DRT_PERMUTE_LUT_BYTES 2048
DRT_HD uint32_t permuted_child_bits(uint32_t m, uint32_t oct_inv, const uint8_t* lut)
{
#if defined(__CUDA_ARCH__)
	if (lut) return lut[__byte_perm(m, oct_inv, 0x5543)]; // byte 3 of m | oct_inv << 8 (oct_inv <= 7: its upper bytes are 0)
#endif
	(void)lut;
	return permute_slots(m, oct_inv) >> 24;
}

#define DRT_F32_CHILD(I, NX, NY, NZ, FX, FY, FZ, C)                                                  \
	{                                                                                                \
		const float tn = fmaxf(fmax3(fmaf(NX, R.idnx, R.bnx), fmaf(NY, R.idny, R.bny), fmaf(NZ, R.idnz, R.bnz)), 0.f); \
		const float tf = fminf(fmin3(fmaf(FX, R.idfx, R.bfx), fmaf(FY, R.idfy, R.bfy), fmaf(FZ, R.idfz, R.bfz)), tlim); \
		DRT_ACC_HIT(hitmask, tn, tf, C)                                                              \
	}

// NODE step on expanded nodes. `nodes` may point to global or shared memory.
template <bool STATS, class STACK>
DRT_HD void node_step_f32(const drt_f4* nodes, LaneTrav& L, STACK& S, const RayBox& R, TraceStats* st, const uint8_t* lut = nullptr)
{
	const uint32_t node_index = trav_pop_child<STATS>(L, S, R.oct_inv, st);
	const drt_f4* np = nodes + (size_t)DRT_F32NODE_F4 * node_index;
	const drt_u4* nu = (const drt_u4*)np;
	const drt_u4 h = ld_u4(nu);
	const float tlim = L.hit.t;
	// near side of axis a: lo if the ray travels towards + (oct_inv bit set), else hi
	const int sx = (R.oct_inv & 4u) ? 0 : 2, sy = (R.oct_inv & 2u) ? 0 : 2, sz = (R.oct_inv & 1u) ? 0 : 2;
	const drt_f4* px = np + 3, * py = np + 7, * pz = np + 11;
	uint32_t hitmask = 0u;
	{
		const drt_u4 c = ld_u4(nu + 1);
		const drt_f4 nx = ld_f4(px + sx), ny = ld_f4(py + sy), nz = ld_f4(pz + sz);
		const drt_f4 fx = ld_f4(px + (2 - sx)), fy = ld_f4(py + (2 - sy)), fz = ld_f4(pz + (2 - sz));
		DRT_F32_CHILD(0, nx.x, ny.x, nz.x, fx.x, fy.x, fz.x, c.x)
		DRT_F32_CHILD(1, nx.y, ny.y, nz.y, fx.y, fy.y, fz.y, c.y)
		DRT_F32_CHILD(2, nx.z, ny.z, nz.z, fx.z, fy.z, fz.z, c.z)
		DRT_F32_CHILD(3, nx.w, ny.w, nz.w, fx.w, fy.w, fz.w, c.w)
	}
	{
		const drt_u4 c = ld_u4(nu + 2);
		const drt_f4 nx = ld_f4(px + sx + 1), ny = ld_f4(py + sy + 1), nz = ld_f4(pz + sz + 1);
		const drt_f4 fx = ld_f4(px + (2 - sx) + 1), fy = ld_f4(py + (2 - sy) + 1), fz = ld_f4(pz + (2 - sz) + 1);
		DRT_F32_CHILD(4, nx.x, ny.x, nz.x, fx.x, fy.x, fz.x, c.x)
		DRT_F32_CHILD(5, nx.y, ny.y, nz.y, fx.y, fy.y, fz.y, c.y)
		DRT_F32_CHILD(6, nx.z, ny.z, nz.z, fx.z, fy.z, fz.z, c.z)
		DRT_F32_CHILD(7, nx.w, ny.w, nz.w, fx.w, fy.w, fz.w, c.w)
	}
	L.ng_base = h.x;
	L.ng_bits = (permuted_child_bits(hitmask, R.oct_inv, lut) << 24) | (h.z & 0xffu);
	L.tri_base = h.y;
	L.tri_bits = hitmask & 0x00ffffffu;
}

// Byte `i` of `w` as a float, exactly. On the device: one PRMT builds 2^23 + q in the mantissa and one FADD
// removes the 2^23 (ALU + FMA pipes) instead of I2F.U8, which runs on the quarter-rate XU pipe (ncu r01a: XU 50 %).
DRT_HD float byte_to_float(uint32_t w, int i)
{
#if defined(__CUDA_ARCH__)
	return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540u | (uint32_t)i)) - 8388608.0f;
#else
	return (float)((w >> (8 * i)) & 0xffu);
#endif
}

// NODE step on the builder's compressed nodes (same box arithmetic as trace_ray_mode, drt_trace.cuh).
template <bool STATS, class STACK>
DRT_HD void node_step_q8(const drt_u4* nodes, LaneTrav& L, STACK& S, const RayBox& R, TraceStats* st, const uint8_t* lut = nullptr)
{
	const uint32_t node_index = trav_pop_child<STATS>(L, S, R.oct_inv, st);
	const drt_u4* np = nodes + 5 * (size_t)node_index;
	const drt_u4 n0 = ld_u4(np), n1 = ld_u4(np + 1), n2 = ld_u4(np + 2), n3 = ld_u4(np + 3), n4 = ld_u4(np + 4);
	const float px = as_float(n0.x), py = as_float(n0.y), pz = as_float(n0.z);
	const float ex = as_float((n0.w & 0xffu) << 23), ey = as_float(((n0.w >> 8) & 0xffu) << 23), ez = as_float(((n0.w >> 16) & 0xffu) << 23);
	const uint32_t imask = n0.w >> 24;
	// t = q * (scale*id) + (p*id + b): scale is a power of two, so q*scale*id carries one rounding
	const float snx = ex * R.idnx, sny = ey * R.idny, snz = ez * R.idnz;
	const float sfx = ex * R.idfx, sfy = ey * R.idfy, sfz = ez * R.idfz;
	const float ka = 2.4e-7f;
	const float enx = (fabsf(px * R.idnx) + 255.f * fabsf(snx)) * ka, eny = (fabsf(py * R.idny) + 255.f * fabsf(sny)) * ka, enz = (fabsf(pz * R.idnz) + 255.f * fabsf(snz)) * ka;
	const float bnx = fmaf(px, R.idnx, R.bnx) - enx, bny = fmaf(py, R.idny, R.bny) - eny, bnz = fmaf(pz, R.idnz, R.bnz) - enz;
	const float bfx = fmaf(px, R.idfx, R.bfx) + enx, bfy = fmaf(py, R.idfy, R.bfy) + eny, bfz = fmaf(pz, R.idfz, R.bfz) + enz;
	const bool posx = (R.oct_inv & 4u) != 0u, posy = (R.oct_inv & 2u) != 0u, posz = (R.oct_inv & 1u) != 0u;
	// near / far plane words per axis, selected once per node by the ray's sign
	const uint32_t qnx[2] = { posx ? n2.x : n3.z, posx ? n2.y : n3.w }, qfx[2] = { posx ? n3.z : n2.x, posx ? n3.w : n2.y };
	const uint32_t qny[2] = { posy ? n2.z : n4.x, posy ? n2.w : n4.y }, qfy[2] = { posy ? n4.x : n2.z, posy ? n4.y : n2.w };
	const uint32_t qnz[2] = { posz ? n3.x : n4.z, posz ? n3.y : n4.w }, qfz[2] = { posz ? n4.z : n3.x, posz ? n4.w : n3.y };
	const uint32_t meta[2] = { n1.z, n1.w };
	const float tlim = L.hit.t;
	uint32_t hitmask = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
	for (int i = 0; i < 8; i++) {
		const int w = i >> 2, sh = (i & 3) * 8;
		const uint32_t mt = (meta[w] >> sh) & 0xffu;
		const int bi = i & 3;
		const float t0x = fmaf(byte_to_float(qnx[w], bi), snx, bnx), t1x = fmaf(byte_to_float(qfx[w], bi), sfx, bfx);
		const float t0y = fmaf(byte_to_float(qny[w], bi), sny, bny), t1y = fmaf(byte_to_float(qfy[w], bi), sfy, bfy);
		const float t0z = fmaf(byte_to_float(qnz[w], bi), snz, bnz), t1z = fmaf(byte_to_float(qfz[w], bi), sfz, bfz);
		const float tn = fmaxf(fmax3(t0x, t0y, t0z), 0.f);
		const float tf = fminf(fmin3(t1x, t1y, t1z), tlim);
		if (mt != 0u && tn <= tf) {
			const bool inner = ((mt & (mt << 1)) & 0x10u) != 0u;
			hitmask |= (inner ? 1u : (mt >> 5)) << (mt & 31u);
		}
	}
	L.ng_base = n1.x;
	L.ng_bits = (permuted_child_bits(hitmask, R.oct_inv, lut) << 24) | imask;
	L.tri_base = n1.y;
	L.tri_bits = hitmask & 0x00ffffffu;
}

// TRI step: tests the lowest pending triangle. Returns true when an any-hit ray is finished (occluded).
template <bool STRICT, bool STATS>
DRT_HD bool tri_step(const SceneView& sc, const drt_f4* tris, LaneTrav& L, V3 o, V3 d, bool anyhit, TraceStats* st)
{
	typedef M<STRICT> m;
	const int k = lo_bit(L.tri_bits);
	L.tri_bits &= L.tri_bits - 1u;
	const int slot = (int)(L.tri_base + (uint32_t)k);
	const drt_f4* tp = tris + 3 * (size_t)slot;
	const drt_f4 r0 = ld_f4(tp), r1 = ld_f4(tp + 1), r2 = ld_f4(tp + 2);
	if (STATS && st) st->tris++;
	const V3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
	// Intersection.cu:8-33, the reference's operation order
	const V3 pvec = m::cross(d, e2);
	const float det = m::dot(e1, pvec);
	if (det > -DRT_TRIANGLE_EPSILON && det < DRT_TRIANGLE_EPSILON) return false;
	const float inv_det = m::div(1.0f, det);
	const V3 tvec = m::sub3(o, v0);
	const float u = m::mul(inv_det, m::dot(tvec, pvec));
	if (u < 0.0f || u > 1.0f) return false;
	const V3 qvec = m::cross(tvec, e1);
	const float v = m::mul(inv_det, m::dot(d, qvec));
	if (v < 0.0f || m::add(u, v) > 1.0f) return false;
	const float t = m::mul(inv_det, m::dot(e2, qvec));
	if (!(t > DRT_TRIANGLE_EPSILON)) return false;
	const uint32_t matw = as_uint(r1.w);
	if (anyhit) {
		if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, u, v, st)) return false;
		L.hit.prim = slot; L.hit.t = t;
		return true;
	}
	if (!(t < L.hit.t)) {
		if (!(t == L.hit.t) || L.hit.prim < 0) return false;
		if (st) st->ties++;
		const uint32_t id = as_uint(r0.w), best_id = as_uint(ld_f4(tris + 3 * (size_t)L.hit.prim).w);
		if (!(id < best_id)) return false;
	}
	if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, u, v, st)) return false;
	L.hit.prim = slot; L.hit.t = t;
	return false;
}

// ---- watertight ray / triangle test (production mode only, DRT_FLAG_WATERTIGHT; north_star subsystem 3) ----
// The reference's Moller-Trumbore (Intersection.cu:4-36, what STRICT mode must keep) has two leaks: an ABSOLUTE epsilon on the
// determinant (|det| < 1e-6, Common/physical_units.hpp:12, Intersection.cu:13), which drops every triangle that is small or
// far away, and per-triangle edge functions, which can round to "outside" on BOTH sides of a shared edge. This is the test of
// Woop, Benthin, Wald, "Watertight Ray/Triangle Intersection" (JCGT 2013): the ray is sheared onto its major axis, the three
// edge functions are differences of two separately rounded products of vertex coordinates that are computed the same way
// whichever triangle the edge belongs to (so a ray cannot pass between two triangles that share the edge: one of them gets
// U, V, W of one sign or zero), with a double-precision recomputation when an edge function is exactly zero.
struct RayShear {
	int kx, ky, kz;
	float sx, sy, sz;
};
DRT_HD float v3_comp(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
DRT_HD RayShear ray_shear_setup(V3 d)
{
	RayShear r;
	const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
	r.kz = ax > ay ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
	r.kx = r.kz == 2 ? 0 : r.kz + 1;
	r.ky = r.kx == 2 ? 0 : r.kx + 1;
	const float dz = v3_comp(d, r.kz);
	if (dz < 0.f) { const int t = r.kx; r.kx = r.ky; r.ky = t; } // keep the winding
	r.sx = v3_comp(d, r.kx) / dz; r.sy = v3_comp(d, r.ky) / dz; r.sz = 1.0f / dz;
	return r;
}
DRT_HD float wt_mul(float a, float b)
{
#if defined(__CUDA_ARCH__)
	return __fmul_rn(a, b); // never contracted into an FMA: a*b - c*d must be the exact negative of c*d - a*b
#else
	return a * b;
#endif
}
DRT_HD float wt_sub(float a, float b)
{
#if defined(__CUDA_ARCH__)
	return __fsub_rn(a, b);
#else
	return a - b;
#endif
}

// TRI step with the watertight predicate on the vertex records (sc.tris_v). Same bookkeeping as tri_step: closest hit with
// exact-t ties to the lowest canonical id, any-hit ends at the first accepted triangle, alpha cut-out on the barycentrics
// (u = weight of v1, v = weight of v2, like Intersection.cu:31). Keeps the reference's t > 1e-6 (self-hit guard), drops its
// determinant epsilon.
template <bool STATS>
DRT_HD bool tri_step_wt(const SceneView& sc, LaneTrav& L, V3 o, V3 d, bool anyhit, TraceStats* st)
{
	const int k = lo_bit(L.tri_bits);
	L.tri_bits &= L.tri_bits - 1u;
	const int slot = (int)(L.tri_base + (uint32_t)k);
	const drt_f4* tp = sc.tris_v + 3 * (size_t)slot;
	const drt_f4 r0 = ld_f4(tp), r1 = ld_f4(tp + 1), r2 = ld_f4(tp + 2);
	if (STATS && st) st->tris++;
	const RayShear R = ray_shear_setup(d);
	const V3 A = mk3(r0.x - o.x, r0.y - o.y, r0.z - o.z), B = mk3(r1.x - o.x, r1.y - o.y, r1.z - o.z), C = mk3(r2.x - o.x, r2.y - o.y, r2.z - o.z);
	const float Akz = v3_comp(A, R.kz), Bkz = v3_comp(B, R.kz), Ckz = v3_comp(C, R.kz);
	const float Ax = wt_sub(v3_comp(A, R.kx), wt_mul(R.sx, Akz)), Ay = wt_sub(v3_comp(A, R.ky), wt_mul(R.sy, Akz));
	const float Bx = wt_sub(v3_comp(B, R.kx), wt_mul(R.sx, Bkz)), By = wt_sub(v3_comp(B, R.ky), wt_mul(R.sy, Bkz));
	const float Cx = wt_sub(v3_comp(C, R.kx), wt_mul(R.sx, Ckz)), Cy = wt_sub(v3_comp(C, R.ky), wt_mul(R.sy, Ckz));
	float U = wt_sub(wt_mul(Cx, By), wt_mul(Cy, Bx));
	float V = wt_sub(wt_mul(Ax, Cy), wt_mul(Ay, Cx));
	float W = wt_sub(wt_mul(Bx, Ay), wt_mul(By, Ax));
	if (U == 0.f || V == 0.f || W == 0.f) { // on an edge in fp32: decide in double (the products are exact there)
		U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
		V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
		W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
	}
	if ((U < 0.f || V < 0.f || W < 0.f) && (U > 0.f || V > 0.f || W > 0.f)) return false;
	const float det = U + V + W;
	if (det == 0.f) return false;
	const float T = U * (R.sz * Akz) + V * (R.sz * Bkz) + W * (R.sz * Ckz);
	const float rdet = 1.0f / det;
	const float t = T * rdet;
	if (!(t > DRT_TRIANGLE_EPSILON)) return false;
	const float u = V * rdet, v = W * rdet;
	const uint32_t matw = as_uint(r1.w);
	if (anyhit) {
		if ((matw & 0x80000000u) && !alpha_accept<false>(sc, slot, matw & 0x7fffffffu, u, v, st)) return false;
		L.hit.prim = slot; L.hit.t = t;
		return true;
	}
	if (!(t < L.hit.t)) {
		if (!(t == L.hit.t) || L.hit.prim < 0) return false;
		if (st) st->ties++;
		const uint32_t id = as_uint(r0.w), best_id = as_uint(ld_f4(sc.tris_v + 3 * (size_t)L.hit.prim).w);
		if (!(id < best_id)) return false;
	}
	if ((matw & 0x80000000u) && !alpha_accept<false>(sc, slot, matw & 0x7fffffffu, u, v, st)) return false;
	L.hit.prim = slot; L.hit.t = t;
	return false;
}

// Geometric part of one Moller-Trumbore test, computed unconditionally (Intersection.cu:8-33, the reference's operation
// order; a rejected triangle may produce inf / NaN in t, u, v, which nobody reads).
struct TriGeo {
	float t, u, v;
	bool ok;
};
template <bool STRICT>
DRT_HD TriGeo tri_geometry(const drt_f4 r0, const drt_f4 r1, const drt_f4 r2, V3 o, V3 d)
{
	typedef M<STRICT> m;
	const V3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
	const V3 pvec = m::cross(d, e2);
	const float det = m::dot(e1, pvec);
	const float inv_det = m::div(1.0f, det);
	const V3 tvec = m::sub3(o, v0);
	TriGeo g;
	g.u = m::mul(inv_det, m::dot(tvec, pvec));
	const V3 qvec = m::cross(tvec, e1);
	g.v = m::mul(inv_det, m::dot(d, qvec));
	g.t = m::mul(inv_det, m::dot(e2, qvec));
	g.ok = !(det > -DRT_TRIANGLE_EPSILON && det < DRT_TRIANGLE_EPSILON) && !(g.u < 0.0f || g.u > 1.0f) && !(g.v < 0.0f || m::add(g.u, g.v) > 1.0f) &&
		(g.t > DRT_TRIANGLE_EPSILON);
	return g;
}

// Closest / any-hit bookkeeping of one geometrically accepted candidate (the tail of tri_step). Returns true when an any-hit
// ray is finished.
template <bool STRICT>
DRT_HD bool tri_apply(const SceneView& sc, const drt_f4* tris, LaneTrav& L, int slot, const drt_f4 r0, const drt_f4 r1, const TriGeo g,
	bool anyhit, TraceStats* st)
{
	if (!g.ok) return false;
	const uint32_t matw = as_uint(r1.w);
	if (anyhit) {
		if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, g.u, g.v, st)) return false;
		L.hit.prim = slot; L.hit.t = g.t;
		return true;
	}
	if (!(g.t < L.hit.t)) {
		if (!(g.t == L.hit.t) || L.hit.prim < 0) return false;
		if (st) st->ties++;
		const uint32_t id = as_uint(r0.w), best_id = as_uint(ld_f4(tris + 3 * (size_t)L.hit.prim).w);
		if (!(id < best_id)) return false;
	}
	if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, g.u, g.v, st)) return false;
	L.hit.prim = slot; L.hit.t = g.t;
	return false;
}

// TRI step on up to TWO pending triangles: their tests are independent instruction streams (the step is bound by dependent
// fixed-latency chains), and a lane with two or more pending triangles needs half the voted steps. The candidates are
// applied in the order tri_step would have met them, the second only if the first did not finish an any-hit ray.
template <bool STRICT, bool STATS>
DRT_HD bool tri_step2(const SceneView& sc, const drt_f4* tris, LaneTrav& L, V3 o, V3 d, bool anyhit, TraceStats* st)
{
	const int k0 = lo_bit(L.tri_bits);
	L.tri_bits &= L.tri_bits - 1u;
	const bool two = L.tri_bits != 0u;
	const int k1 = two ? lo_bit(L.tri_bits) : k0;
	const int slot0 = (int)(L.tri_base + (uint32_t)k0), slot1 = (int)(L.tri_base + (uint32_t)k1);
	const drt_f4* tp0 = tris + 3 * (size_t)slot0;
	const drt_f4* tp1 = tris + 3 * (size_t)slot1;
	const drt_f4 a0 = ld_f4(tp0), a1 = ld_f4(tp0 + 1), a2 = ld_f4(tp0 + 2);
	const drt_f4 b0 = ld_f4(tp1), b1 = ld_f4(tp1 + 1), b2 = ld_f4(tp1 + 2);
	const TriGeo ga = tri_geometry<STRICT>(a0, a1, a2, o, d);
	const TriGeo gb = tri_geometry<STRICT>(b0, b1, b2, o, d);
	if (STATS && st) st->tris++;
	if (tri_apply<STRICT>(sc, tris, L, slot0, a0, a1, ga, anyhit, st)) return true;
	if (!two) return false;
	L.tri_bits &= L.tri_bits - 1u;
	if (STATS && st) st->tris++;
	return tri_apply<STRICT>(sc, tris, L, slot1, b0, b1, gb, anyhit, st);
}

// Barycentrics of the accepted hit, recomputed with the operations of tri_step (Intersection.cu:8-27) when a
// textured material needs them: cheaper than carrying u, v through every step of every lane.
template <bool STRICT>
DRT_HD V2 hit_barycentrics(const drt_f4* tris, int slot, V3 o, V3 d)
{
	typedef M<STRICT> m;
	const drt_f4* tp = tris + 3 * (size_t)slot;
	const drt_f4 r0 = ld_f4(tp), r1 = ld_f4(tp + 1), r2 = ld_f4(tp + 2);
	const V3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
	const V3 pvec = m::cross(d, e2);
	const float det = m::dot(e1, pvec);
	const float inv_det = m::div(1.0f, det);
	const V3 tvec = m::sub3(o, v0);
	const float u = m::mul(inv_det, m::dot(tvec, pvec));
	const V3 qvec = m::cross(tvec, e1);
	const float v = m::mul(inv_det, m::dot(d, qvec));
	return mk2(u, v);
}

// Path state of one lane between steps (everything RayGen keeps in locals across its bounce loop).
struct LanePath {
	V3 o, d;              // current ray
	V3 light, throughput;
	V3 fn;                // face-forwarded normal of the last hit (needed again after the sun shadow ray)
	V3 acc;               // running per-pixel sum (RenderKernel.cu:29)
	uint32_t seed;
	uint32_t k;           // sample of this launch the lane is working on
	int bounce;
	bool shadow;          // current ray is the sun shadow ray (RayTest) of the last hit
};

template <bool STRICT>
DRT_HD void path_begin_sample(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, LanePath& S)
{
	S.light = mk3(0.f, 0.f, 0.f); S.throughput = mk3(1.f, 1.f, 1.f);
	S.bounce = 0; S.shadow = false;
	S.seed = (x + y * (uint32_t)P.width) * (P.frame_first + S.k * P.frame_stride); // RayGen.cuh:74-75
	const RayOD r = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, S.seed, rtl);
	S.o = r.o; S.d = r.d;
}

// SHADE step: consumes the finished ray of this lane. Returns false when the lane has no more work
// (all samples of its pixel are accumulated); otherwise S.o / S.d hold the next ray to trace.
// Block for block this is the body of render_pixel_flat() (drt_shade.cuh), i.e. RayGen.cuh:88-169.
// shade_core: everything of the step up to the decision "this sample is complete" (returned); S.o / S.d hold the next ray otherwise.
template <bool STRICT, bool STATS>
DRT_HD bool shade_core(const RenderParams& P, bool rtl, const Hit& hit, LanePath& S, PathCounters& pc, TraceStats* st)
{
	typedef M<STRICT> m;
	const SceneView& sc = P.scene;
	const ShadeParams& SP = P.shade;
	const bool found = hit.prim >= 0;
	bool path_done = false, scatter = false;
	if (S.shadow) {
		pc.shadow++;
		if (!found) S.light = m::add3(S.light, m::mul3(SP.suncol, S.throughput)); // RayGen.cuh:126-127
		S.shadow = false;
		scatter = true;
	}
	else {
		pc.closest++;
		S.seed += (uint32_t)S.bounce; // RayGen.cuh:91
		if (!found) {
			const V3 sky = sky_model<STRICT>(S.d, SP.sky_color);
			S.light = m::add3(S.light, m::scale(m::mul3(sky, S.throughput), SP.sky_intensity));
			path_done = true;
		}
		else {
			const drt_f4 r1 = ld_f4(sc.tris + 3 * (size_t)hit.prim + 1);
			const uint32_t mat = as_uint(r1.w) & 0x7fffffffu;
			const drt_f4 mrec = ld_f4(sc.materials + mat);
			const int tex = (int)as_uint(mrec.w);
			if (tex < 0) S.throughput = m::mul3(S.throughput, mk3(mrec.x, mrec.y, mrec.z));
			else {
				const V2 bc = hit_barycentrics<STRICT>(sc.tris, hit.prim, S.o, S.d);
			const V2 tuv = interp_uv<STRICT>(sc, hit.prim, bc.x, bc.y);
				S.throughput = m::mul3(S.throughput, texture_get_pixel<STRICT>(sc, tex, tuv, st));
			}
			const drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)hit.prim);
			V3 fn = mk3(s0.x, s0.y, s0.z);
			const V3 pos = m::add3(S.o, m::scale(S.d, hit.t));
			// ClosestHit.cuh:17-25 flips against normalize(d); the sign does not need the normalisation (production mode)
			if ((STRICT ? m::dot(fn, m::normalize(S.d)) : m::dot(fn, S.d)) > 0.f) fn = m::scale(fn, -1.f);
			S.fn = fn;
			S.o = m::add3(pos, m::scale(fn, 0.001f));                          // RayGen.cuh:121
			if (SP.enable_sunlight) {
				S.d = m::add3(SP.sunpos, m::scale(random_unit_vec3<STRICT>(S.seed, rtl), 1.5f)); // RayGen.cuh:124-125
				S.shadow = true;
			}
			else scatter = true;
		}
	}
	if (scatter) {
		S.d = m::add3(S.fn, random_unit_sphere_vec3<STRICT>(S.seed, rtl)); // RayGen.cuh:133-134
		S.bounce++;
		if (S.bounce > SP.bounce_limit) path_done = true;
	}
	return path_done;
}

template <bool STRICT, bool STATS>
DRT_HD bool shade_step(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, const Hit& hit, LanePath& S,
	PathCounters& pc, TraceStats* st)
{
	typedef M<STRICT> m;
	if (shade_core<STRICT, STATS>(P, rtl, hit, S, pc, st)) {
		S.acc = m::add3(S.acc, post_process<STRICT>(P, S.light));
		S.k++;
		if (S.k >= P.n_samples) return false;
		path_begin_sample<STRICT>(P, x, y, rtl, S);
	}
	return true;
}

// One lane, serial schedule: the CPU logic tests (tests/hostemu) drive the very same steps with it.
template <bool STRICT, bool STATS, int FMT, bool TRI2 = false>
DRT_HD V3 render_pixel_steps(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, V3 acc, PathCounters& pc, TraceStats* st)
{
	if (P.n_samples == 0) return acc;
	LanePath S;
	LaneTrav L;
	LaneStack K;
	RayBox R;
	S.acc = acc; S.k = 0; S.fn = mk3(0.f, 0.f, 0.f);
	path_begin_sample<STRICT>(P, x, y, rtl, S);
	ray_box_setup(S.o, S.d, R);
	trav_begin(L, FLT_MAX);
	int phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
	while (phase != PH_DONE) {
		if (phase == PH_NODE) {
			if (FMT == NODE_FMT_F32) node_step_f32<STATS>(P.scene.nodes_f32, L, K, R, st);
			else node_step_q8<STATS>(P.scene.nodes, L, K, R, st);
			phase = trav_advance(L, K);
		}
		else if (phase == PH_TRI) {
			if (TRI2 ? tri_step2<STRICT, STATS>(P.scene, P.scene.tris, L, S.o, S.d, S.shadow, st)
			         : tri_step<STRICT, STATS>(P.scene, P.scene.tris, L, S.o, S.d, S.shadow, st)) phase = PH_SHADE;
			else phase = trav_advance(L, K);
		}
		else {
			if (!shade_step<STRICT, STATS>(P, x, y, rtl, L.hit, S, pc, st)) phase = PH_DONE;
			else {
				ray_box_setup(S.o, S.d, R);
				trav_begin(L, FLT_MAX);
				phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
			}
		}
	}
	return S.acc;
}

#if defined(__CUDACC__)
// Scheduling weights of the vote (launch parameters so that they can be swept without recompiling).
struct PhasedTuning {
	int stick;  // a step is re-issued without a full vote while at least this many lanes want it again (33 = never)
	int refill; // pixel-queue kernel: idle lanes take new pixels once at least this many of them wait (1..32)
	int pool;   // sample-pool kernel: the last `pool` samples of every pixel of a tile are handed out dynamically (0 = none)
};

// The permutation table, constant-initialised in device memory (2 KB, L2/L1 resident), and the per-block copy to
// shared memory: one 16-byte load + store per thread (all threads of the block call it, then __syncthreads()).
struct alignas(16) PermuteLut {
	uint8_t v[DRT_PERMUTE_LUT_BYTES];
	constexpr PermuteLut() : v()
	{
		for (int i = 0; i < DRT_PERMUTE_LUT_BYTES; i++) {
			unsigned b = (unsigned)(i & 255), o = (unsigned)(i >> 8);
			if (o & 1u) b = ((b & 0x55u) << 1) | ((b >> 1) & 0x55u);
			if (o & 2u) b = ((b & 0x33u) << 2) | ((b >> 2) & 0x33u);
			if (o & 4u) b = ((b & 0x0fu) << 4) | (b >> 4);
			v[i] = (uint8_t)b;
		}
	}
};
__device__ const PermuteLut g_permute_lut{};

// Thread i < 128 moves 16 bytes (smaller blocks loop over the rest at the store). The load is issued first and the store + barrier come
// after the first camera ray has been set up, so that a one-sample launch does not wait for the table.
__device__ __forceinline__ uint4 permute_lut_load()
{
	return threadIdx.x < DRT_PERMUTE_LUT_BYTES / 16 ? __ldg((const uint4*)g_permute_lut.v + threadIdx.x) : make_uint4(0u, 0u, 0u, 0u);
}
__device__ __forceinline__ void permute_lut_store(uint8_t* lut, uint4 v)
{
	if (threadIdx.x < DRT_PERMUTE_LUT_BYTES / 16) ((uint4*)lut)[threadIdx.x] = v;
	for (int i = threadIdx.x + blockDim.x; i < DRT_PERMUTE_LUT_BYTES / 16; i += blockDim.x) // blocks of fewer than 128 threads
		((uint4*)lut)[i] = __ldg((const uint4*)g_permute_lut.v + i);
	__syncthreads();
}

__device__ __forceinline__ void stack_bind(LaneStack&, uint32_t*) {}
__device__ __forceinline__ void stack_bind(SharedLaneStack& K, uint32_t* words) { K.s = words + threadIdx.x; }

// One warp: every lane owns one pixel (valid or not) for all samples of the launch.
// POOL (sample pool): a warp of the fixed-tile kernel lives until its SLOWEST lane has finished; a lane's work is a sum of spp
// random path lengths, so that wait grows like sqrt(spp): measured t(spp) = 0.481 spp + 0.525 sqrt(spp) ms on room.glb 1080p
// (profiles/r02c_strided_tiles_and_spp_model.txt) -- 6 % of a 256-spp launch, 16 % of the 32-spp launches of 8 GPUs sharing a frame.
// Here every lane renders only the first n - R samples of its own pixel; the last R samples of all 32 pixels of the tile
// form a pool of 32 R (pixel, sample) items that the lanes take one at a time as they become free (one shared-memory atomic
// per item), so fast lanes take more of them and the warp ends within about one sample of its average lane. Pool samples stay
// in the warp's own tile (the rays stay coherent). Their colours go to a scratch array and every pixel adds its R pool
// colours in sample order at the end: per pixel the same values are added in the same order as always (same bits).
struct SamplePool {
	unsigned* counters;  // shared memory, one word per warp of the block: next pool item
	uint32_t* cold;      // shared memory, 4 words per thread ([word][thread], DRT_POOL_COLD_THREADS threads): the pixel's running sum
	                     // and its coordinates -- touched once per sample, kept out of the registers of the NODE / TRI steps
	int warps_per_block;
};
#define DRT_POOL_COLD_THREADS 128
// (pointers and tile coordinates are recomputed where they are needed -- once per pool sample -- instead of being carried in
// registers through the NODE / TRI steps: the kernel runs at its register limit)
__device__ __forceinline__ float4* pool_scratch_of_warp(const RenderParams& P, const SamplePool& pool, uint32_t pool_r)
{
	return (float4*)P.pool_scratch + ((size_t)blockIdx.x * pool.warps_per_block + (threadIdx.x >> 5)) * 32u * (size_t)pool_r;
}

template <bool STRICT, bool STATS, int FMT, bool TRI2 = false, class STACK = LaneStack, bool WT = false, bool POOL = false>
__device__ __forceinline__ V3 render_warp_phased(const RenderParams& P, const PhasedTuning T, uint32_t x, uint32_t y, bool valid,
	bool rtl, V3 acc, PathCounters& pc, TraceStats* st, uint8_t* lut, uint32_t* stack_words = nullptr, SamplePool pool = SamplePool())
{
	const uint4 lut_part = permute_lut_load();
#define DRT_POOL_R (POOL ? ((uint32_t)T.pool < P.n_samples ? (uint32_t)T.pool : P.n_samples) : 0u) /* R */
#define DRT_POOL_OWN_END (P.n_samples - DRT_POOL_R)                                                 /* samples of the own pixel */
	// POOL keeps the lane's own pixel and the pixel it is sampling in ONE register each (x | y << 16) and no item number: a lane is
	// on its own pixel while S.k < DRT_POOL_OWN_END, and a pool item is (S.k - DRT_POOL_OWN_END) * 32 + position in the tile
	uint32_t xy = x | (y << 16);
	uint32_t* const cold = POOL ? pool.cold + threadIdx.x : nullptr;
#define DRT_POOL_XY_OWN (cold[3 * DRT_POOL_COLD_THREADS])
	if (POOL) {
		cold[0] = __float_as_uint(acc.x); cold[DRT_POOL_COLD_THREADS] = __float_as_uint(acc.y); cold[2 * DRT_POOL_COLD_THREADS] = __float_as_uint(acc.z);
		DRT_POOL_XY_OWN = xy;
	}
	LanePath S;
	LaneTrav L;
	STACK K;
	stack_bind(K, stack_words);
	RayBox R;
	S.acc = acc; S.k = 0; S.fn = mk3(0.f, 0.f, 0.f);
	S.o = mk3(0.f, 0.f, 0.f); S.d = mk3(0.f, 0.f, 1.f);
	S.light = S.throughput = mk3(0.f, 0.f, 0.f);
	S.seed = 0; S.bounce = 0; S.shadow = false;
	int phase = PH_DONE;
	const int first_phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
	trav_begin(L, FLT_MAX);
	R.oct_inv = 0u;
	R.idnx = R.idny = R.idnz = R.bnx = R.bny = R.bnz = R.idfx = R.idfy = R.idfz = R.bfx = R.bfy = R.bfz = 0.f;
	// POOL: the next pool item whose pixel lies inside the image; false when the pool is empty
	// POOL: takes the next pool item whose pixel lies inside the image (xy, S.k = its pixel and sample); false when the pool is empty
	auto pool_take = [&]() -> bool {
		for (;;) {
			const uint32_t item = atomicAdd(pool.counters + (threadIdx.x >> 5), 1u);
			if (item >= 32u * DRT_POOL_R) return false;
			const uint32_t lane = threadIdx.x & 31u, xy_own = DRT_POOL_XY_OWN;
			const uint32_t px = (xy_own & 0xffffu) - (lane & 7u) + (item & 7u), py = (xy_own >> 16) - (lane >> 3) + ((item >> 3) & 3u);
			if (px < (uint32_t)P.width && py < (uint32_t)P.height) { xy = px | (py << 16); S.k = DRT_POOL_OWN_END + (item >> 5); return true; }
		}
	};
	bool start = valid && P.n_samples > 0; // this lane has a sample to start
	if (POOL) {
		// (a lane without a pixel of its own -- ragged image edge -- still works on the pool of its tile)
		const bool tile_live = __ballot_sync(0xffffffffu, valid) != 0u && P.n_samples > 0;
		start = tile_live && ((valid && DRT_POOL_OWN_END > 0u) || pool_take());
	}
	if (start) {
		path_begin_sample<STRICT>(P, POOL ? (xy & 0xffffu) : x, POOL ? (xy >> 16) : y, rtl, S);
		ray_box_setup(S.o, S.d, R);
		phase = first_phase;
	}
	permute_lut_store(lut, lut_part); // every thread of the block gets here (no early exit above)
	// Full vote: two ballots over the bits of `phase` (NODE 0, TRI 1, SHADE 2, DONE 3) give the three lane sets; an
	// empty set can only win the arg-max when all three are empty, which is the exit condition. Sticky re-issue:
	// after a step, ONE ballot asks how many lanes want the same step again; while at least T.stick do, it is
	// re-issued without a full vote (6 instead of 28 instructions per iteration).
	for (;;) {
		const unsigned b0 = __ballot_sync(0xffffffffu, (phase & 1) != 0), b1 = __ballot_sync(0xffffffffu, (phase & 2) != 0);
		if ((b0 & b1) == 0xffffffffu) break;
		const int cN = __popc(~(b0 | b1)), cT = __popc(b0 & ~b1), cS = __popc(b1 & ~b0);
		const int pick = (cN >= cT && cN >= cS) ? PH_NODE : (cT >= cS ? PH_TRI : PH_SHADE);
		do {
			if (phase == pick) {
				if (pick == PH_NODE) {
					if (FMT == NODE_FMT_F32) node_step_f32<STATS>(P.scene.nodes_f32, L, K, R, st, lut);
					else node_step_q8<STATS>(P.scene.nodes, L, K, R, st, lut);
					phase = trav_advance(L, K);
				}
				else if (pick == PH_TRI) {
					if (WT ? tri_step_wt<STATS>(P.scene, L, S.o, S.d, S.shadow, st)
					       : (TRI2 ? tri_step2<STRICT, STATS>(P.scene, P.scene.tris, L, S.o, S.d, S.shadow, st)
					               : tri_step<STRICT, STATS>(P.scene, P.scene.tris, L, S.o, S.d, S.shadow, st))) phase = PH_SHADE;
					else phase = trav_advance(L, K);
				}
				else {
					bool more;
					if (POOL) {
						more = true;
						if (shade_core<STRICT, STATS>(P, rtl, L.hit, S, pc, st)) { // the sample is complete
							const V3 c = post_process<STRICT>(P, S.light);
							bool take = true;
							if (S.k < DRT_POOL_OWN_END) { // a sample of the own pixel
								const V3 a = M<STRICT>::add3(mk3(__uint_as_float(cold[0]), __uint_as_float(cold[DRT_POOL_COLD_THREADS]), __uint_as_float(cold[2 * DRT_POOL_COLD_THREADS])), c);
								cold[0] = __float_as_uint(a.x); cold[DRT_POOL_COLD_THREADS] = __float_as_uint(a.y); cold[2 * DRT_POOL_COLD_THREADS] = __float_as_uint(a.z);
								S.k++;
								take = S.k >= DRT_POOL_OWN_END; // (xy is still the own pixel)
							}
							else { // a pool sample: its colour goes to [sample of the pool][position of its pixel in the tile]
								const uint32_t lane = threadIdx.x & 31u, xy_own = DRT_POOL_XY_OWN;
								const uint32_t pos = ((xy & 0xffffu) - ((xy_own & 0xffffu) - (lane & 7u))) + 8u * ((xy >> 16) - ((xy_own >> 16) - (lane >> 3)));
								pool_scratch_of_warp(P, pool, DRT_POOL_R)[(S.k - DRT_POOL_OWN_END) * 32u + pos] = make_float4(c.x, c.y, c.z, 0.f);
							}
							if (take) more = pool_take();
							if (more) path_begin_sample<STRICT>(P, xy & 0xffffu, xy >> 16, rtl, S);
						}
					}
					else more = shade_step<STRICT, STATS>(P, x, y, rtl, L.hit, S, pc, st);
					if (!more) phase = PH_DONE;
					else {
						ray_box_setup(S.o, S.d, R);
						trav_begin(L, FLT_MAX);
						phase = first_phase;
					}
				}
			}
		} while (__popc(__ballot_sync(0xffffffffu, phase == pick)) >= T.stick);
	}
	if (POOL) S.acc = mk3(__uint_as_float(cold[0]), __uint_as_float(cold[DRT_POOL_COLD_THREADS]), __uint_as_float(cold[2 * DRT_POOL_COLD_THREADS]));
	if (POOL && DRT_POOL_R > 0u) {
		// every pool colour of the tile is in the scratch array: each pixel adds its own in sample order (RenderKernel.cu:29)
		__syncwarp();
		if (valid) {
			const float4* sc = pool_scratch_of_warp(P, pool, DRT_POOL_R) + (threadIdx.x & 31u);
			for (uint32_t j = 0; j < DRT_POOL_R; j++) {
				const float4 c = sc[j * 32u];
				S.acc = M<STRICT>::add3(S.acc, mk3(c.x, c.y, c.z));
			}
		}
	}
#undef DRT_POOL_R
#undef DRT_POOL_OWN_END
#undef DRT_POOL_XY_OWN
	return S.acc;
}

// ---------------------------------------------------------------------------------------------
// The same loop with the SHADE-only part of the lane state parked in shared memory.
// light, throughput, normal, pixel sum, RNG state, sample / bounce counters, ray counters and the pixel coordinates
// (18 words per lane) are dead weight in the registers of the NODE and TRI steps, which issue ~75 % of the
// instructions. They live in a [field][thread] array of shared memory (conflict-free) and are loaded / stored around
// the SHADE step; what stays in registers is the ray (o, d, shadow flag), the slab constants and the traversal
// cursor. That fits 64 registers without spills in the hot steps -> 8 resident blocks per SM instead of 6.
// QUEUE adds the pixel queue of the persistent kernel (see below); per pixel the arithmetic is unchanged.
// ---------------------------------------------------------------------------------------------
enum { CF_LIGHT = 0, CF_THRU = 3, CF_FN = 6, CF_ACC = 9, CF_SEED = 12, CF_K = 13, CF_BOUNCE = 14, CF_CLOSEST = 15, CF_SHADOW = 16, CF_XY = 17, CF_COUNT = 18 };
#define DRT_COLD_THREADS 128
#define DRT_COLD_WORDS (CF_COUNT * DRT_COLD_THREADS)
#define DRT_QUEUE_NO_PIXEL 0xffffffffu

__device__ __forceinline__ void cold_put3(uint32_t* c, int f, V3 v)
{
	c[(f + 0) * DRT_COLD_THREADS] = __float_as_uint(v.x); c[(f + 1) * DRT_COLD_THREADS] = __float_as_uint(v.y); c[(f + 2) * DRT_COLD_THREADS] = __float_as_uint(v.z);
}
__device__ __forceinline__ V3 cold_get3(const uint32_t* c, int f)
{
	return mk3(__uint_as_float(c[(f + 0) * DRT_COLD_THREADS]), __uint_as_float(c[(f + 1) * DRT_COLD_THREADS]), __uint_as_float(c[(f + 2) * DRT_COLD_THREADS]));
}

// Pixel queue of one persistent warp (warp-uniform registers): the 8x4 tile it is handing out and how many of its pixels
// are gone. Tiles come from ONE global counter (lane 0, one atomic per 32 pixels).
struct WarpQueue {
	unsigned org;    // tile origin, x | y << 16
	unsigned cursor; // pixels of the tile already handed out (32 = none left)
	unsigned pending; // lane 0: index of the next tile, requested ahead of its use
};

// REFILL step (warp-uniform call): the lanes of `idle` take the next pixels of the warp's queue in rank order, load the
// pixel's sum and start its first sample. Returns false when the queue is empty and the global counter has run out.
template <bool STRICT>
__device__ __forceinline__ bool queue_refill(const RenderParams& P, WarpQueue& q, unsigned idle, bool rtl, uint32_t* c,
	V3& o, V3& d, RayBox& R, LaneTrav& L, int& phase, int first_phase)
{
	const unsigned lane = threadIdx.x & 31u;
	const unsigned tiles_x = (unsigned)(P.width + 7) >> 3, tiles_y = (unsigned)(P.height + 3) >> 2;
	uint32_t xy = DRT_QUEUE_NO_PIXEL;
	bool progress = false;
	for (int it = 0; it < 2 && idle != 0u; it++) {
		if (q.cursor == 32u) {
			// the tile index was requested when the previous tile was opened: the atomic's round trip is off the critical path
			const unsigned long long gt = (unsigned long long)__shfl_sync(0xffffffffu, q.pending, 0) * P.tile_stride + P.tile_first; // k-th tile this launch owns
			if (gt >= (unsigned long long)tiles_x * tiles_y) break;
			const unsigned t = (unsigned)gt;
			if (lane == 0) q.pending = atomicAdd(P.work_queue, 1u);
			const unsigned ty = t / tiles_x, tx = t - ty * tiles_x;
			q.org = (tx * 8u) | ((ty * 4u) << 16);
			q.cursor = 0u;
		}
		const unsigned rank = (unsigned)__popc(idle & ((1u << lane) - 1u));
		const bool mine = ((idle >> lane) & 1u) != 0u && rank < 32u - q.cursor;
		if (mine) {
			const unsigned i = q.cursor + rank;
			const unsigned x = (q.org & 0xffffu) + (i & 7u), y = (q.org >> 16) + (i >> 3);
			if (x < (unsigned)P.width && y < (unsigned)P.height) xy = x | (y << 16);
		}
		const unsigned taken = __ballot_sync(0xffffffffu, mine);
		idle &= ~taken;
		q.cursor += (unsigned)__popc(taken);
		progress = true;
	}
	if (xy != DRT_QUEUE_NO_PIXEL) {
		const uint32_t x = xy & 0xffffu, y = xy >> 16;
		LanePath S;
		const float4 a4 = P.accum[(size_t)x + (size_t)y * (size_t)P.width];
		S.acc = mk3(a4.x, a4.y, a4.z); S.k = 0; S.fn = mk3(0.f, 0.f, 0.f);
		path_begin_sample<STRICT>(P, x, y, rtl, S);
		cold_put3(c, CF_LIGHT, S.light); cold_put3(c, CF_THRU, S.throughput); cold_put3(c, CF_FN, S.fn); cold_put3(c, CF_ACC, S.acc);
		c[CF_SEED * DRT_COLD_THREADS] = S.seed; c[CF_K * DRT_COLD_THREADS] = S.k; c[CF_BOUNCE * DRT_COLD_THREADS] = (uint32_t)S.bounce;
		c[CF_XY * DRT_COLD_THREADS] = xy;
		o = S.o; d = S.d;
		ray_box_setup(o, d, R);
		trav_begin(L, FLT_MAX);
		phase = first_phase;
	}
	return progress;
}

template <bool STRICT, bool STATS, int FMT, bool QUEUE, bool TRI2 = false>
__device__ __forceinline__ void render_warp_cold(const RenderParams& P, const PhasedTuning T, uint32_t x0, uint32_t y0, bool valid0,
	bool rtl, TraceStats* st, uint8_t* lut, uint32_t* cold)
{
	WarpQueue q;
	q.org = 0u; q.cursor = 32u; q.pending = 0xffffffffu;
	if (QUEUE && (threadIdx.x & 31) == 0 && P.n_samples > 0) q.pending = atomicAdd(P.work_queue, 1u);
	int refill_min = (QUEUE && P.n_samples > 0) ? (T.refill > 32 ? 32 : (T.refill < 1 ? 1 : T.refill)) : 33;
	const uint4 lut_part = permute_lut_load();
	uint32_t* c = cold + threadIdx.x;
	LaneTrav L;
	LaneStack K;
	RayBox R;
	V3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
	bool shadow = false;
	int phase = PH_DONE;
	const int first_phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
	trav_begin(L, FLT_MAX);
	R.oct_inv = 0u;
	R.idnx = R.idny = R.idnz = R.bnx = R.bny = R.bnz = R.idfx = R.idfy = R.idfz = R.bfx = R.bfy = R.bfz = 0.f;
	{
		LanePath S;
		S.acc = mk3(0.f, 0.f, 0.f); S.k = 0; S.fn = mk3(0.f, 0.f, 0.f);
		S.o = o; S.d = d;
		S.light = S.throughput = mk3(0.f, 0.f, 0.f);
		S.seed = 0; S.bounce = 0; S.shadow = false;
		uint32_t xy = DRT_QUEUE_NO_PIXEL;
		if (!QUEUE && valid0) {
			xy = x0 | (y0 << 16);
			const float4 a4 = P.accum[(size_t)x0 + (size_t)y0 * (size_t)P.width];
			S.acc = mk3(a4.x, a4.y, a4.z);
			if (P.n_samples > 0) {
				path_begin_sample<STRICT>(P, x0, y0, rtl, S);
				o = S.o; d = S.d;
				ray_box_setup(o, d, R);
				phase = first_phase;
			}
		}
		cold_put3(c, CF_LIGHT, S.light); cold_put3(c, CF_THRU, S.throughput); cold_put3(c, CF_FN, S.fn); cold_put3(c, CF_ACC, S.acc);
		c[CF_SEED * DRT_COLD_THREADS] = S.seed; c[CF_K * DRT_COLD_THREADS] = S.k; c[CF_BOUNCE * DRT_COLD_THREADS] = (uint32_t)S.bounce;
		c[CF_CLOSEST * DRT_COLD_THREADS] = 0u; c[CF_SHADOW * DRT_COLD_THREADS] = 0u; c[CF_XY * DRT_COLD_THREADS] = xy;
	}
	permute_lut_store(lut, lut_part);
	for (;;) {
		const unsigned b0 = __ballot_sync(0xffffffffu, (phase & 1) != 0), b1 = __ballot_sync(0xffffffffu, (phase & 2) != 0);
		if (QUEUE) {
			// REFILL step: once at least T.refill lanes idle, they take the next pixels of the warp's queue TOGETHER (rank order
			// = pixel order of the tile), so that neighbouring pixels start their primary rays in the same steps.
			const unsigned idle = b0 & b1;
			if (__popc(idle) >= refill_min) {
				if (queue_refill<STRICT>(P, q, idle, rtl, c, o, d, R, L, phase, first_phase)) continue;
				refill_min = 33; // the queue has run out: never again
			}
		}
		if ((b0 & b1) == 0xffffffffu) break;
		const int cN = __popc(~(b0 | b1)), cT = __popc(b0 & ~b1), cS = __popc(b1 & ~b0);
		const int pick = (cN >= cT && cN >= cS) ? PH_NODE : (cT >= cS ? PH_TRI : PH_SHADE);
		do {
			if (phase == pick) {
				if (pick == PH_NODE) {
					if (FMT == NODE_FMT_F32) node_step_f32<STATS>(P.scene.nodes_f32, L, K, R, st, lut);
					else node_step_q8<STATS>(P.scene.nodes, L, K, R, st, lut);
					phase = trav_advance(L, K);
				}
				else if (pick == PH_TRI) {
					if (TRI2 ? tri_step2<STRICT, STATS>(P.scene, P.scene.tris, L, o, d, shadow, st)
					         : tri_step<STRICT, STATS>(P.scene, P.scene.tris, L, o, d, shadow, st)) phase = PH_SHADE;
					else phase = trav_advance(L, K);
				}
				else {
					LanePath S;
					PathCounters pc;
					S.o = o; S.d = d; S.shadow = shadow;
					S.light = cold_get3(c, CF_LIGHT); S.throughput = cold_get3(c, CF_THRU); S.fn = cold_get3(c, CF_FN); S.acc = cold_get3(c, CF_ACC);
					S.seed = c[CF_SEED * DRT_COLD_THREADS]; S.k = c[CF_K * DRT_COLD_THREADS]; S.bounce = (int)c[CF_BOUNCE * DRT_COLD_THREADS];
					pc.closest = c[CF_CLOSEST * DRT_COLD_THREADS]; pc.shadow = c[CF_SHADOW * DRT_COLD_THREADS];
					const uint32_t xy = c[CF_XY * DRT_COLD_THREADS];
					const bool more = shade_step<STRICT, STATS>(P, xy & 0xffffu, xy >> 16, rtl, L.hit, S, pc, st);
					if (QUEUE && !more) {
						// all samples of the pixel are in (RenderKernel.cu:29-31); the lane idles until the next refill
						const size_t p = (size_t)(xy & 0xffffu) + (size_t)(xy >> 16) * (size_t)P.width;
						P.accum[p] = make_float4(S.acc.x, S.acc.y, S.acc.z, 0.f);
						P.display[p] = make_float4(S.acc.x / P.total_samples, S.acc.y / P.total_samples, S.acc.z / P.total_samples, 1.f);
					}
					cold_put3(c, CF_LIGHT, S.light); cold_put3(c, CF_THRU, S.throughput); cold_put3(c, CF_FN, S.fn); cold_put3(c, CF_ACC, S.acc);
					c[CF_SEED * DRT_COLD_THREADS] = S.seed; c[CF_K * DRT_COLD_THREADS] = S.k; c[CF_BOUNCE * DRT_COLD_THREADS] = (uint32_t)S.bounce;
					c[CF_CLOSEST * DRT_COLD_THREADS] = pc.closest; c[CF_SHADOW * DRT_COLD_THREADS] = pc.shadow;
					o = S.o; d = S.d; shadow = S.shadow;
					if (!more) phase = PH_DONE;
					else {
						ray_box_setup(o, d, R);
						trav_begin(L, FLT_MAX);
						phase = first_phase;
					}
				}
			}
		} while (__popc(__ballot_sync(0xffffffffu, phase == pick)) >= T.stick);
	}
}

#endif


// =============================================================================================
// Five-step decomposition (NODE, TRI, HIT, UVEC, END) with overlapping NODE/TRI eligibility.
//   * A lane in traversal may hold BOTH pending triangles and pending child nodes. It can take part in a
//     NODE step (its pending triangles are parked on the traversal stack as a triangle group, the way
//     compressed-wide-BVH traversal postpones leaves) or in a TRI step, whichever the warp runs next.
//   * The per-ray shading of RayGen.cuh:88-169 is cut where its control flow forks, so that the
//     expensive pieces run for lanes that all need them:
//       HIT  : closest-hit ray finished  -> sky on a miss, else material / normal / next origin
//       UVEC : one random_unit_vec3 draw -> the sun shadow direction, or ONE try of the scatter rejection
//              loop (Random.cu:50-58; a rejected try leaves the lane in UVEC for the next vote)
//       END  : sample finished           -> per-sample post-process, accumulate, next camera ray
// =============================================================================================
enum { ST_NODE = 0, ST_TRI = 1, ST_HIT = 2, ST_UVEC = 3, ST_END = 4, ST_COUNT = 5 };
enum { LP_TRAV = 0, LP_HIT = 1, LP_UVEC = 2, LP_END = 3, LP_DONE = 4 };
enum { UV_SHADOWGEN = 0, UV_SCATTER = 1 };

struct LanePath5 {
	V3 o, d;
	V3 light, throughput;
	V3 fn;
	V3 acc;
	uint32_t seed;
	uint32_t k;
	int bounce;
	bool shadow;   // the ray being traced is a sun shadow ray
	bool sunflag;  // a shadow ray just finished: UVEC applies its result before drawing
	int uvmode;
};

// After a NODE or TRI step: true while the lane still has traversal work (pops the stack if needed).
DRT_HD bool trav_settle(LaneTrav& L, LaneStack& S)
{
	if (L.tri_bits != 0u || L.ng_bits > 0x00ffffffu) return true;
	for (;;) {
		if (L.sp == 0) return false;
		--L.sp;
		if (L.sp < DRT_PSTACK) break;
	}
	const uint32_t x = S.x[L.sp], y = S.y[L.sp];
	if (y > 0x00ffffffu) { L.ng_base = x; L.ng_bits = y; }
	else { L.tri_base = x; L.tri_bits = y; }
	return true;
}

// Parks the pending triangle group so that the lane can run a NODE step first.
template <bool STATS>
DRT_HD void trav_postpone_tris(LaneTrav& L, LaneStack& S, TraceStats* st)
{
	if (L.tri_bits == 0u) return;
	if (L.sp < DRT_PSTACK) { S.x[L.sp] = L.tri_base; S.y[L.sp] = L.tri_bits; }
	++L.sp;
	if (STATS && st && (uint32_t)L.sp > st->max_stack) st->max_stack = (uint32_t)L.sp;
	L.tri_bits = 0u;
}

template <bool STRICT>
DRT_HD void path5_begin_sample(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, LanePath5& S)
{
	S.light = mk3(0.f, 0.f, 0.f); S.throughput = mk3(1.f, 1.f, 1.f);
	S.bounce = 0; S.shadow = false; S.sunflag = false;
	S.seed = (x + y * (uint32_t)P.width) * (P.frame_first + S.k * P.frame_stride);
	const RayOD r = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, S.seed, rtl);
	S.o = r.o; S.d = r.d;
}

// HIT: a closest-hit ray finished. Returns the lane's next phase (LP_END on a miss, else LP_UVEC).
template <bool STRICT, bool STATS>
DRT_HD int hit_step(const RenderParams& P, const Hit& hit, LanePath5& S, PathCounters& pc, TraceStats* st)
{
	typedef M<STRICT> m;
	const SceneView& sc = P.scene;
	const ShadeParams& SP = P.shade;
	pc.closest++;
	S.seed += (uint32_t)S.bounce;
	if (hit.prim < 0) {
		const V3 sky = sky_model<STRICT>(S.d, SP.sky_color);
		S.light = m::add3(S.light, m::scale(m::mul3(sky, S.throughput), SP.sky_intensity));
		return LP_END;
	}
	const drt_f4 r1 = ld_f4(sc.tris + 3 * (size_t)hit.prim + 1);
	const uint32_t mat = as_uint(r1.w) & 0x7fffffffu;
	const drt_f4 mrec = ld_f4(sc.materials + mat);
	const int tex = (int)as_uint(mrec.w);
	if (tex < 0) S.throughput = m::mul3(S.throughput, mk3(mrec.x, mrec.y, mrec.z));
	else {
		const V2 bc = hit_barycentrics<STRICT>(sc.tris, hit.prim, S.o, S.d);
			const V2 tuv = interp_uv<STRICT>(sc, hit.prim, bc.x, bc.y);
		S.throughput = m::mul3(S.throughput, texture_get_pixel<STRICT>(sc, tex, tuv, st));
	}
	const drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)hit.prim);
	V3 fn = mk3(s0.x, s0.y, s0.z);
	const V3 pos = m::add3(S.o, m::scale(S.d, hit.t));
	if (m::dot(fn, m::normalize(S.d)) > 0.f) fn = m::scale(fn, -1.f);
	S.fn = fn;
	S.o = m::add3(pos, m::scale(fn, 0.001f));
	S.uvmode = SP.enable_sunlight ? UV_SHADOWGEN : UV_SCATTER;
	return LP_UVEC;
}

// UVEC: one draw. Returns LP_TRAV (S.o/S.d hold a new ray), LP_UVEC (rejected try) or LP_END (bounce limit).
template <bool STRICT>
DRT_HD int uvec_step(const RenderParams& P, bool rtl, bool shadow_found, LanePath5& S, PathCounters& pc)
{
	typedef M<STRICT> m;
	const ShadeParams& SP = P.shade;
	if (S.sunflag) {
		pc.shadow++;
		if (!shadow_found) S.light = m::add3(S.light, m::mul3(SP.suncol, S.throughput));
		S.sunflag = false; S.shadow = false;
		S.uvmode = UV_SCATTER;
	}
	const V3 v = random_unit_vec3<STRICT>(S.seed, rtl);
	if (S.uvmode == UV_SHADOWGEN) {
		S.d = m::add3(SP.sunpos, m::scale(v, 1.5f));
		S.shadow = true;
		return LP_TRAV;
	}
	const float len = m::length(v);
	if (!(m::mul(len, len) < 1.f)) return LP_UVEC;
	S.d = m::add3(S.fn, v);
	S.bounce++;
	return S.bounce > SP.bounce_limit ? LP_END : LP_TRAV;
}

// END: the sample is complete. Returns LP_DONE or LP_TRAV (next sample's camera ray in S.o/S.d).
template <bool STRICT>
DRT_HD int end_step(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, LanePath5& S)
{
	typedef M<STRICT> m;
	S.acc = m::add3(S.acc, post_process<STRICT>(P, S.light));
	S.k++;
	if (S.k >= P.n_samples) return LP_DONE;
	path5_begin_sample<STRICT>(P, x, y, rtl, S);
	return LP_TRAV;
}

// Phase of a lane whose traversal just ran out of work.
DRT_HD int trav_finished_phase(LanePath5& S)
{
	if (S.shadow) { S.sunflag = true; return LP_UVEC; }
	return LP_HIT;
}

// All state of one lane for the five-step kernels.
struct Lane5 {
	LanePath5 S;
	LaneTrav L;
	LaneStack K;
	RayBox R;
	int phase;
};

template <bool STRICT>
DRT_HD void lane5_init(const RenderParams& P, uint32_t x, uint32_t y, bool valid, bool rtl, V3 acc, Lane5& A)
{
	A.S.acc = acc; A.S.k = 0; A.S.fn = mk3(0.f, 0.f, 0.f);
	A.S.o = mk3(0.f, 0.f, 0.f); A.S.d = mk3(0.f, 0.f, 1.f);
	A.S.light = A.S.throughput = mk3(0.f, 0.f, 0.f);
	A.S.seed = 0; A.S.bounce = 0; A.S.shadow = false; A.S.sunflag = false; A.S.uvmode = UV_SCATTER;
	trav_begin(A.L, FLT_MAX);
	A.R.oct_inv = 0u;
	A.R.idnx = A.R.idny = A.R.idnz = A.R.bnx = A.R.bny = A.R.bnz = 0.f;
	A.R.idfx = A.R.idfy = A.R.idfz = A.R.bfx = A.R.bfy = A.R.bfz = 0.f;
	A.phase = LP_DONE;
	if (valid && P.n_samples > 0) {
		path5_begin_sample<STRICT>(P, x, y, rtl, A.S);
		ray_box_setup(A.S.o, A.S.d, A.R);
		A.phase = P.scene.num_nodes > 0 ? LP_TRAV : LP_HIT;
	}
}

// A new ray is in S.o / S.d: reset the traversal.
DRT_HD void lane5_new_ray(const RenderParams& P, Lane5& A)
{
	ray_box_setup(A.S.o, A.S.d, A.R);
	trav_begin(A.L, FLT_MAX);
	if (P.scene.num_nodes <= 0) A.phase = trav_finished_phase(A.S);
}

DRT_HD bool lane5_wants(const Lane5& A, int step)
{
	switch (step) {
	case ST_NODE: return A.phase == LP_TRAV && A.L.ng_bits > 0x00ffffffu;
	case ST_TRI: return A.phase == LP_TRAV && A.L.tri_bits != 0u;
	case ST_HIT: return A.phase == LP_HIT;
	case ST_UVEC: return A.phase == LP_UVEC;
	default: return A.phase == LP_END;
	}
}

// Executes `step` for a lane that wants it.
template <bool STRICT, bool STATS, int FMT>
DRT_HD void lane5_run(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, int step, Lane5& A, PathCounters& pc, TraceStats* st)
{
	if (step == ST_NODE) {
		trav_postpone_tris<STATS>(A.L, A.K, st);
		if (FMT == NODE_FMT_F32) node_step_f32<STATS>(P.scene.nodes_f32, A.L, A.K, A.R, st);
		else node_step_q8<STATS>(P.scene.nodes, A.L, A.K, A.R, st);
		if (!trav_settle(A.L, A.K)) A.phase = trav_finished_phase(A.S);
	}
	else if (step == ST_TRI) {
		if (tri_step<STRICT, STATS>(P.scene, P.scene.tris, A.L, A.S.o, A.S.d, A.S.shadow, st)) A.phase = trav_finished_phase(A.S);
		else if (!trav_settle(A.L, A.K)) A.phase = trav_finished_phase(A.S);
	}
	else if (step == ST_HIT) {
		A.phase = hit_step<STRICT, STATS>(P, A.L.hit, A.S, pc, st);
	}
	else if (step == ST_UVEC) {
		A.phase = uvec_step<STRICT>(P, rtl, A.L.hit.prim >= 0, A.S, pc);
		if (A.phase == LP_TRAV) lane5_new_ray(P, A);
	}
	else {
		A.phase = end_step<STRICT>(P, x, y, rtl, A.S);
		if (A.phase == LP_TRAV) lane5_new_ray(P, A);
	}
}

// One lane, serial schedule (triangles before nodes, like the reference-shaped traversal).
template <bool STRICT, bool STATS, int FMT>
DRT_HD V3 render_pixel_steps5(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, V3 acc, PathCounters& pc, TraceStats* st, bool nodes_first)
{
	Lane5 A;
	lane5_init<STRICT>(P, x, y, true, rtl, acc, A);
	while (A.phase != LP_DONE) {
		int step;
		if (A.phase == LP_TRAV) {
			const bool wn = lane5_wants(A, ST_NODE), wt = lane5_wants(A, ST_TRI);
			step = (wn && (nodes_first || !wt)) ? ST_NODE : ST_TRI;
		}
		else step = A.phase == LP_HIT ? ST_HIT : (A.phase == LP_UVEC ? ST_UVEC : ST_END);
		lane5_run<STRICT, STATS, FMT>(P, x, y, rtl, step, A, pc, st);
	}
	return A.S.acc;
}

} // namespace drt
