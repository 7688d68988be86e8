// drt_trace.cuh -- closest-hit / any-hit traversal of the compressed 8-wide BVH.
//
// Replaces, for one ray, the reference's
//   traverseBVH / traverseBVH_raytest   DustRayTracer/src/Core/BVH/BVHTraversal.cuh:14-134
//   Bounds3f::intersect                 DustRayTracer/src/Core/Bounds.cu:18-41
//   Intersection (Moller-Trumbore)      DustRayTracer/src/Core/Kernel/Shaders/Intersection.cu:4-36
//   AnyHit (alpha cut-out)              DustRayTracer/src/Core/Kernel/Shaders/AnyHit.cuh:8-28
//   Texture::getAlpha                   DustRayTracer/src/Core/Scene/Texture.cu:60-75
// The tree and the box test are ours (conservative: a box test may only ever ADMIT more triangles);
// the accept/reject predicate and t,u,v of a triangle are the reference's Moller-Trumbore in the
// reference's operation order, so the winning triangle is the same one: the minimum-t accepted
// candidate. Exact-t ties go to the lowest canonical id (the reference takes the first VISITED,
// BVHTraversal.cuh:51, which depends on its own tree) and are counted in Counters::tie_events.
#pragma once
#include "drt_types.h"

namespace drt {

#ifndef DRT_STACK_SIZE
#define DRT_STACK_SIZE 48
#endif

#define DRT_TRIANGLE_EPSILON 0.000001f // Common/physical_units.hpp:12

struct TraceStats {
	uint32_t nodes, tris, ties, alpha, max_stack, clamps;
};

DRT_HD drt_u4 ld_u4(const drt_u4* p)
{
#if defined(__CUDA_ARCH__)
	return __ldg(p);
#else
	return *p;
#endif
}
DRT_HD drt_f4 ld_f4(const drt_f4* p)
{
#if defined(__CUDA_ARCH__)
	return __ldg(p);
#else
	return *p;
#endif
}
DRT_HD uint32_t ld_u32(const uint32_t* p)
{
#if defined(__CUDA_ARCH__)
	return __ldg(p);
#else
	return *p;
#endif
}

DRT_HD int hi_bit(uint32_t v) // index of highest set bit, v != 0
{
#if defined(__CUDA_ARCH__)
	return 31 - __clz(v);
#else
	return 31 - __builtin_clz(v);
#endif
}
DRT_HD int lo_bit(uint32_t v) // index of lowest set bit, v != 0
{
#if defined(__CUDA_ARCH__)
	return __ffs(v) - 1;
#else
	return __builtin_ctz(v);
#endif
}
DRT_HD int pop_count(uint32_t v)
{
#if defined(__CUDA_ARCH__)
	return __popc(v);
#else
	return __builtin_popcount(v);
#endif
}

// Nearest texel with repeat wrap exactly as Texture.cu:35-36,65-66 index it. The reference can index
// one past a row when frac() rounds to 1.0; a linear index past the pool is clamped (and counted).
DRT_HD uint32_t texel_index(const TexDesc& tx, float u, float v, uint32_t* clamps)
{
	int x = (int)((u - floorf(u)) * (float)tx.width);
	int y = (int)((v - floorf(v)) * (float)tx.height);
	int idx = y * tx.width + x;
	int last = tx.width * tx.height - 1;
	if (idx > last || idx < 0) { idx = idx < 0 ? 0 : last; if (clamps) ++*clamps; }
	return tx.offset + (uint32_t)idx;
}

template <bool STRICT>
DRT_HD V2 interp_uv(const SceneView& sc, int slot, float bu, float bv)
{
	typedef M<STRICT> m;
	// UVW = (1-u-v, u, v); UV = UVW.x*uv0 + UVW.y*uv1 + UVW.z*uv2 (AnyHit.cuh:20-22, RayGen.cuh:116)
	drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)slot);
	drt_f4 s1 = ld_f4(sc.shade + 3 * (size_t)slot + 1);
	drt_f4 s2 = ld_f4(sc.shade + 3 * (size_t)slot + 2);
	float bw = m::sub(m::sub(1.0f, bu), bv);
	V2 r;
	r.x = m::add(m::add(m::mul(bw, s0.w), m::mul(bu, s1.y)), m::mul(bv, s1.w));
	r.y = m::add(m::add(m::mul(bw, s1.x), m::mul(bu, s1.z)), m::mul(bv, s2.x));
	return r;
}

// AnyHit.cuh:8-28 for a triangle whose material is flagged "albedo texture has 4 components".
template <bool STRICT>
DRT_HD bool alpha_accept(const SceneView& sc, int slot, uint32_t mat, float bu, float bv, TraceStats* st)
{
	drt_f4 mrec = ld_f4(sc.materials + mat);
	int tex = (int)as_uint(mrec.w);
	TexDesc tx = sc.textures[tex];
	V2 uv = interp_uv<STRICT>(sc, slot, bu, bv);
	uint32_t texel = ld_u32(sc.texels + texel_index(tx, uv.x, uv.y, st ? &st->clamps : nullptr));
	if (st) st->alpha++;
	return (texel >> 24) == 255u; // !(a/255 < 1)
}

// Traces one ray. ANYHIT: returns at the first accepted triangle (RayTest semantics: t > 1e-6,
// unbounded). Otherwise finds the closest accepted triangle with t < tmax.
// hit.prim is the LEAF-ORDER slot; the canonical id is as_uint(tris[3*slot].w).
// MODE: 0 = closest hit, 1 = any hit, 2 = decided per ray at run time by `anyhit_rt` (one call site
// serving both ray kinds keeps a warp converged inside the traversal loop).
template <bool STRICT, int MODE, bool STATS>
DRT_HD bool trace_ray_mode(const SceneView& sc, V3 o, V3 d, float tmax, Hit& hit, TraceStats* st, bool anyhit_rt)
{
	typedef M<STRICT> m;
	const bool ANYHIT = MODE == 2 ? anyhit_rt : (MODE == 1);
	hit.prim = -1;
	hit.t = tmax;
	hit.u = 0.f;
	hit.v = 0.f;
	if (sc.num_nodes <= 0) return false;
	uint32_t best_id = 0xffffffffu;

	// Box-test setup (ours, not parity-relevant): clamp tiny direction components so 1/d is finite.
	const float tiny = 1e-20f;
	float dx = fabsf(d.x) > tiny ? d.x : (d.x < 0.f ? -tiny : tiny);
	float dy = fabsf(d.y) > tiny ? d.y : (d.y < 0.f ? -tiny : tiny);
	float dz = fabsf(d.z) > tiny ? d.z : (d.z < 0.f ? -tiny : tiny);
	const float idx = 1.0f / dx, idy = 1.0f / dy, idz = 1.0f / dz;
	const uint32_t oct_inv = (d.x >= 0.f ? 4u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 1u : 0u);

	uint32_t stack_x[DRT_STACK_SIZE], stack_y[DRT_STACK_SIZE];
	int sp = 0;
	uint32_t ng_base = 0u, ng_bits = 0x80000000u; // root: one pending inner slot, imask 0 -> relative index 0

	for (;;) {
		uint32_t tri_base = 0u, tri_bits = 0u;
		if (ng_bits > 0x00ffffffu) {
			const int bit = hi_bit(ng_bits);
			ng_bits &= ~(1u << bit);
			const uint32_t slot = (uint32_t)(bit - 24) ^ oct_inv;
			const uint32_t rel = (uint32_t)pop_count(ng_bits & 0xffu & ((1u << slot) - 1u));
			const uint32_t node_index = ng_base + rel;
			if (ng_bits > 0x00ffffffu) {
				if (sp < DRT_STACK_SIZE) { stack_x[sp] = ng_base; stack_y[sp] = ng_bits; }
				++sp;
				if (STATS && st && (uint32_t)sp > st->max_stack) st->max_stack = (uint32_t)sp;
			}
			const drt_u4* np = sc.nodes + 5 * (size_t)node_index;
			const drt_u4 n0 = ld_u4(np), n1 = ld_u4(np + 1), n2 = ld_u4(np + 2), n3 = ld_u4(np + 3), n4 = ld_u4(np + 4);
			if (STATS && st) st->nodes++;

			const float px = as_float(n0.x), py = as_float(n0.y), pz = as_float(n0.z);
			const uint32_t ex = n0.w & 0xffu, ey = (n0.w >> 8) & 0xffu, ez = (n0.w >> 16) & 0xffu;
			const uint32_t imask = n0.w >> 24;
			const float sx = as_float(ex << 23) * idx, sy = as_float(ey << 23) * idy, sz = as_float(ez << 23) * idz;
			const float bx = (px - o.x) * idx, by = (py - o.y) * idy, bz = (pz - o.z) * idz;
			// near/far quantised planes per axis by ray sign
			const uint32_t qlx[2] = { n2.x, n2.y }, qly[2] = { n2.z, n2.w }, qlz[2] = { n3.x, n3.y };
			const uint32_t qhx[2] = { n3.z, n3.w }, qhy[2] = { n4.x, n4.y }, qhz[2] = { n4.z, n4.w };
			const uint32_t meta[2] = { n1.z, n1.w };
			const bool negx = idx < 0.f, negy = idy < 0.f, negz = idz < 0.f;
			const float tfar_lim = hit.t;
			uint32_t hitmask = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
			for (int i = 0; i < 8; i++) {
				const int w = i >> 2, sh = (i & 3) * 8;
				const uint32_t mt = (meta[w] >> sh) & 0xffu;
				if (mt == 0u) continue;
				const float lx = (float)((qlx[w] >> sh) & 0xffu), hx = (float)((qhx[w] >> sh) & 0xffu);
				const float ly = (float)((qly[w] >> sh) & 0xffu), hy = (float)((qhy[w] >> sh) & 0xffu);
				const float lz = (float)((qlz[w] >> sh) & 0xffu), hz = (float)((qhz[w] >> sh) & 0xffu);
				const float t0x = (negx ? hx : lx) * sx + bx, t1x = (negx ? lx : hx) * sx + bx;
				const float t0y = (negy ? hy : ly) * sy + by, t1y = (negy ? ly : hy) * sy + by;
				const float t0z = (negz ? hz : lz) * sz + bz, t1z = (negz ? lz : hz) * sz + bz;
				float tn = fmaxf(fmax3(t0x, t0y, t0z), 0.f);
				float tf = fminf(fmin3(t1x, t1y, t1z) * 1.000001f, tfar_lim);
				tn *= 0.999999f; // a few ulp of slack on both ends: the box test must never lose a triangle
				if (tn <= tf) {
					const bool inner = ((mt & (mt << 1)) & 0x10u) != 0u;
					const uint32_t bits = inner ? 1u : (mt >> 5);
					const uint32_t pos = inner ? ((mt & 31u) ^ oct_inv) : (mt & 31u);
					hitmask |= bits << pos;
				}
			}
			ng_base = n1.x;
			ng_bits = (hitmask & 0xff000000u) | imask;
			tri_base = n1.y;
			tri_bits = hitmask & 0x00ffffffu;
		}

		while (tri_bits != 0u) {
			const int k = lo_bit(tri_bits);
			tri_bits &= tri_bits - 1u;
			const int slot = (int)(tri_base + (uint32_t)k);
			const drt_f4* tp = sc.tris + 3 * (size_t)slot;
			const drt_f4 r0 = ld_f4(tp), r1 = ld_f4(tp + 1), r2 = ld_f4(tp + 2);
			if (STATS && st) st->tris++;
			const V3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
			// Intersection.cu:8-33
			const V3 pvec = m::cross(d, e2);
			const float det = m::dot(e1, pvec);
			if (det > -DRT_TRIANGLE_EPSILON && det < DRT_TRIANGLE_EPSILON) continue;
			const float inv_det = m::div(1.0f, det);
			const V3 tvec = m::sub3(o, v0);
			const float u = m::mul(inv_det, m::dot(tvec, pvec));
			if (u < 0.0f || u > 1.0f) continue;
			const V3 qvec = m::cross(tvec, e1);
			const float v = m::mul(inv_det, m::dot(d, qvec));
			if (v < 0.0f || m::add(u, v) > 1.0f) continue;
			const float t = m::mul(inv_det, m::dot(e2, qvec));
			if (!(t > DRT_TRIANGLE_EPSILON)) continue;
			const uint32_t matw = as_uint(r1.w);
			if (ANYHIT) {
				if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, u, v, st)) continue;
				hit.prim = slot; hit.t = t; hit.u = u; hit.v = v;
				return true;
			}
			else {
				const uint32_t id = as_uint(r0.w);
				if (!(t < hit.t)) {
					if (!(t == hit.t) || hit.prim < 0) continue;
					if (st) st->ties++;
					if (!(id < best_id)) continue;
				}
				if ((matw & 0x80000000u) && !alpha_accept<STRICT>(sc, slot, matw & 0x7fffffffu, u, v, st)) continue;
				hit.prim = slot; hit.t = t; hit.u = u; hit.v = v;
				best_id = id;
			}
		}

		if (ng_bits <= 0x00ffffffu) {
			if (sp == 0) break;
			--sp;
			if (sp >= DRT_STACK_SIZE) continue; // overflowed entry was dropped (counted via max_stack)
			ng_base = stack_x[sp];
			ng_bits = stack_y[sp];
		}
	}
	return hit.prim >= 0;
}

template <bool STRICT, bool ANYHIT, bool STATS>
DRT_HD bool trace_ray(const SceneView& sc, V3 o, V3 d, float tmax, Hit& hit, TraceStats* st)
{
	return trace_ray_mode<STRICT, ANYHIT ? 1 : 0, STATS>(sc, o, d, tmax, hit, st, ANYHIT);
}

} // namespace drt
