// drt_shade.cuh -- one pixel-sample of the bounce / shade loop.
//
// Semantics follow the reference's RayGen (DustRayTracer/src/Core/Kernel/Shaders/RayGen.cuh:63-172)
// step for step: seed schedule, sky on miss, albedo / texture throughput,
// sun shadow ray with the un-normalised jittered direction, un-normalised diffuse scatter direction,
// debug views, per-sample Uncharted-2 tonemap + sqrt gamma BEFORE accumulation.
//   SkyModel                         RayGen.cuh:54-61
//   uncharted2_* / gammaCorrection   RayGen.cuh:23-52
//   ClosestHit (P, face-forward N)   Kernel/Shaders/ClosestHit.cuh:4-28
//   Texture::getPixel                Scene/Texture.cu:33-58
#pragma once
#include "drt_types.h"
#include "drt_rng.cuh"
#include "drt_trace.cuh"

namespace drt {

enum { DRT_RENDER_NORMAL = 0, DRT_RENDER_DEBUG = 1 };
enum { DRT_DEBUG_ALBEDO = 0, DRT_DEBUG_NORMAL = 1, DRT_DEBUG_BARYCENTRIC = 2, DRT_DEBUG_UVS = 3, DRT_DEBUG_MESHBVH = 4, DRT_DEBUG_WORLDBVH = 5 };

struct PathCounters {
	uint32_t closest, shadow;
};

template <bool STRICT>
DRT_HD float tonemap_partial(float x)
{
	typedef M<STRICT> m;
	// ((x*(A*x + C*B) + D*E) / (x*(A*x + B) + D*F)) - E/F with A=.15 B=.5 C=.1 D=.2 E=.02 F=.3
	const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
	const float CB = C * B, DE = D * E, DF = D * F, EF = E / F; // scalar constants (folded identically on host)
	float num = m::add(m::mul(x, m::add(m::mul(A, x), CB)), DE);
	float den = m::add(m::mul(x, m::add(m::mul(A, x), B)), DF);
	return m::sub(m::div(num, den), EF);
}

template <bool STRICT>
DRT_HD V3 sky_model(V3 dir, V3 sky_color)
{
	typedef M<STRICT> m;
	float g = m::mul(0.5f, m::add(1.0f, m::normalize(dir).y));
	float w = m::sub(1.0f, g);
	V3 c = mk3(m::add(w, m::mul(g, sky_color.x)), m::add(w, m::mul(g, sky_color.y)), m::add(w, m::mul(g, sky_color.z)));
	return m::mul3(c, c); // powf(c, 2)
}

template <bool STRICT>
DRT_HD V3 texture_get_pixel(const SceneView& sc, int tex, V2 uv, TraceStats* st)
{
	typedef M<STRICT> m;
	TexDesc tx = sc.textures[tex];
	if (tx.comps != 3 && tx.comps != 4) return mk3(0.f, 0.f, 1.f); // uchar4{0,0,255} -> /255 -> squared
	uint32_t texel = ld_u32(sc.texels + texel_index(tx, uv.x, uv.y, st ? &st->clamps : nullptr));
	float r = m::div((float)(texel & 0xffu), 255.f);
	float g = m::div((float)((texel >> 8) & 0xffu), 255.f);
	float b = m::div((float)((texel >> 16) & 0xffu), 255.f);
	return mk3(m::mul(r, r), m::mul(g, g), m::mul(b, b));
}

// Schlick reflectance and refraction: the reference's dead helpers (Random.cu:26-40), used by the extended materials.
DRT_HD float schlick_reflectance(float cosine, float ri)
{
	float r0 = (1.f - ri) / (1.f + ri);
	r0 = r0 * r0;
	const float k = 1.f - cosine, k2 = k * k;
	return r0 + (1.f - r0) * (k2 * k2 * k);
}
DRT_HD V3 refract_dir(V3 uv, V3 n, float etai_over_etat)
{
	typedef M<false> m;
	const float cos_theta = fminf(-m::dot(uv, n), 1.0f);
	const V3 perp = m::scale(m::add3(uv, m::scale(n, cos_theta)), etai_over_etat);
	const V3 par = m::scale(n, -sqrtf(fabsf(1.0f - m::dot(perp, perp))));
	return m::add3(perp, par);
}

// One call of the reference's RayGen: radiance of pixel (x, y) for frame index `frameidx`.
// EXT (DRT_FLAG_EXTENDED_MATERIALS, beyond the reference): emission is added at every hit, metals reflect about the
// face normal with a roughness-scaled fuzz, dielectrics reflect or refract by Schlick's approximation; specular hits
// take no sun sample. Diffuse, non-emissive materials shade exactly as without EXT (same random draws).
template <bool STRICT, bool STATS, bool EXT = false>
DRT_HD V3 path_radiance(const RenderParams& P, uint32_t x, uint32_t y, uint32_t frameidx, bool rtl,
	PathCounters& pc, TraceStats* st)
{
	typedef M<STRICT> m;
	const SceneView& sc = P.scene;
	const ShadeParams& S = P.shade;
	uint32_t seed = x + y * (uint32_t)P.width;
	seed *= frameidx;
	RayOD ray = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, seed, rtl);

	V3 light = mk3(0.f, 0.f, 0.f);
	V3 throughput = mk3(1.f, 1.f, 1.f);
	V2 tex_uv = mk2(0.f, 1.f);
	const bool debug = S.render_mode == DRT_RENDER_DEBUG;

	for (int i = 0; i <= S.bounce_limit; i++) {
		Hit hit;
		uint32_t nodes_before = (STATS && st) ? st->nodes : 0u;
		bool found = trace_ray<STRICT, false, STATS>(sc, ray.o, ray.d, FLT_MAX, hit, st);
		pc.closest++;
		seed += (uint32_t)i;

		if (!found) {
			if (debug && S.debug_mode == DRT_DEBUG_MESHBVH) {
				// heat colour: 0.05 per visited node (BVHTraversal.cuh:43); ours counts wide nodes
				float heat = (STATS && st) ? 0.05f * (float)(st->nodes - nodes_before) : 0.f;
				light = mk3(heat, heat, heat);
			}
			else {
				V3 sky = sky_model<STRICT>(ray.d, S.sky_color);
				light = m::add3(light, m::scale(m::mul3(sky, throughput), S.sky_intensity));
			}
			break;
		}

		const drt_f4 r1 = ld_f4(sc.tris + 3 * (size_t)hit.prim + 1);
		const uint32_t mat = as_uint(r1.w) & 0x7fffffffu;
		const drt_f4 mrec = ld_f4(sc.materials + mat);
		const int tex = (int)as_uint(mrec.w);
		const V3 throughput_in = throughput;
		if (tex < 0) throughput = m::mul3(throughput, mk3(mrec.x, mrec.y, mrec.z));
		else {
			tex_uv = interp_uv<STRICT>(sc, hit.prim, hit.u, hit.v);
			throughput = m::mul3(throughput, texture_get_pixel<STRICT>(sc, tex, tex_uv, st));
		}

		// ClosestHit.cuh:17-25: P = o + d*t; N = face normal flipped against the ray
		const drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)hit.prim);
		V3 fn = mk3(s0.x, s0.y, s0.z);
		const V3 pos = m::add3(ray.o, m::scale(ray.d, hit.t));
		if (m::dot(fn, m::normalize(ray.d)) > 0.f) fn = m::scale(fn, -1.f);
		const V3 new_o = m::add3(pos, m::scale(fn, 0.001f));

		if (EXT && !debug) {
			const drt_f4 e0 = ld_f4(sc.materials_ext + 2 * (size_t)mat), e1 = ld_f4(sc.materials_ext + 2 * (size_t)mat + 1);
			light = m::add3(light, m::mul3(mk3(e0.x, e0.y, e0.z), throughput_in));
			const uint32_t kind = as_uint(e0.w);
			if (kind != 0u) {
				const V3 dn = m::normalize(ray.d);
				const V3 mirrored = m::sub3(dn, m::scale(fn, 2.f * m::dot(dn, fn)));
				if (kind == 1u) { // metal: fuzzy mirror, absorbed when the fuzz pushes it below the surface
					V3 dir = mirrored;
					if (e1.x > 0.f) dir = m::add3(dir, m::scale(random_unit_sphere_vec3<STRICT>(seed, rtl), e1.x));
					if (!(m::dot(dir, fn) > 0.f)) break;
					ray.o = new_o; ray.d = dir;
				}
				else { // dielectric: fn faces the incoming ray; it was flipped when the ray leaves the medium
					const bool entering = !(m::dot(mk3(s0.x, s0.y, s0.z), dn) > 0.f);
					const float ri = entering ? 1.f / e1.y : e1.y;
					const float cos_t = fminf(-m::dot(dn, fn), 1.f);
					const float sin_t = sqrtf(fmaxf(0.f, 1.f - cos_t * cos_t));
					if (ri * sin_t > 1.f || schlick_reflectance(cos_t, ri) > random_float(seed)) { ray.o = new_o; ray.d = mirrored; }
					else { ray.o = m::sub3(pos, m::scale(fn, 0.001f)); ray.d = refract_dir(dn, fn, ri); }
				}
				continue;
			}
		}

		if (S.enable_sunlight && S.render_mode == DRT_RENDER_NORMAL) {
			V3 sd = m::add3(S.sunpos, m::scale(random_unit_vec3<STRICT>(seed, rtl), 1.5f));
			Hit sh;
			bool occluded = trace_ray<STRICT, true, STATS>(sc, new_o, sd, FLT_MAX, sh, st);
			pc.shadow++;
			if (!occluded) light = m::add3(light, m::mul3(S.suncol, throughput));
		}

		ray.o = new_o;
		ray.d = m::add3(fn, random_unit_sphere_vec3<STRICT>(seed, rtl));

		if (debug) {
			switch (S.debug_mode) {
			case DRT_DEBUG_ALBEDO: light = throughput; break;
			case DRT_DEBUG_NORMAL: light = fn; break;
			case DRT_DEBUG_BARYCENTRIC: light = mk3(m::sub(m::sub(1.0f, hit.u), hit.v), hit.u, hit.v); break;
			case DRT_DEBUG_UVS: light = mk3(tex_uv.x, tex_uv.y, 0.f); break;
			case DRT_DEBUG_MESHBVH: {
				float heat = (STATS && st) ? 0.05f * (float)(st->nodes - nodes_before) : 0.f;
				light = mk3(heat, 0.1f + heat, 0.1f + heat);
				break;
			}
			default: break;
			}
			break;
		}
	}

	if (S.render_mode == DRT_RENDER_NORMAL || S.debug_mode == DRT_DEBUG_ALBEDO) {
		if (S.tone_mapping) {
			V3 e = m::scale(light, P.cam.exposure);
			light = m::scale(mk3(tonemap_partial<STRICT>(e.x), tonemap_partial<STRICT>(e.y), tonemap_partial<STRICT>(e.z)), S.white_scale);
		}
		if (S.gamma_correction) light = mk3(m::sqrt(light.x), m::sqrt(light.y), m::sqrt(light.z));
	}
	return light;
}

template <bool STRICT>
DRT_HD V3 post_process(const RenderParams& P, V3 light)
{
	typedef M<STRICT> m;
	const ShadeParams& S = P.shade;
	if (S.render_mode == DRT_RENDER_NORMAL || S.debug_mode == DRT_DEBUG_ALBEDO) {
		if (S.tone_mapping) {
			V3 e = m::scale(light, P.cam.exposure);
			light = m::scale(mk3(tonemap_partial<STRICT>(e.x), tonemap_partial<STRICT>(e.y), tonemap_partial<STRICT>(e.z)), S.white_scale);
		}
		if (S.gamma_correction) light = mk3(m::sqrt(light.x), m::sqrt(light.y), m::sqrt(light.z));
	}
	return light;
}

// All samples of one pixel as a flat state machine with ONE traversal call site: every loop trip
// traces the lane's current ray -- a path segment (closest hit) or a sun shadow ray (any hit) -- then
// advances that lane's path; a lane whose path ends starts its next sample at once instead of
// waiting for the longest path of the warp. Per-lane results are identical to calling
// path_radiance() once per sample (same RNG stream, same operation order, same accumulation order).
// NORMALMODE only (debug views go through path_radiance).
template <bool STRICT, bool STATS>
DRT_HD V3 render_pixel_flat(const RenderParams& P, uint32_t x, uint32_t y, bool rtl, V3 acc, PathCounters& pc, TraceStats* st)
{
	typedef M<STRICT> m;
	const SceneView& sc = P.scene;
	const ShadeParams& S = P.shade;
	const uint32_t pixel_seed = x + y * (uint32_t)P.width;
	uint32_t k = 0;            // sample in this launch
	uint32_t seed = 0;
	int bounce = 0;
	bool shadow_phase = false; // current ray is a sun shadow ray
	V3 light = mk3(0.f, 0.f, 0.f), throughput = mk3(1.f, 1.f, 1.f);
	V3 fn = mk3(0.f, 0.f, 0.f);
	RayOD ray;
	ray.o = mk3(0.f, 0.f, 0.f); ray.d = mk3(0.f, 0.f, 1.f);
	V3 next_o = ray.o;
	if (P.n_samples == 0) return acc;
	seed = pixel_seed * P.frame_first;
	ray = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, seed, rtl);

	while (k < P.n_samples) {
		Hit hit;
		const bool found = trace_ray_mode<STRICT, 2, STATS>(sc, ray.o, ray.d, FLT_MAX, hit, st, shadow_phase);
		bool path_done = false;
		if (shadow_phase) {
			pc.shadow++;
			if (!found) light = m::add3(light, m::mul3(S.suncol, throughput));
			shadow_phase = false;
		}
		else {
			pc.closest++;
			seed += (uint32_t)bounce;
			if (!found) {
				V3 sky = sky_model<STRICT>(ray.d, S.sky_color);
				light = m::add3(light, m::scale(m::mul3(sky, throughput), S.sky_intensity));
				path_done = true;
			}
			else {
				const drt_f4 r1 = ld_f4(sc.tris + 3 * (size_t)hit.prim + 1);
				const uint32_t mat = as_uint(r1.w) & 0x7fffffffu;
				const drt_f4 mrec = ld_f4(sc.materials + mat);
				const int tex = (int)as_uint(mrec.w);
				if (tex < 0) throughput = m::mul3(throughput, mk3(mrec.x, mrec.y, mrec.z));
				else {
					V2 tuv = interp_uv<STRICT>(sc, hit.prim, hit.u, hit.v);
					throughput = m::mul3(throughput, texture_get_pixel<STRICT>(sc, tex, tuv, st));
				}
				const drt_f4 s0 = ld_f4(sc.shade + 3 * (size_t)hit.prim);
				fn = mk3(s0.x, s0.y, s0.z);
				const V3 pos = m::add3(ray.o, m::scale(ray.d, hit.t));
				if (m::dot(fn, m::normalize(ray.d)) > 0.f) fn = m::scale(fn, -1.f);
				next_o = m::add3(pos, m::scale(fn, 0.001f));
				if (S.enable_sunlight) {
					ray.o = next_o;
					ray.d = m::add3(S.sunpos, m::scale(random_unit_vec3<STRICT>(seed, rtl), 1.5f));
					shadow_phase = true;
				}
			}
		}
		if (!shadow_phase && !path_done) {
			// diffuse scatter (RayGen.cuh:133-134), then the loop counter of RayGen.cuh:88
			ray.o = next_o;
			ray.d = m::add3(fn, random_unit_sphere_vec3<STRICT>(seed, rtl));
			bounce++;
			if (bounce > S.bounce_limit) path_done = true;
		}
		if (path_done) {
			acc = m::add3(acc, post_process<STRICT>(P, light));
			k++;
			if (k < P.n_samples) {
				light = mk3(0.f, 0.f, 0.f); throughput = mk3(1.f, 1.f, 1.f);
				bounce = 0;
				seed = pixel_seed * (P.frame_first + k * P.frame_stride);
				ray = camera_get_ray<STRICT>(P.cam, x, y, (uint32_t)P.width, (uint32_t)P.height, seed, rtl);
			}
		}
	}
	return acc;
}

} // namespace drt
