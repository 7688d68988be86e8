// drt_rng.cuh -- per-path RNG and primary-ray generation.
// Follows the reference's stream exactly (one u32 state per path, advanced by pcg_hash per draw):
//   pcg_hash / randomFloat / randomUnitVec3 / randomUnitSphereVec3 / random_in_unit_disk
//     DustRayTracer/src/Core/CudaMath/Random.cu:6-17,42-66
//   Camera::GetRay   DustRayTracer/src/Core/Scene/Camera.cu:82-123 (pixel-independent terms are in CameraParams)
#pragma once
#include "drt_types.h"

namespace drt {

DRT_HD uint32_t pcg_hash(uint32_t input)
{
	uint32_t state = input * 747796405u + 2891336453u;
	uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
	return (word >> 22u) ^ word;
}

// (float)seed / (float)UINT32_MAX: the divisor rounds to 2^32, so the quotient is an exact scaling.
DRT_HD float random_float(uint32_t& seed)
{
	seed = pcg_hash(seed);
	return (float)seed * 2.3283064365386963e-10f;
}

// Random.cu:42-48 draws the three components as ARGUMENTS of make_float3, whose evaluation order the
// language leaves open: nvcc device code evaluates them left to right, g++ (the host-compiled oracle)
// right to left. `rtl` selects the oracle's order so that parity runs see the identical stream.
// randomFloat(seed) * 2 - 1. Production mode folds the exact scalings into one fused multiply-add.
template <bool STRICT>
DRT_HD float random_signed(uint32_t& seed)
{
	typedef M<STRICT> m;
	if (STRICT) return m::sub(m::mul(random_float(seed), 2.f), 1.f);
	seed = pcg_hash(seed);
	return fmaf((float)seed, 4.6566128730773926e-10f, -1.f);
}

template <bool STRICT>
DRT_HD V3 random_unit_vec3(uint32_t& seed, bool rtl)
{
	typedef M<STRICT> m;
	float a = random_signed<STRICT>(seed);
	float b = random_signed<STRICT>(seed);
	float c = random_signed<STRICT>(seed);
	V3 v = rtl ? mk3(c, b, a) : mk3(a, b, c);
	return m::normalize(v);
}

// Random.cu:50-58: rejection on |p|^2 < 1 of an already normalised p, i.e. a coin flip on rounding.
template <bool STRICT>
DRT_HD V3 random_unit_sphere_vec3(uint32_t& seed, bool rtl)
{
	typedef M<STRICT> m;
	for (;;) {
		V3 p = random_unit_vec3<STRICT>(seed, rtl);
		if (STRICT) {
			float len = m::length(p);
			if (m::mul(len, len) < 1.f) return p;
		}
		else if (m::dot(p, p) < 1.f) return p; // with a correctly rounded sqrt, sqrt(x)^2 < 1 <=> x < 1: no MUFU.SQRT needed
	}
}

template <bool STRICT>
DRT_HD V2 random_in_unit_disk(uint32_t& seed, bool rtl)
{
	typedef M<STRICT> m;
	for (;;) {
		float a = m::sub(m::mul(random_float(seed), 2.f), 1.f);
		float b = m::sub(m::mul(random_float(seed), 2.f), 1.f);
		V2 p = rtl ? mk2(b, a) : mk2(a, b);
		if (m::add(m::mul(p.x, p.x), m::mul(p.y, p.y)) < 1.0f) return p;
	}
}

struct RayOD {
	V3 o, d;
};

// Primary ray of pixel (x, y) for the RNG state `seed` (RayGen.cuh:65-66,77; Camera.cu:98-122).
template <bool STRICT>
DRT_HD RayOD camera_get_ray(const CameraParams& cam, uint32_t x, uint32_t y, uint32_t width, uint32_t height,
	uint32_t& seed, bool rtl)
{
	typedef M<STRICT> m;
	// screen_uv = (x/W, y/H) * 2 - 1
	float u = m::sub(m::mul(m::div((float)x, (float)width), 2.f), 1.f);
	float v = m::sub(m::mul(m::div((float)y, (float)height), 2.f), 1.f);
	// braced initialiser: left-to-right on every compiler
	float ox = m::sub(random_float(seed), 0.5f);
	float oy = m::sub(random_float(seed), 0.5f);
	ox = m::mul(ox, 0.0035f);
	oy = m::mul(oy, 0.0035f);
	V3 rorig = cam.position;
	if (!(cam.defocus_angle <= 0.f)) {
		V2 p = random_in_unit_disk<STRICT>(seed, rtl);
		rorig = m::add3(m::add3(cam.position, m::scale(cam.defocus_u, p.x)), m::scale(cam.defocus_v, p.y));
	}
	V3 dir = m::add3(m::sub3(m::add3(m::add3(cam.forward_focus, m::scale(cam.plane_h, m::add(u, ox))),
		m::scale(cam.plane_v, m::add(v, oy))), rorig), cam.position);
	RayOD r;
	r.o = rorig;
	r.d = m::normalize(dir);
	return r;
}

} // namespace drt
