// drt_math.cuh -- scalar/vector fp32 helpers for the path-tracing hot path.
//
// Two arithmetic policies, selected by a template flag:
//   STRICT = true : every multiply/add/sub/div/sqrt is a separately rounded IEEE-754 operation
//                   (__fmul_rn & co. are never contracted into FMA by nvcc). With the operation ORDER
//                   of the reference's helpers (helper_math.cuh:1264-1267 dot, :1436-1439 cross,
//                   :1325-1329 normalize with the host rsqrtf = 1/sqrtf of :78-81) this reproduces the
//                   host-compiled reference bit for bit -- it is the parity mode.
//   STRICT = false: plain operators; nvcc is free to contract a*b+c into FFMA and normalize uses the
//                   hardware rsqrt. This is the production mode (fewer issue slots per ray).
// The header also compiles as plain C++ (DRT_HOSTEMU) for CPU-side logic tests of the device
// functions; there both policies are separately rounded (g++ -ffp-contract=off).
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define DRT_HD __host__ __device__ __forceinline__
#define DRT_D __device__ __forceinline__
#else
#define DRT_HD inline
#define DRT_D inline
#endif

namespace drt {

struct V3 { float x, y, z; };
struct V2 { float x, y; };

DRT_HD V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
DRT_HD V2 mk2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }

DRT_HD float as_float(uint32_t u)
{
#if defined(__CUDA_ARCH__)
	return __uint_as_float(u);
#else
	union { uint32_t u; float f; } c; c.u = u; return c.f;
#endif
}
DRT_HD uint32_t as_uint(float f)
{
#if defined(__CUDA_ARCH__)
	return __float_as_uint(f);
#else
	union { uint32_t u; float f; } c; c.f = f; return c.u;
#endif
}

template <bool STRICT> struct M {
	static DRT_HD float mul(float a, float b)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fmul_rn(a, b);
#endif
		return a * b;
	}
	static DRT_HD float add(float a, float b)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fadd_rn(a, b);
#endif
		return a + b;
	}
	static DRT_HD float sub(float a, float b)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fsub_rn(a, b);
#endif
		return a - b;
	}
	// production mode uses the hardware approximations (MUFU.RCP / MUFU.SQRT, <= 2 ulp), like the reference's
	// Release build (--use_fast_math, premake5.lua:85); parity mode keeps IEEE round-to-nearest
	static DRT_HD float div(float a, float b)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fdiv_rn(a, b);
		// a * MUFU.RCP(b): __fdividef() adds a rescue for |b| outside [2^-126, 2^126] (FSETP + two predicated FMUL per quotient)
		// that no divisor on this path needs (image size, 255, tonemap denominators >= 0.06, |det| >= 1e-6)
		float r;
		asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
		return a * r;
#else
		return a / b;
#endif
	}
	static DRT_HD float sqrt(float a)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fsqrt_rn(a);
		float r;
		// the approximate tonemap quotient can land a few ulp below zero where the exact one is +0: clamp, never NaN
		asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaxf(a, 0.f)));
		return r;
#else
		return sqrtf(a);
#endif
	}
	// reciprocal square root: host helper_math defines rsqrtf(x) = 1.0f / sqrtf(x) (helper_math.cuh:78-81)
	static DRT_HD float rsqrt(float a)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fdiv_rn(1.0f, __fsqrt_rn(a));
		// one MUFU.RSQ: rsqrtf() wraps it in a denormal-input rescue (FSETP + two predicated FMUL) that the squared lengths on
		// this path never need (they are sums of squares of O(1) components; a denormal one would be a zero vector anyway)
		float r;
		asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
		return r;
#else
		return 1.0f / sqrtf(a);
#endif
	}
	static DRT_HD float rcp(float a)
	{
#if defined(__CUDA_ARCH__)
		if (STRICT) return __fdiv_rn(1.0f, a);
		float r;
		asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
		return r;
#else
		return 1.0f / a;
#endif
	}

	static DRT_HD V3 add3(V3 a, V3 b) { return mk3(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
	static DRT_HD V3 sub3(V3 a, V3 b) { return mk3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
	static DRT_HD V3 mul3(V3 a, V3 b) { return mk3(mul(a.x, b.x), mul(a.y, b.y), mul(a.z, b.z)); }
	static DRT_HD V3 scale(V3 a, float s) { return mk3(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
	static DRT_HD V3 divs(V3 a, float s) { return mk3(div(a.x, s), div(a.y, s), div(a.z, s)); }
	// a.x*b.x + a.y*b.y + a.z*b.z, left to right (helper_math.cuh:1264-1267)
	static DRT_HD float dot(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
	// (a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x) (helper_math.cuh:1436-1439)
	static DRT_HD V3 cross(V3 a, V3 b)
	{
		return mk3(sub(mul(a.y, b.z), mul(a.z, b.y)), sub(mul(a.z, b.x), mul(a.x, b.z)), sub(mul(a.x, b.y), mul(a.y, b.x)));
	}
	// v * rsqrt(dot(v,v)) (helper_math.cuh:1325-1329)
	static DRT_HD V3 normalize(V3 v) { return scale(v, rsqrt(dot(v, v))); }
	static DRT_HD float length(V3 v) { return sqrt(dot(v, v)); }
};

DRT_HD float fmin3(float a, float b, float c) { return fminf(fminf(a, b), c); }
DRT_HD float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

} // namespace drt
