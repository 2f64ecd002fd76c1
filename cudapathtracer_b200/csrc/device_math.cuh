// Device arithmetic helpers.
//
// Parity policy: everything that feeds a primary-hit decision or the value of t is written
// with the explicit round-to-nearest intrinsics below, which nvcc never contracts into FMA, so the bits equal
// the reference's host pass (x86-64 without FMA contraction).  Box tests are free to use FMA:
// they only have to be conservative.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace cupt {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }

// ---- exact-op vec3 (reference vec3.h:13-84) ---------------------------------------------------------------
__device__ __forceinline__ float rmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float radd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float rsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float rdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ f3 radd3(f3 a, f3 b) { return mk3(radd(a.x, b.x), radd(a.y, b.y), radd(a.z, b.z)); }
__device__ __forceinline__ f3 rsub3(f3 a, f3 b) { return mk3(rsub(a.x, b.x), rsub(a.y, b.y), rsub(a.z, b.z)); }
__device__ __forceinline__ f3 rmul3(f3 a, float s) { return mk3(rmul(a.x, s), rmul(a.y, s), rmul(a.z, s)); }
__device__ __forceinline__ f3 rdiv3(f3 a, float s) { return mk3(rdiv(a.x, s), rdiv(a.y, s), rdiv(a.z, s)); }
__device__ __forceinline__ float rdot(f3 a, f3 b)  // vec3.h:63-66: (ax*bx + ay*by) + az*bz
{
	return radd(radd(rmul(a.x, b.x), rmul(a.y, b.y)), rmul(a.z, b.z));
}
__device__ __forceinline__ f3 rcross(f3 a, f3 b)   // vec3.h:67-70
{
	return mk3(rsub(rmul(a.y, b.z), rmul(a.z, b.y)), rsub(rmul(a.z, b.x), rmul(a.x, b.z)), rsub(rmul(a.x, b.y), rmul(a.y, b.x)));
}
__device__ __forceinline__ float rlength(f3 v) { return __fsqrt_rn(radd(radd(rmul(v.x, v.x), rmul(v.y, v.y)), rmul(v.z, v.z))); }
__device__ __forceinline__ f3 rnormalized(f3 v)    // vec3.h:58-62
{
	float len = rlength(v);
	return mk3(rdiv(v.x, len), rdiv(v.y, len), rdiv(v.z, len));
}

// ---- Philox4x32-10 (Salmon, Moraes, Dror, Shaw, SC'11), counter/key layout of cuRAND's Philox state ----------
// stream (seed; x, y, sample): counter = (draw >> 2, 0, y << 16 | x, sample), key = seed; draw & 3 selects the word.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
	for (int r = 0; r < 10; ++r) {
		const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
		const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
		c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	return c;
}
// cuRAND's uniform mapping: x * 2^-32 + 2^-33, in (0, 1] (the product is exact, so FMA contraction is harmless)
__device__ __forceinline__ float u32_to_uniform(uint32_t x) { return radd(rmul((float)x, 2.3283064e-10f), 2.3283064e-10f / 2.0f); }

struct PhiloxStream {
	uint32_t k0, k1, sub_lo, sub_hi;
	uint32_t block;
	uint4 out;
	__device__ __forceinline__ void init(uint64_t seed, uint32_t x, uint32_t y, uint32_t sample)
	{
		k0 = (uint32_t)seed; k1 = (uint32_t)(seed >> 32); sub_lo = (y << 16) | x; sub_hi = sample; block = 0xffffffffu;
	}
	__device__ __forceinline__ float draw(uint32_t index)   // index-th uniform of the stream
	{
		const uint32_t b = index >> 2;
		if (b != block) { out = philox4x32_10(make_uint4(b, 0u, sub_lo, sub_hi), k0, k1); block = b; }
		const uint32_t w = index & 3u;
		const uint32_t v = w == 0 ? out.x : (w == 1 ? out.y : (w == 2 ? out.z : out.w));
		return u32_to_uniform(v);
	}
};

// ---- packed single precision (sm_100: fma.rn.f32x2 / add.rn.f32x2, SASS FFMA2 / FADD2) ----------------------------
// One instruction works on the two f32 halves of a 64-bit register pair; each half rounds exactly like the scalar
// fma / add, so a packed evaluation gives the bits of the scalar one in half the issue slots.  A pair whose halves are
// the same scalar costs nothing: the SASS forms take a broadcast operand (`R.F32`).
#ifndef CUPT_HOST_EMUL
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo2(f32x2 v) { return __uint_as_float((uint32_t)v); }           // (register-pair halves: no instruction)
__device__ __forceinline__ float hi2(f32x2 v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
#endif

// ---- 128-bit read-only fetches -----------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_nc_u4(const void* p)
{
#ifdef CUPT_HOST_EMUL   // tests/emul only: the traversal logic compiled for the host to check the builder
	return *reinterpret_cast<const uint4*>(p);
#else
	uint4 v;
	asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
#endif
}
__device__ __forceinline__ float4 ldg_nc_f4(const void* p)
{
#ifdef CUPT_HOST_EMUL
	return *reinterpret_cast<const float4*>(p);
#else
	float4 v;
	asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
	return v;
#endif
}

}  // namespace cupt
