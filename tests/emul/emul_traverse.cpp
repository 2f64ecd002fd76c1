// TEST INFRASTRUCTURE (CPU suite): compiles the product's traversal header (csrc/traverse.cuh) for the host with
// a shim for the CUDA intrinsics it uses, so that the wide-BVH builder and the traversal logic (slot order,
// quantisation, tie-break, stack discipline) can be checked against the oracle without a GPU.  This is a
// structural test of host-built data; it is never used by the product and proves nothing about the kernels'
// device arithmetic (the GPU tests do that).
#define CUPT_HOST_EMUL 1
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#undef __device__
#undef __forceinline__
#define __device__
#define __forceinline__ inline
static inline float __fmul_rn(float a, float b) { return a * b; }   // built with -ffp-contract=off
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline int __clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static inline int __popc(uint32_t x) { return __builtin_popcount(x); }
static inline int __ffs(uint32_t x) { return __builtin_ffs((int)x); }
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s) { return (uint32_t)(((((uint64_t)hi << 32) | lo) << (s & 31u)) >> 32); }
#include "../../cudapathtracer_b200/csrc/traverse.cuh"

using namespace cupt;

extern "C" int emul_trace_wide(const void* nodes80, const void* prims48, int depth, const float* rays6, long long n,
                               int32_t* ids, float* ts, unsigned long long* counts)
{
	unsigned long long nn = 0, nt = 0;
	int max_sp_ok = 1;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nn, nt)
	for (long long i = 0; i < n; ++i) {
		std::vector<uint2> stack((size_t)depth + 1);
		Hit h;
		uint32_t a = 0, b = 0;
		traverse_wide<true>((const uint4*)nodes80, (const float4*)prims48, mk3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]),
		                    mk3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]), h, stack.data(), 1, depth, a, b);
		ids[i] = h.id; ts[i] = h.t;
		nn += a; nt += b;
	}
	if (counts) { counts[0] = nn; counts[1] = nt; }
	return max_sp_ok ? 0 : -1;
}

// the same rays through the hot kernel's node step (wide_node_step_fast: sign-bit hit mask, order table, sparse
// primitive groups) with the two look-up tables built the way k_extend builds them
extern "C" int emul_trace_wide_fast(const void* nodes80, const void* prims48, int depth, const float* rays6, long long n,
                                    int32_t* ids, float* ts, unsigned long long* counts)
{
	std::vector<uint8_t> perm(8 * 256);
	std::vector<uint32_t> nib(256);
	for (uint32_t i = 0; i < 8u * 256u; ++i) {
		uint32_t v = 0u;
		for (uint32_t sl = 0; sl < 8u; ++sl)
			if ((i >> sl) & 1u) v |= 1u << (sl ^ (i >> 8));
		perm[i] = (uint8_t)v;
	}
	for (uint32_t i = 0; i < 256u; ++i) {
		uint32_t v = 0u;
		for (uint32_t sl = 0; sl < 8u; ++sl)
			if ((i >> sl) & 1u) v |= 7u << (4u * sl);
		nib[i] = v;
	}
	unsigned long long nn = 0, nt = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nn, nt)
	for (long long i = 0; i < n; ++i) {
		std::vector<uint2> stack((size_t)depth + 1);
		Hit h;
		uint32_t a = 0, b = 0;
		traverse_wide_fast<true>((const uint4*)nodes80, (const float4*)prims48, mk3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]),
		                         mk3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]), h, stack.data(), 1, depth, perm.data(), nib.data(), a, b);
		ids[i] = h.id; ts[i] = h.t;
		nn += a; nt += b;
	}
	if (counts) { counts[0] = nn; counts[1] = nt; }
	return 0;
}

// the loop of traverse_wide_fast once more, counting the node steps that find NOTHING (no inner child and no leaf record hit):
// counts = {node steps, primitive tests, empty node steps, node steps in nodes that CONTAIN the ray origin, empty ones among those}
extern "C" int emul_trace_wide_count_empty(const void* nodes80, const void* prims48, int depth, const float* rays6, long long n, unsigned long long* counts)
{
	std::vector<uint8_t> perm(8 * 256);
	std::vector<uint32_t> nib(256);
	for (uint32_t i = 0; i < 8u * 256u; ++i) { uint32_t v = 0u; for (uint32_t sl = 0; sl < 8u; ++sl) if ((i >> sl) & 1u) v |= 1u << (sl ^ (i >> 8)); perm[i] = (uint8_t)v; }
	for (uint32_t i = 0; i < 256u; ++i) { uint32_t v = 0u; for (uint32_t sl = 0; sl < 8u; ++sl) if ((i >> sl) & 1u) v |= 7u << (4u * sl); nib[i] = v; }
	unsigned long long nn = 0, nt = 0, ne = 0, no = 0, neo = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nn, nt, ne, no, neo)
	for (long long i = 0; i < n; ++i) {
		std::vector<uint2> stack((size_t)depth + 1);
		WideRay r;
		wide_ray_init(r, mk3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]), mk3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]));
		uint2 ngroup = make_uint2(0u, 0x80000000u);
		int sp = 0;
		for (;;) {
			uint2 rest, tgroup;
			// which node will this step fetch (same arithmetic as the step): is the ray's ORIGIN inside the union of its child boxes?
			bool inside;
			{
				const uint32_t hits = ngroup.y, bit = 31u - __clz(hits), slot = (bit - 24u) ^ r.oct_inv;
				const uint32_t rel = __popc(hits & 0xffu & ~(0xffffffffu << slot));
				const Node8& nd = ((const Node8*)nodes80)[ngroup.x + rel];
				const float o3[3] = {r.o.x, r.o.y, r.o.z};
				inside = true;
				for (int a = 0; a < 3; ++a) {
					int lo = 255, hi = 0;
					for (int sl = 0; sl < 8; ++sl) if (((nd.imask >> sl) & 1u) || ((nd.pvalid >> (4 * sl)) & 7u)) { lo = std::min(lo, (int)nd.qlo[a][sl]); hi = std::max(hi, (int)nd.qhi[a][sl]); }
					const float sc = __uint_as_float((uint32_t)nd.e[a] << 23);
					if (o3[a] < nd.p[a] + lo * sc - 1e-3f || o3[a] > nd.p[a] + hi * sc + 1e-3f) inside = false;
				}
			}
			wide_node_step_fast((const uint4*)nodes80, r, ngroup, rest, tgroup, CUPT_QBIAS_BITS, perm.data() + (r.oct_inv << 8), nib.data());
			if (group_has_node(rest)) stack[(size_t)sp++] = rest;
			++nn;
			if (inside) ++no;
			if (!group_has_node(ngroup) && tgroup.y == 0u) { ++ne; if (inside) ++neo; }
			if (tgroup.y != 0u) {
				const uint2 tv = node_records((const uint4*)nodes80, tgroup.x);
				while (tgroup.y != 0u) {
					const uint32_t k = __ffs(tgroup.y) - 1u;
					tgroup.y &= tgroup.y - 1u;
					uint2 one = make_uint2(sparse_record(tv, k), 1u);
					wide_prim_step((const float4*)prims48, r, one);
					++nt;
				}
			}
			if (!group_has_node(ngroup)) {
				if (sp == 0) break;
				ngroup = stack[(size_t)--sp];
			}
		}
	}
	counts[0] = nn; counts[1] = nt; counts[2] = ne; counts[3] = no; counts[4] = neo;
	return 0;
}

extern "C" int emul_trace_bvh2(const void* nodes32, const float* verts, const void* tris28, const float* rays6, long long n,
                               int32_t* ids, float* ts)
{
#pragma omp parallel for schedule(dynamic, 256)
	for (long long i = 0; i < n; ++i) {
		Hit h;
		uint32_t a = 0, b = 0;
		traverse_bvh2_ref<false>((const cupt_bvh2_node*)nodes32, verts, (const cupt_triangle*)tris28,
		                         mk3(rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]),
		                         mk3(rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]), h, a, b);
		ids[i] = h.id; ts[i] = h.t;
	}
	return 0;
}
