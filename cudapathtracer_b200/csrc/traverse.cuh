// Closest-hit traversal: the replacement of trace() (reference kernel.cu:112-161).
//
//  * traverse_wide  — ordered, culled traversal of the 80-byte 8-wide quantised BVH (layout.h) with a per-thread
//                     stack in shared memory.  Returns argmin t over every triangle the ray can hit with
//                     0 < t < 1e5 (kernel.cu:117,127), equal t resolved by `rank` (the reference's left-first
//                     leaf order), so it equals the reference's result whenever the reference's own
//                     non-conservative slab test (BVH.h:65,79) did not drop the winning triangle.
//  * traverse_bvh2_ref — the reference's algorithm op for op on its own 32-byte node array: same slab
//                     arithmetic (6 IEEE divisions per node, no t range), same left-first order, same strict <.
//                     Parity mode / drop-in for a caller that built its BVH with buildBVH().
#pragma once
#include "device_math.cuh"
#include "layout.h"
#include "../../include/cupt.h"

namespace cupt {

struct Hit {
	float t;
	int32_t id;
	uint32_t rank;
};

// smallest float f with (double)f >= 0.00001: `abs(a) < 0.00001` (modelLoader.h:64) compares the float against a
// double literal; |a| < 0.00001 (double)  <=>  |a| < kParallelEps (float).  Pinned by tests/test_abi.py.
#define CUPT_PARALLEL_EPS_BITS 0x3727C5ADu

// triIntersect (modelLoader.h:49-82) on a pre-gathered record: a = v0, b = v1 - v0, c = v2 - v0
__device__ __forceinline__ float tri_refexact(const f3 o, const f3 d, const float4 r0, const float4 r1, const float4 r2)
{
	const f3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r0.w, r1.x, r1.y), e2 = mk3(r1.z, r1.w, r2.x);
	const f3 q = rcross(d, e2);
	const float a = rdot(e1, q);
	if (fabsf(a) < __uint_as_float(CUPT_PARALLEL_EPS_BITS))
		return CUPT_MAX_FLOAT;
	const f3 s = rdiv3(rsub3(o, v0), a);
	const f3 r = rcross(s, e1);
	const float b0 = rdot(s, q), b1 = rdot(r, d), b2 = rsub(rsub(1.0f, b0), b1);
	if (b0 < 0.0f || b1 < 0.0f || b2 < 0.0f)
		return CUPT_MAX_FLOAT;
	return rdot(e2, r);
}

// analytic ray-sphere (no reference code exists, sphere.h:7-13 is a POD only: parity unpinned)
__device__ __forceinline__ float sphere_nearest(const f3 o, const f3 d, const float4 r0, const float4 r1)
{
	const float cx = r0.x - o.x, cy = r0.y - o.y, cz = r0.z - o.z, rad = r0.w;
	const float dd = d.x * d.x + d.y * d.y + d.z * d.z;
	const float b = cx * d.x + cy * d.y + cz * d.z;
	// numerically robust discriminant (distance from the centre to the ray, squared)
	const float px = cx - (b / dd) * d.x, py = cy - (b / dd) * d.y, pz = cz - (b / dd) * d.z;
	const float disc = rad * rad - (px * px + py * py + pz * pz);
	if (!(disc >= 0.0f))
		return CUPT_MAX_FLOAT;
	const float root = sqrtf(disc * dd);
	const float qv = b + copysignf(root, b);
	const float cc = cx * cx + cy * cy + cz * cz - rad * rad;
	float t0 = cc / qv, t1 = qv / dd;
	if (t0 > t1) { const float tt = t0; t0 = t1; t1 = tt; }
	(void)r1;
	if (t0 > 0.0f) return t0;
	if (t1 > 0.0f) return t1;
	return CUPT_MAX_FLOAT;
}

__device__ __forceinline__ float u8f(uint32_t w, int j) { return (float)((w >> (8 * j)) & 0xffu); }

// stack entry i of this thread lives at stack[i * stride]
template <bool STATS>
__device__ __forceinline__ void traverse_wide(const uint4* __restrict__ nodes, const float4* __restrict__ prims,
                                              const f3 o, const f3 d, Hit& hit, uint2* stack, const int stride,
                                              const int stack_cap, uint32_t& n_nodes, uint32_t& n_tris)
{
	const float big = 1.0e30f;
	const float ix = fabsf(d.x) > 1.0e-30f ? 1.0f / d.x : copysignf(big, d.x);
	const float iy = fabsf(d.y) > 1.0e-30f ? 1.0f / d.y : copysignf(big, d.y);
	const float iz = fabsf(d.z) > 1.0e-30f ? 1.0f / d.z : copysignf(big, d.z);
	const bool negx = ix < 0.0f, negy = iy < 0.0f, negz = iz < 0.0f;
	const uint32_t oct_inv = (negx ? 0u : 4u) | (negy ? 0u : 2u) | (negz ? 0u : 1u);
	const uint32_t oct_inv4 = oct_inv * 0x01010101u;

	float best = CUPT_MAX_FLOAT;
	int32_t best_id = -1;
	uint32_t best_rank = 0u;   // rank < 0 never holds: a tie can only replace an existing hit

	uint2 ngroup = make_uint2(0u, 0x80000000u);
	uint2 tgroup = make_uint2(0u, 0u);
	int sp = 0;

	for (;;) {
		if (ngroup.y > 0x00ffffffu) {
			const uint32_t hits = ngroup.y;
			const uint32_t bit = 31u - __clz(hits);
			ngroup.y &= ~(1u << bit);
			if (ngroup.y > 0x00ffffffu) {
				if (sp < stack_cap) stack[sp * stride] = ngroup;
				++sp;   // entries beyond the capacity are dropped only if the builder's depth bound is violated (checked on upload)
			}
			const uint32_t slot = (bit - 24u) ^ oct_inv;
			const uint32_t rel = __popc(hits & 0xffu & ~(0xffffffffu << slot));
			const uint4* np = nodes + (size_t)(ngroup.x + rel) * 5u;
			const uint4 n0 = ldg_nc_u4(np + 0), n1 = ldg_nc_u4(np + 1), n2 = ldg_nc_u4(np + 2), n3 = ldg_nc_u4(np + 3),
			            n4 = ldg_nc_u4(np + 4);
			if (STATS) ++n_nodes;
			const float sx = __uint_as_float((n0.w & 0xffu) << 23) * ix;
			const float sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * iy;
			const float sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * iz;
			const float ox = (__uint_as_float(n0.x) - o.x) * ix;
			const float oy = (__uint_as_float(n0.y) - o.y) * iy;
			const float oz = (__uint_as_float(n0.z) - o.z) * iz;
			// rounding slack of t = q * s + o', both sides (2 * 2^-21 relative to the magnitudes involved)
			const float slack = fmaxf(fmaxf(fmaf(255.0f, fabsf(sx), fabsf(ox)), fmaf(255.0f, fabsf(sy), fabsf(oy))),
			                          fmaf(255.0f, fabsf(sz), fabsf(oz))) * 9.5367431640625e-7f;
			const uint32_t imask = n0.w >> 24;
			uint32_t hitmask = 0u;
#pragma unroll
			for (int h = 0; h < 2; ++h) {
				const uint32_t meta4 = h ? n1.w : n1.z;
				const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
				const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
				const uint32_t bit_index4 = (meta4 ^ (oct_inv4 & inner_mask4)) & 0x1f1f1f1fu;
				const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
				const uint32_t qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
				const uint32_t qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
				const uint32_t nx = negx ? qhx : qlx, fx = negx ? qlx : qhx;
				const uint32_t ny = negy ? qhy : qly, fy = negy ? qly : qhy;
				const uint32_t nz = negz ? qhz : qlz, fz = negz ? qlz : qhz;
#pragma unroll
				for (int j = 0; j < 4; ++j) {
					const float tnx = fmaf(u8f(nx, j), sx, ox), tny = fmaf(u8f(ny, j), sy, oy), tnz = fmaf(u8f(nz, j), sz, oz);
					const float tfx = fmaf(u8f(fx, j), sx, ox), tfy = fmaf(u8f(fy, j), sy, oy), tfz = fmaf(u8f(fz, j), sz, oz);
					const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
					const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, best));
					if (tmin <= tmax + slack) {
						const uint32_t cb = (child_bits4 >> (8 * j)) & 0xffu, bi = (bit_index4 >> (8 * j)) & 0xffu;
						hitmask |= cb << bi;
					}
				}
			}
			ngroup = make_uint2(n1.x, (hitmask & 0xff000000u) | imask);
			tgroup = make_uint2(n1.y, hitmask & 0x00ffffffu);
		} else {
			tgroup = ngroup;
			ngroup = make_uint2(0u, 0u);
		}

		while (tgroup.y != 0u) {
			const uint32_t k = __ffs(tgroup.y) - 1u;
			tgroup.y &= tgroup.y - 1u;
			const float4* pp = prims + (size_t)(tgroup.x + k) * 3u;
			const float4 r0 = ldg_nc_f4(pp + 0), r1 = ldg_nc_f4(pp + 1), r2 = ldg_nc_f4(pp + 2);
			if (STATS) ++n_tris;
			const uint32_t kind = __float_as_uint(r2.w);
			float t;
			if (kind == 0u) t = tri_refexact(o, d, r0, r1, r2);
			else t = sphere_nearest(o, d, r0, r1);
			const uint32_t rank = __float_as_uint(r2.z);
			if (0.0f < t && (t < best || (t == best && rank < best_rank))) {
				best = t; best_rank = rank;
				best_id = (int32_t)(__float_as_uint(r2.y) | (kind << 31));
			}
		}

		if (ngroup.y <= 0x00ffffffu) {
			if (sp == 0) break;
			--sp;
			ngroup = stack[sp * stride];
		}
	}
	hit.t = best; hit.id = best_id; hit.rank = best_rank;
}

// trace(), kernel.cu:112-161, with rayAABBIntersect (BVH.h:51-83) and triIntersect (modelLoader.h:49-82) in the
// reference's arithmetic; the visit counters (kernel.cu:133) are not kept.
template <bool STATS>
__device__ __forceinline__ void traverse_bvh2_ref(const cupt_bvh2_node* __restrict__ nodes, const float* __restrict__ verts,
                                                  const cupt_triangle* __restrict__ tris, const f3 o, const f3 d, Hit& hit,
                                                  uint32_t& n_nodes, uint32_t& n_tris)
{
	uint32_t stack[CUPT_MAX_BVH2_DEPTH];
	stack[0] = 0u;
	float closestT = CUPT_MAX_FLOAT;
	int32_t trisID = -1;
	int i = 0;
	while (i >= 0) {
		const uint32_t e = stack[i];
		if (e & CUPT_BVH_LEAF_FLAG) {
			const uint32_t id = e ^ CUPT_BVH_LEAF_FLAG;
			const cupt_triangle tr = tris[id];
			const f3 v0 = mk3(verts[3 * tr.v0], verts[3 * tr.v0 + 1], verts[3 * tr.v0 + 2]);
			const f3 v1 = mk3(verts[3 * tr.v1], verts[3 * tr.v1 + 1], verts[3 * tr.v1 + 2]);
			const f3 v2 = mk3(verts[3 * tr.v2], verts[3 * tr.v2 + 1], verts[3 * tr.v2 + 2]);
			const f3 e1 = rsub3(v1, v0), e2 = rsub3(v2, v0);
			const float t = tri_refexact(o, d, make_float4(v0.x, v0.y, v0.z, e1.x), make_float4(e1.y, e1.z, e2.x, e2.y),
			                             make_float4(e2.z, 0.f, 0.f, 0.f));
			if (STATS) ++n_tris;
			if (0.0f < t && t < closestT) { closestT = t; trisID = (int32_t)id; }
			--i;
		} else {
			const cupt_bvh2_node nd = nodes[e];
			if (STATS) ++n_nodes;
			bool b = true;
			float tmin = rdiv(rsub(nd.lo.x, o.x), d.x), tmax = rdiv(rsub(nd.hi.x, o.x), d.x);
			if (tmin > tmax) { const float s = tmin; tmin = tmax; tmax = s; }
			float tymin = rdiv(rsub(nd.lo.y, o.y), d.y), tymax = rdiv(rsub(nd.hi.y, o.y), d.y);
			if (tymin > tymax) { const float s = tymin; tymin = tymax; tymax = s; }
			if ((tmin > tymax) || (tymin > tmax)) b = false;
			if (b) {
				if (tymin > tmin) tmin = tymin;
				if (tymax < tmax) tmax = tymax;
				float tzmin = rdiv(rsub(nd.lo.z, o.z), d.z), tzmax = rdiv(rsub(nd.hi.z, o.z), d.z);
				if (tzmin > tzmax) { const float s = tzmin; tzmin = tzmax; tzmax = s; }
				if ((tmin > tzmax) || (tzmin > tmax)) b = false;
			}
			if (b) { stack[i] = nd.right; stack[i + 1] = nd.left; ++i; }
			else --i;
		}
	}
	hit.t = closestT; hit.id = trisID; hit.rank = 0u;
}

}  // namespace cupt
