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

// Watertight ray-triangle test after Woop, Benthin, Wald, "Watertight Ray/Triangle Intersection" (JCGT 2013), for
// records of kind 2 (a = v0, b = v1, c = v2: the exact vertices, not edges, so that two triangles sharing an edge see
// the same numbers).  CUPT_TRI_WATERTIGHT only: NOT the reference's test (modelLoader.h:49-82 lets 5 of 1 048 576
// primary rays of config 2 slip between triangles), so this mode is outside the bit-exact parity statement.
// Every product and difference is individually rounded (an FMA would make the edge function of a shared edge
// differ between its two triangles); an edge function that comes out as exactly 0 is re-evaluated in double.
__device__ __forceinline__ float tri_watertight(const f3 o, const f3 d, const float4 r0, const float4 r1, const float4 r2)
{
	const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
	const int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
	const float dzc = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
	const bool swp = dzc < 0.0f;   // keep the winding when the dominant component is negative
#define CUPT_PERM(v, X, Y, Z)                                                   \
	{                                                                           \
		const float c0 = kz == 0 ? (v).y : (kz == 1 ? (v).z : (v).x);           \
		const float c1 = kz == 0 ? (v).z : (kz == 1 ? (v).x : (v).y);           \
		X = swp ? c1 : c0; Y = swp ? c0 : c1;                                   \
		Z = kz == 0 ? (v).x : (kz == 1 ? (v).y : (v).z);                        \
	}
	float dx, dy, dz;
	CUPT_PERM(d, dx, dy, dz)
	const float Sx = rdiv(dx, dz), Sy = rdiv(dy, dz), Sz = rdiv(1.0f, dz);
	const f3 A = rsub3(mk3(r0.x, r0.y, r0.z), o), B = rsub3(mk3(r0.w, r1.x, r1.y), o), C = rsub3(mk3(r1.z, r1.w, r2.x), o);
	float Akx, Aky, Akz, Bkx, Bky, Bkz, Ckx, Cky, Ckz;
	CUPT_PERM(A, Akx, Aky, Akz)
	CUPT_PERM(B, Bkx, Bky, Bkz)
	CUPT_PERM(C, Ckx, Cky, Ckz)
#undef CUPT_PERM
	const float Ax = rsub(Akx, rmul(Sx, Akz)), Ay = rsub(Aky, rmul(Sy, Akz));
	const float Bx = rsub(Bkx, rmul(Sx, Bkz)), By = rsub(Bky, rmul(Sy, Bkz));
	const float Cx = rsub(Ckx, rmul(Sx, Ckz)), Cy = rsub(Cky, rmul(Sy, Ckz));
	float U = rsub(rmul(Cx, By), rmul(Cy, Bx));
	float V = rsub(rmul(Ax, Cy), rmul(Ay, Cx));
	float W = rsub(rmul(Bx, Ay), rmul(By, Ax));
	if (U == 0.0f || V == 0.0f || W == 0.0f) {
		U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
		V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
		W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
	}
	if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f))
		return CUPT_MAX_FLOAT;
	const float det = radd(radd(U, V), W);
	if (det == 0.0f)
		return CUPT_MAX_FLOAT;
	const float Az = rmul(Sz, Akz), Bz = rmul(Sz, Bkz), Cz = rmul(Sz, Ckz);
	const float T = radd(radd(rmul(U, Az), rmul(V, Bz)), rmul(W, Cz));
	return rdiv(T, det);   // <= 0 (behind the origin) is rejected by the caller's 0 < t
}

// analytic ray-sphere, nearest root with 0 < t (no reference code exists, sphere.h:7-13 is a POD only: PARITY
// UNPINNED; the test-side CPU restatement repeats exactly this arithmetic).  Every operation is individually rounded.
//   c = centre - o, dd = d.d, b = c.d, p = c - (b/dd) d (closest approach), disc = r^2 - p.p,
//   q = b + sign(b) sqrt(disc dd), roots (c.c - r^2)/q and q/dd
__device__ __forceinline__ float sphere_nearest(const f3 o, const f3 d, const float4 r0, const float4 r1)
{
	const f3 c = rsub3(mk3(r0.x, r0.y, r0.z), o);
	const float rad = r0.w;
	const float dd = rdot(d, d), b = rdot(c, d);
	const float k = rdiv(b, dd);
	const f3 p = rsub3(c, rmul3(d, k));
	const float disc = rsub(rmul(rad, rad), rdot(p, p));
	if (!(disc >= 0.0f))
		return CUPT_MAX_FLOAT;
	const float root = __fsqrt_rn(rmul(disc, dd));
	const float qv = radd(b, copysignf(root, b));
	const float cc = rsub(rdot(c, c), rmul(rad, rad));
	float t0 = rdiv(cc, qv), t1 = rdiv(qv, dd);
	if (t0 > t1) { const float tt = t0; t0 = t1; t1 = tt; }
	(void)r1;
	if (t0 > 0.0f) return t0;
	if (t1 > 0.0f) return t1;
	return CUPT_MAX_FLOAT;
}

// Byte j of w as the float 2^15 + b without I2F (which runs on the quarter-rate XU pipe and was the top stall source
// of the first profile, profiles/r1a_extend_opcodes.md): one PRMT drops the byte into mantissa bits 8..15 of 2^15.
// The constant 2^15 is folded into the per-node offset (o' = o - 2^15 s), which costs 2^-9 of a grid cell of
// accuracy and is covered by the per-axis slack below.
#define CUPT_QBIAS 32768.0f
// `bias_bits` must hold 0x47000000 (the bits of 2^15) but come from a kernel parameter, not a literal: with a literal
// ptxas keeps the constant as PRMT's immediate and re-materialises the SELECTOR in a register before almost every
// PRMT (60 extra IMAD/MOV per node step in profiles/r1d); with an opaque register the selector is the immediate.
#define CUPT_QBIAS_BITS 0x47000000u
__device__ __forceinline__ float u8f_biased(uint32_t w, int j, uint32_t bias_bits)
{
#ifdef CUPT_HOST_EMUL
	(void)bias_bits;
	return CUPT_QBIAS + (float)((w >> (8 * j)) & 0xffu);
#else
	return __uint_as_float(__byte_perm(w, bias_bits, 0x7604u + ((uint32_t)j << 4)));
#endif
}

// 1: the child tests of wide_node_step_fast with sm_100's packed fma.rn.f32x2 / add.rn.f32x2 (FFMA2 / FADD2): same bits, 28 fewer
// instructions per node step — and 0.85 ms per config 2 frame SLOWER (34.35 against 33.50 ms of extend, same box, profiles/
// r2_extend_variants.md): FFMA2 has the FMA pipe for twice as long as FFMA (tools/ffma2_probe.cu: 37.1 against 35.9 TFMA/s), so it
// only saves issue slots, and the child test is bound by the ALU pipe (6 PRMT + 2 FMNMX3 + LOP3 + SHF per child), not by issue.
#ifndef CUPT_NODE_F32X2
#define CUPT_NODE_F32X2 0
#endif

// Per-ray traversal state of the wide BVH (registers).
struct WideRay {
	f3 o, d;
	float ix, iy, iz;      // 1/d, +-1e20 for a (near-)zero component
	uint32_t oct_inv;      // bit 2/1/0 set when d.x/y/z is non-negative
	float best;            // current closest t (kernel.cu:117: starts at MAX_FLOAT)
	int32_t best_id;
	uint32_t best_rank;    // rank < 0 never holds: a tie can only replace an existing hit
};

// 1/x for the box tests only: MUFU.RCP (1 ulp) instead of the ~10-instruction IEEE division; the error scales every
// plane distance of an axis by (1 + 2^-23) at most, a quarter of the 2^-21 relative slack of the node step
__device__ __forceinline__ float box_rcp(float x)
{
#ifdef CUPT_HOST_EMUL
	return 1.0f / x;
#else
	float y;
	asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
	return y;
#endif
}

__device__ __forceinline__ void wide_ray_init(WideRay& r, const f3 o, const f3 d)
{
	const float big = 1.0e20f;   // keeps 2^15 * scale / d finite for every scene inside the reference's +-1e4 world (BVH.h:21-25)
	r.o = o; r.d = d;
	r.ix = fabsf(d.x) > 1.0e-20f ? box_rcp(d.x) : copysignf(big, d.x);
	r.iy = fabsf(d.y) > 1.0e-20f ? box_rcp(d.y) : copysignf(big, d.y);
	r.iz = fabsf(d.z) > 1.0e-20f ? box_rcp(d.z) : copysignf(big, d.z);
	r.oct_inv = (r.ix < 0.0f ? 0u : 4u) | (r.iy < 0.0f ? 0u : 2u) | (r.iz < 0.0f ? 0u : 1u);
	r.best = CUPT_MAX_FLOAT; r.best_id = -1; r.best_rank = 0u;
}

// A "group" is {base index, bit mask}: node group = {child_base, hits << 24 | imask} (hit bits in front-to-back
// order, highest first), primitive group = {tri_base, one bit per 48-byte record}; a group whose mask has no bit
// above 23 is a primitive group.
__device__ __forceinline__ bool group_has_node(const uint2 g) { return g.y > 0x00ffffffu; }

// Takes the nearest untested child out of `ngroup` (the caller pushes what is left), fetches its 80-byte node with
// five 128-bit ld.global.nc and tests the eight quantised child boxes against [0, best]; returns the children hit
// as a new node group and a primitive group.
__device__ __forceinline__ void wide_node_step(const uint4* __restrict__ nodes, const WideRay& r, uint2& ngroup, uint2& rest, uint2& tgroup,
                                               const uint32_t bias_bits = CUPT_QBIAS_BITS)
{
	const uint32_t hits = ngroup.y;
	const uint32_t bit = 31u - __clz(hits);
	rest = make_uint2(ngroup.x, hits & ~(1u << bit));
	const uint32_t slot = (bit - 24u) ^ r.oct_inv;
	const uint32_t rel = __popc(hits & 0xffu & ~(0xffffffffu << slot));
	const uint4* np = nodes + (size_t)(ngroup.x + rel) * 5u;
	const uint4 n0 = ldg_nc_u4(np + 0), n1 = ldg_nc_u4(np + 1), n2 = ldg_nc_u4(np + 2), n3 = ldg_nc_u4(np + 3), n4 = ldg_nc_u4(np + 4);
	const bool negx = r.ix < 0.0f, negy = r.iy < 0.0f, negz = r.iz < 0.0f;
	const float sx = __uint_as_float((n0.w & 0xffu) << 23) * r.ix;
	const float sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * r.iy;
	const float sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * r.iz;
	const float ox = (__uint_as_float(n0.x) - r.o.x) * r.ix;
	const float oy = (__uint_as_float(n0.y) - r.o.y) * r.iy;
	const float oz = (__uint_as_float(n0.z) - r.o.z) * r.iz;
	// Plane distances are t = (2^15 + q) * s + (o - 2^15 s).  Rounding slack PER AXIS (a common slack lets an axis
	// with a zero direction component, |o| ~ 1e30, open every box of the tree): 2^-21 relative to the magnitudes
	// involved plus 2^-6 of a grid cell, subtracted from the near planes and added to the far planes through the
	// two offsets, so it costs nothing per child.
	const float kx = fmaf(fabsf(sx), 0.015625f, fmaf(255.0f, fabsf(sx), fabsf(ox)) * 4.76837158203125e-7f);
	const float ky = fmaf(fabsf(sy), 0.015625f, fmaf(255.0f, fabsf(sy), fabsf(oy)) * 4.76837158203125e-7f);
	const float kz = fmaf(fabsf(sz), 0.015625f, fmaf(255.0f, fabsf(sz), fabsf(oz)) * 4.76837158203125e-7f);
	const float bx = fmaf(-CUPT_QBIAS, sx, ox), by = fmaf(-CUPT_QBIAS, sy, oy), bz = fmaf(-CUPT_QBIAS, sz, oz);
	const float onx = bx - kx, ony = by - ky, onz = bz - kz;
	const float ofx = bx + kx, ofy = by + ky, ofz = bz + kz;
	const uint32_t imask = n0.w >> 24;
	const uint32_t pvalid = n1.w;
	uint32_t hitmask = 0u;
#pragma unroll
	for (int h = 0; h < 2; ++h) {
		const uint32_t qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
		const uint32_t qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
		const uint32_t nx = negx ? qhx : qlx, fx = negx ? qlx : qhx;
		const uint32_t ny = negy ? qhy : qly, fy = negy ? qly : qhy;
		const uint32_t nz = negz ? qhz : qlz, fz = negz ? qlz : qhz;
#pragma unroll
		for (int j = 0; j < 4; ++j) {
			const float tnx = fmaf(u8f_biased(nx, j, bias_bits), sx, onx), tny = fmaf(u8f_biased(ny, j, bias_bits), sy, ony), tnz = fmaf(u8f_biased(nz, j, bias_bits), sz, onz);
			const float tfx = fmaf(u8f_biased(fx, j, bias_bits), sx, ofx), tfy = fmaf(u8f_biased(fy, j, bias_bits), sy, ofy), tfz = fmaf(u8f_biased(fz, j, bias_bits), sz, ofz);
			const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
			const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, r.best));
			if (tmin <= tmax) {
				// inner child: bit 24 + (slot ^ octant), i.e. front-to-back order from the top; leaf child: its unary record
				// count at the position of its first record (records are consecutive in slot order: layout.h)
				const uint32_t s = 4u * h + j;
				if ((imask >> s) & 1u) hitmask |= 1u << (24u + (s ^ r.oct_inv));
				else hitmask |= ((pvalid >> (4u * s)) & 7u) << __popc(pvalid & ((1u << (4u * s)) - 1u));
			}
		}
	}
	ngroup = make_uint2(n1.x, (hitmask & 0xff000000u) | imask);
	tgroup = make_uint2(n1.z, hitmask & 0x00ffffffu);
}

// The node step of the hot kernel (k_extend).  Same boxes, same slack and same result groups as wide_node_step, but
// the hit mask is built differently: wide_node_step spends an FSETP and five predicated ALU instructions per child on
// `hitmask |= cb << bi` (profiles/r1i_extend_opcodes.md: the ALU pipe — PRMT, FMNMX, LOP3, SHF, SEL — carries 45 % of
// the kernel's instructions and the FMA pipe 31 %).  Here
//   * a child is hit iff none of  tmax - tmin,  best - tmin,  tmax  is negative (tmin/tmax over the three axes only):
//     two FADDs on the FMA pipe, one LOP3 that ORs the three sign bits and one funnel shift that appends the sign to an
//     8-bit miss mask — 4 ALU + 2 FMA instructions per child instead of 10 ALU;
//   * the hit inner children are put in front-to-back order (slot XOR octant) by one look-up in a 2 KB shared-memory
//     table `perm` (8 octants x 256 masks) instead of per-child variable shifts;
//   * the hit LEAF children are not expanded at all: the primitive group is the node's valid-record word (one nibble
//     per slot) ANDed with the hit nibbles from a second 1 KB table `nib`, and it is `tgroup = {node index, sparse
//     mask}` instead of {first record, compact mask}; the record index of a bit is tri_base + popcount(valid bits
//     below it), computed when the pair is listed (sparse_record).  The first version of this function expanded the
//     hit leaves in a loop (count << first record per leaf) and gained nothing on diffuse bounces: the loop ran at
//     3-16 active lanes as often as the warp's worst lane had hit leaves (profiles/r1j_extend_lines.txt).
// No value here can be an infinity or NaN (see wide_ray_init), and without flush-to-zero a - b is negative exactly
// when a < b, so the decisions are those of `max(tmin, 0) <= min(tmax, best)`.
__device__ __forceinline__ void wide_node_step_fast(const uint4* __restrict__ nodes, const WideRay& r, uint2& ngroup, uint2& rest, uint2& tgroup,
                                                    const uint32_t bias_bits, const uint8_t* __restrict__ perm, const uint32_t* __restrict__ nib)
{
	const uint32_t hits = ngroup.y;
	const uint32_t bit = 31u - __clz(hits);
	rest = make_uint2(ngroup.x, hits & ~(1u << bit));
	const uint32_t slot = (bit - 24u) ^ r.oct_inv;
	const uint32_t rel = __popc(hits & 0xffu & ~(0xffffffffu << slot));
	const uint32_t node_index = ngroup.x + rel;
	const uint4* np = nodes + (size_t)node_index * 5u;
	const uint4 n0 = ldg_nc_u4(np + 0), n1 = ldg_nc_u4(np + 1), n2 = ldg_nc_u4(np + 2), n3 = ldg_nc_u4(np + 3), n4 = ldg_nc_u4(np + 4);
	const bool negx = r.ix < 0.0f, negy = r.iy < 0.0f, negz = r.iz < 0.0f;
	const float sx = __uint_as_float((n0.w & 0xffu) << 23) * r.ix;
	const float sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * r.iy;
	const float sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * r.iz;
	const float ox = (__uint_as_float(n0.x) - r.o.x) * r.ix;
	const float oy = (__uint_as_float(n0.y) - r.o.y) * r.iy;
	const float oz = (__uint_as_float(n0.z) - r.o.z) * r.iz;
	const float kx = fmaf(fabsf(sx), 0.015625f, fmaf(255.0f, fabsf(sx), fabsf(ox)) * 4.76837158203125e-7f);   // per-axis slack, see wide_node_step
	const float ky = fmaf(fabsf(sy), 0.015625f, fmaf(255.0f, fabsf(sy), fabsf(oy)) * 4.76837158203125e-7f);
	const float kz = fmaf(fabsf(sz), 0.015625f, fmaf(255.0f, fabsf(sz), fabsf(oz)) * 4.76837158203125e-7f);
	const float bx = fmaf(-CUPT_QBIAS, sx, ox), by = fmaf(-CUPT_QBIAS, sy, oy), bz = fmaf(-CUPT_QBIAS, sz, oz);
	const float onx = bx - kx, ony = by - ky, onz = bz - kz;
	const float ofx = bx + kx, ofy = by + ky, ofz = bz + kz;
	const uint32_t imask = n0.w >> 24;
	uint32_t miss = 0u;
#if !defined(CUPT_HOST_EMUL) && CUPT_NODE_F32X2
	// Packed evaluation (FFMA2 / FADD2): the near and the far plane of an axis share the scale, so one fma.rn.f32x2 per
	// child and axis yields (-tnear, tfar) = (qn, qf) * (-s, s) + (-on, of) — the near distance NEGATED, so that
	// -tmin = min3 of the negated near distances and both bounds come out of FMNMX3.min — and one add.rn.f32x2 per PAIR
	// of children each yields (tmax_a - tmin_a, tmax_b - tmin_b) and (best - tmin_a, best - tmin_b).  Negation commutes
	// with round-to-nearest, so every value is the scalar path's bit for bit (or its exact negative): 3 FFMA2 + 1 FADD2 per
	// child instead of 6 FFMA + 2 FADD.
	const f32x2 s2x = pack2(-sx, sx), s2y = pack2(-sy, sy), s2z = pack2(-sz, sz);
	const f32x2 o2x = pack2(-onx, ofx), o2y = pack2(-ony, ofy), o2z = pack2(-onz, ofz);
	const f32x2 best2 = pack2(r.best, r.best);
#pragma unroll
	for (int h = 1; h >= 0; --h) {   // slots 7..0, so that slot j ends up in bit j
		const uint32_t qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
		const uint32_t qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
		const uint32_t nx = negx ? qhx : qlx, fx = negx ? qlx : qhx;
		const uint32_t ny = negy ? qhy : qly, fy = negy ? qly : qhy;
		const uint32_t nz = negz ? qhz : qlz, fz = negz ? qlz : qhz;
#pragma unroll
		for (int j = 3; j >= 1; j -= 2) {   // children j and j - 1
			float ntmin[2], tmax[2];
#pragma unroll
			for (int c = 0; c < 2; ++c) {
				const int jj = j - c;
				const f32x2 tx = ffma2(pack2(u8f_biased(nx, jj, bias_bits), u8f_biased(fx, jj, bias_bits)), s2x, o2x);
				const f32x2 ty = ffma2(pack2(u8f_biased(ny, jj, bias_bits), u8f_biased(fy, jj, bias_bits)), s2y, o2y);
				const f32x2 tz = ffma2(pack2(u8f_biased(nz, jj, bias_bits), u8f_biased(fz, jj, bias_bits)), s2z, o2z);
				ntmin[c] = fminf(fminf(lo2(tx), lo2(ty)), lo2(tz));
				tmax[c] = fminf(fminf(hi2(tx), hi2(ty)), hi2(tz));
			}
			const f32x2 nt2 = pack2(ntmin[0], ntmin[1]);
			const f32x2 w2 = fadd2(pack2(tmax[0], tmax[1]), nt2);    // tmax - tmin
			const f32x2 b2 = fadd2(best2, nt2);                      // best - tmin
			miss = __funnelshift_l(__float_as_uint(lo2(w2)) | __float_as_uint(lo2(b2)) | __float_as_uint(tmax[0]), miss, 1);
			miss = __funnelshift_l(__float_as_uint(hi2(w2)) | __float_as_uint(hi2(b2)) | __float_as_uint(tmax[1]), miss, 1);
		}
	}
#else
#pragma unroll
	for (int h = 1; h >= 0; --h) {   // slots 7..0, so that slot j ends up in bit j
		const uint32_t qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
		const uint32_t qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
		const uint32_t nx = negx ? qhx : qlx, fx = negx ? qlx : qhx;
		const uint32_t ny = negy ? qhy : qly, fy = negy ? qly : qhy;
		const uint32_t nz = negz ? qhz : qlz, fz = negz ? qlz : qhz;
#pragma unroll
		for (int j = 3; j >= 0; --j) {
			const float tnx = fmaf(u8f_biased(nx, j, bias_bits), sx, onx), tny = fmaf(u8f_biased(ny, j, bias_bits), sy, ony), tnz = fmaf(u8f_biased(nz, j, bias_bits), sz, onz);
			const float tfx = fmaf(u8f_biased(fx, j, bias_bits), sx, ofx), tfy = fmaf(u8f_biased(fy, j, bias_bits), sy, ofy), tfz = fmaf(u8f_biased(fz, j, bias_bits), sz, ofz);
			const float tmin = fmaxf(fmaxf(tnx, tny), tnz);
			const float tmax = fminf(fminf(tfx, tfy), tfz);
			const uint32_t neg = __float_as_uint(tmax - tmin) | __float_as_uint(r.best - tmin) | __float_as_uint(tmax);
			miss = __funnelshift_l(neg, miss, 1);
		}
	}
#endif
	const uint32_t inner = ~miss & imask;
	const uint32_t leaf = ~miss & ~imask & 0xffu;   // (an empty slot has no valid bits)
	ngroup = make_uint2(n1.x, ((uint32_t)perm[inner] << 24) | imask);
	// the hit leaves' records as a SPARSE mask (one nibble per slot, layout.h): no per-child work at all; the record
	// index of a bit is worked out by sparse_record() when the pair is listed
	tgroup = make_uint2(node_index, n1.w & nib[leaf]);
}

// record index of bit b of a sparse primitive group; tv = {tri_base, pvalid} of its node (node_records)
__device__ __forceinline__ uint2 node_records(const uint4* __restrict__ nodes, const uint32_t node_index)
{
#ifdef CUPT_HOST_EMUL
	return *(reinterpret_cast<const uint2*>(nodes + (size_t)node_index * 5u + 1u) + 1);
#else
	return __ldg(reinterpret_cast<const uint2*>(nodes + (size_t)node_index * 5u + 1u) + 1);
#endif
}
__device__ __forceinline__ uint32_t sparse_record(const uint2 tv, const uint32_t b) { return tv.x + (uint32_t)__popc(tv.y & ((1u << b) - 1u)); }

// Tests the lowest-numbered pending record of `tgroup` (three 128-bit fetches) and updates the closest hit with the
// reference's acceptance rule `0 < t && t < closestT` (kernel.cu:127), equal t resolved by rank.
__device__ __forceinline__ void wide_prim_step(const float4* __restrict__ prims, WideRay& r, uint2& tgroup)
{
	const uint32_t k = __ffs(tgroup.y) - 1u;
	tgroup.y &= tgroup.y - 1u;
	const float4* pp = prims + (size_t)(tgroup.x + k) * 3u;
	const float4 r0 = ldg_nc_f4(pp + 0), r1 = ldg_nc_f4(pp + 1), r2 = ldg_nc_f4(pp + 2);
	const uint32_t kind = __float_as_uint(r2.w);
	float t;
	if (kind == 0u) t = tri_refexact(r.o, r.d, r0, r1, r2);
	else if (kind == 2u) t = tri_watertight(r.o, r.d, r0, r1, r2);
	else t = sphere_nearest(r.o, r.d, r0, r1);
	const uint32_t rank = __float_as_uint(r2.z);
	if (0.0f < t && (t < r.best || (t == r.best && rank < r.best_rank))) {
		r.best = t; r.best_rank = rank;
		r.best_id = (int32_t)(__float_as_uint(r2.y) | ((kind & 1u) << 31));
	}
}

// One ray, start to end (host emulation harness, and the definition the warp-synchronous kernel loop in kernels.cuh
// must agree with).  Stack entry i of this thread lives at stack[i * stride].
template <bool STATS>
__device__ __forceinline__ void traverse_wide(const uint4* __restrict__ nodes, const float4* __restrict__ prims,
                                              const f3 o, const f3 d, Hit& hit, uint2* stack, const int stride,
                                              const int stack_cap, uint32_t& n_nodes, uint32_t& n_tris)
{
	WideRay r;
	wide_ray_init(r, o, d);
	uint2 ngroup = make_uint2(0u, 0x80000000u);
	uint2 tgroup = make_uint2(0u, 0u);
	int sp = 0;
	for (;;) {
		if (group_has_node(ngroup)) {
			uint2 rest;
			wide_node_step(nodes, r, ngroup, rest, tgroup);
			if (group_has_node(rest)) {
				if (sp < stack_cap) stack[sp * stride] = rest;
				++sp;   // entries beyond the capacity are dropped only if the builder's depth bound is violated (checked on upload)
			}
			if (STATS) ++n_nodes;
		} else {
			tgroup = ngroup;
			ngroup = make_uint2(0u, 0u);
		}
		while (tgroup.y != 0u) {
			wide_prim_step(prims, r, tgroup);
			if (STATS) ++n_tris;
		}
		if (!group_has_node(ngroup)) {
			if (sp == 0) break;
			--sp;
			ngroup = stack[sp * stride];
		}
	}
	hit.t = r.best; hit.id = r.best_id; hit.rank = r.best_rank;
}

// One ray through wide_node_step_fast and the sparse primitive groups: the single-ray definition of what k_extend's
// warp-synchronous loop does per lane (host emulation harness; the kernel shares the two functions it calls).
// perm: 8 x 256 bytes, perm[oct << 8 | mask] has bit (s ^ oct) set for every slot s in mask; nib: 256 words,
// nib[mask] = the nibbles of the slots in mask (both built by k_extend in shared memory).
template <bool STATS>
__device__ __forceinline__ void traverse_wide_fast(const uint4* __restrict__ nodes, const float4* __restrict__ prims,
                                                   const f3 o, const f3 d, Hit& hit, uint2* stack, const int stride,
                                                   const int stack_cap, const uint8_t* __restrict__ perm, const uint32_t* __restrict__ nib,
                                                   uint32_t& n_nodes, uint32_t& n_tris)
{
	WideRay r;
	wide_ray_init(r, o, d);
	uint2 ngroup = make_uint2(0u, 0x80000000u);
	int sp = 0;
	for (;;) {
		uint2 rest, tgroup;
		wide_node_step_fast(nodes, r, ngroup, rest, tgroup, CUPT_QBIAS_BITS, perm + (r.oct_inv << 8), nib);
		if (group_has_node(rest)) {
			if (sp < stack_cap) stack[sp * stride] = rest;
			++sp;
		}
		if (STATS) ++n_nodes;
		if (tgroup.y != 0u) {
			const uint2 tv = node_records(nodes, tgroup.x);
			while (tgroup.y != 0u) {
				const uint32_t k = __ffs(tgroup.y) - 1u;
				tgroup.y &= tgroup.y - 1u;
				uint2 one = make_uint2(sparse_record(tv, k), 1u);
				wide_prim_step(prims, r, one);
				if (STATS) ++n_tris;
			}
		}
		if (!group_has_node(ngroup)) {
			if (sp == 0) break;
			--sp;
			ngroup = stack[sp * stride];
		}
	}
	hit.t = r.best; hit.id = r.best_id; hit.rank = r.best_rank;
}

// rayAABBIntersect, BVH.h:51-83, in the reference's arithmetic: infinite-line slab test, six IEEE divisions, no t range,
// the final z update skipped.  Not conservative (no rounding margin), which is why it can lose a triangle the ray hits.
__device__ __forceinline__ bool ref_slab_hit(const f3 o, const f3 d, const cupt_bvh2_node& nd)
{
	float tmin = rdiv(rsub(nd.lo.x, o.x), d.x), tmax = rdiv(rsub(nd.hi.x, o.x), d.x);
	if (tmin > tmax) { const float s = tmin; tmin = tmax; tmax = s; }
	float tymin = rdiv(rsub(nd.lo.y, o.y), d.y), tymax = rdiv(rsub(nd.hi.y, o.y), d.y);
	if (tymin > tymax) { const float s = tymin; tymin = tymax; tymax = s; }
	if ((tmin > tymax) || (tymin > tmax)) return false;
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = rdiv(rsub(nd.lo.z, o.z), d.z), tzmax = rdiv(rsub(nd.hi.z, o.z), d.z);
	if (tzmin > tzmax) { const float s = tzmin; tzmin = tzmax; tzmax = s; }
	if ((tmin > tzmax) || (tzmin > tmax)) return false;
	return true;
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
			const bool b = ref_slab_hit(o, d, nd);
			if (b) { stack[i] = nd.right; stack[i + 1] = nd.left; ++i; }
			else --i;
		}
	}
	hit.t = closestT; hit.id = trisID; hit.rank = 0u;
}

}  // namespace cupt
