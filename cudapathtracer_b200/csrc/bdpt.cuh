// Wavefront form of the reference's LIVE integrator radianceAlongSingleStep (kernel.cu:217-415): a fixed-length
// bidirectional path tracer with two light vertices and three eye vertices, four unweighted connections (no MIS), the
// primary hit's emission added four times (kernel.cu:409-410), seven uniforms per sample.  The megakernel's 3 + <= 4
// trace() calls become three rounds of the shared extend kernel:
//
//   k_bdpt_gen      light vertex 0 (area-sampled light point, kernel.cu:230-267), its uniform-hemisphere ray
//                   (randRay, kernel.cu:60-77) and the camera ray                         -> 2 rays per path
//   extend
//   k_bdpt_mid      eye vertex 3 = primary hit, cosine-weighted bounce (kernel.cu:306-333) -> 1 ray per path
//   extend
//   k_bdpt_connect  light vertex 1, eye vertex 2, the 2 x 2 connection weights in f64 like `color`
//                   (kernel.cu:353-395), a visibility ray for every weight above 0.01      -> <= 4 rays per path
//   extend
//   k_bdpt_accumulate  visible iff |t - len| <= 0.01 (kernel.cu:396-405), per-pixel sum
//
// Ray records live at index 4 * slot + k of rayA / rayB; st0[4 * slot + c] holds connection c's weight and length,
// st1[slot] four times the primary hit's emission (xyz) and the sampled light triangle (w).  No vertex array is kept
// between the stages: every stage recomputes the vertices it needs from the ray records and their hits with the
// reference's operation order, so the geometry is bit-identical to the megakernel's.
#pragma once
#include "kernels.cuh"

namespace cupt {

// kernel.cu:60-77
__device__ __forceinline__ f3 rand_ray_dir(const f3 n, const float u1, const float u2)
{
	const float r = __fsqrt_rn(rsub(1.0f, rmul(u1, u1)));
	const float phi = (float)(2 * 3.14159 * (double)u2);
	const float cx = rmul(r, cosf(phi)), cz = rmul(r, sinf(phi));
	const f3 tangent = get_tangent(n), bitangent = rcross(n, tangent);
	return rnormalized(radd3(radd3(rmul3(n, u1), rmul3(tangent, cx)), rmul3(bitangent, cz)));
}
// kernel.cu:78-99
__device__ __forceinline__ f3 cosine_dir(const f3 n, const float u1, const float u2)
{
	const float r = __fsqrt_rn(u1);
	const float theta = (float)(2 * 3.14159 * (double)u2);
	const float cx = rmul(r, cosf(theta)), cz = rmul(r, sinf(theta));
	const float cy = __fsqrt_rn(fmaxf(0.0f, rsub(1.0f, u1)));
	const f3 tangent = get_tangent(n), bitangent = rcross(n, tangent);
	return rnormalized(radd3(radd3(rmul3(n, cy), rmul3(tangent, cx)), rmul3(bitangent, cz)));
}

// scene.tris[id]: normal and material; id = -1 is the escaped bounce's read of tris[-1] (kernel.cu:336,345-346),
// defined here as the zeroed guard triangle of the reference build (normal 0, material 0)
__device__ __forceinline__ void tri_shade_at(const DevScene& sc, const int32_t id, f3& n, int& mat)
{
	if (id < 0) { n = mk3(0.f, 0.f, 0.f); mat = 0; return; }
	const float4 t = sc.trishade[id];
	n = mk3(t.x, t.y, t.z); mat = __float_as_int(t.w);
}

// the hit of a traced segment the way kernel.cu:276-283,311-318 post-processes it: t -= 0.001, miss -> triangle 0 at t = 0
__device__ __forceinline__ void hit_with_miss_rule(const float4 rb, int32_t& id, float& t)
{
	id = __float_as_int(rb.w);
	t = (float)((double)rb.z - 0.001);
	if (t > CUPT_MAX_FLOAT - 1) { id = 0; t = 0.f; }
}

__device__ __forceinline__ int pick_light(const DevScene& sc, float randArea)
{
	int sel = sc.numLights;   // fallback entry = triangle 0 (`selectedTri = 0`, kernel.cu:232)
	if (sc.numLights <= CUPT_LIGHT_SCAN_MAX) {
		for (int j = 0; j < sc.numLights; ++j) {   // kernel.cu:233-244
			const float area = sc.lights[3 * j].w;
			if (randArea < area && randArea > 0.f) sel = j;
			randArea = rsub(randArea, area);
		}
	} else {
		int lo = 0, hi = sc.numLights;
		while (lo < hi) {
			const int mid = (lo + hi) >> 1;
			if (randArea < sc.light_cdf[mid]) hi = mid; else lo = mid + 1;
		}
		sel = lo;
	}
	return sel;
}

__global__ void __launch_bounds__(256) k_bdpt_gen(DevScene sc, PassParams pp, PathBuffers pb)
{
	__shared__ uint32_t s_warp[9];
	const uint32_t total = (uint32_t)pp.p_pad * (uint32_t)pp.samples;
	for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += gridDim.x * blockDim.x) {
		const uint32_t slot = base + threadIdx.x;
		bool valid = false;
		if (slot < total) {
			int x, y, sl;
			slot_to_pixel(pp, slot, x, y, sl);
			valid = x < pp.cam.pxlWidth && y < pp.cam.pxlHeight;
			if (valid) {
				PhiloxStream rng;
				rng.init(pp.seed, (uint32_t)x, (uint32_t)y, (uint32_t)(pp.first_sample + sl));
				const float l1 = rng.draw(0), l2 = rng.draw(1);                     // kernel.cu:547
				f3 co, cd;
				camera_ray(pp.cam, x, y, l1, l2, co, cd);
				// light vertex 0, kernel.cu:230-267
				const int sel = pick_light(sc, rmul(sc.totalLightArea, rng.draw(2)));
				float u = rng.draw(3), v = rng.draw(4);
				const float4 l0 = sc.lights[3 * sel], la = sc.lights[3 * sel + 1], lb = sc.lights[3 * sel + 2];
				if ((double)radd(u, v) > 1.0) {
					u = (float)((double)u + 2 * (0.5 - (double)u));
					v = (float)((double)v + 2 * (0.5 - (double)v));
				}
				const int32_t selTri = __float_as_int(la.w);
				f3 n0; int m0;
				tri_shade_at(sc, selTri, n0, m0);
				const f3 x0 = radd3(radd3(radd3(mk3(l0.x, l0.y, l0.z), rmul3(mk3(la.x, la.y, la.z), u)), rmul3(mk3(lb.x, lb.y, lb.z), v)), rmul3(n0, 0.001f));
				const f3 od = rand_ray_dir(n0, rng.draw(5), rng.draw(6));
				const size_t r = (size_t)slot * 4u;
				pb.rayA[r + 0] = make_float4(x0.x, x0.y, x0.z, od.x);
				pb.rayB[r + 0] = make_float4(od.y, od.z, CUPT_MAX_FLOAT, __int_as_float(-1));
				pb.rayA[r + 1] = make_float4(co.x, co.y, co.z, cd.x);
				pb.rayB[r + 1] = make_float4(cd.y, cd.z, CUPT_MAX_FLOAT, __int_as_float(-1));
				pb.st1[slot] = make_float4(0.f, 0.f, 0.f, __int_as_float(selTri));
			}
		}
		const uint32_t p0 = block_queue_reserve(valid, &pb.counters[0], s_warp);
		if (valid) pb.queue[0][p0] = slot * 4u + 0u;
		const uint32_t p1 = block_queue_reserve(valid, &pb.counters[0], s_warp);
		if (valid) pb.queue[0][p1] = slot * 4u + 1u;
	}
	if (blockIdx.x == 0 && threadIdx.x == 0)
		atomicAdd(&pb.stats[3], (unsigned long long)pp.cam.pxlWidth * pp.cam.pxlHeight * pp.samples);
}

__global__ void __launch_bounds__(256) k_bdpt_mid(DevScene sc, PassParams pp, PathBuffers pb)
{
	__shared__ uint32_t s_warp[9];
	const uint32_t total = (uint32_t)pp.p_pad * (uint32_t)pp.samples;
	for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += gridDim.x * blockDim.x) {
		const uint32_t slot = base + threadIdx.x;
		bool valid = false;
		if (slot < total) {
			int x, y, sl;
			slot_to_pixel(pp, slot, x, y, sl);
			valid = x < pp.cam.pxlWidth && y < pp.cam.pxlHeight;
			if (valid) {
				const size_t r = (size_t)slot * 4u;
				const float4 ca = pb.rayA[r + 1], cb = pb.rayB[r + 1];
				int32_t id; float t;
				hit_with_miss_rule(cb, id, t);                                      // kernel.cu:306-318
				f3 n3; int m3;
				tri_shade_at(sc, id, n3, m3);
				const f3 x3 = radd3(mk3(ca.x, ca.y, ca.z), rmul3(mk3(ca.w, cb.x, cb.y), t));
				PhiloxStream rng;
				rng.init(pp.seed, (uint32_t)x, (uint32_t)y, (uint32_t)(pp.first_sample + sl));
				const f3 bd = cosine_dir(n3, rng.draw(7), rng.draw(8));             // kernel.cu:332
				pb.rayA[r + 2] = make_float4(x3.x, x3.y, x3.z, bd.x);
				pb.rayB[r + 2] = make_float4(bd.y, bd.z, CUPT_MAX_FLOAT, __int_as_float(-1));
			}
		}
		const uint32_t p = block_queue_reserve(valid, &pb.counters[1], s_warp);
		if (valid) pb.queue[1][p] = slot * 4u + 2u;
	}
}

struct dcol { double r, g, b; };   // `color` (color.h:4-6): the reference keeps radiance in f64

#ifndef CUPT_BDPT_CONNECT_BLOCKS
#define CUPT_BDPT_CONNECT_BLOCKS 4   // 64 registers + 76 spilled bytes instead of 112 / none: mid + connect 11.3 -> 9.3 ms per frame (1 / 3 / 4 / 5 / 6 / 8 blocks: 11.3 / 10.4 / 9.3 / 11.7 / 14.3 / 14.0)
#endif
__global__ void __launch_bounds__(256, CUPT_BDPT_CONNECT_BLOCKS) k_bdpt_connect(DevScene sc, PassParams pp, PathBuffers pb, const double* __restrict__ mat_fs /* albedo / 3.14159f per material, f64 (cupt_set_scene) */)
{
	__shared__ uint32_t s_warp[33];
	const uint32_t total = (uint32_t)pp.p_pad * (uint32_t)pp.samples;
	for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += gridDim.x * blockDim.x) {
		const uint32_t slot = base + threadIdx.x;
		bool valid = false;
		bool want[4] = {false, false, false, false};
		if (slot < total) {
			int px, py, sl;
			slot_to_pixel(pp, slot, px, py, sl);
			valid = px < pp.cam.pxlWidth && py < pp.cam.pxlHeight;
		}
		if (valid) {
			const size_t r = (size_t)slot * 4u;
			f3 x[5], norm[4];
			int mat[4];
			float invProb[4];
			// light vertex 0 (k_bdpt_gen) and 1 (kernel.cu:268-298)
			const float4 la = pb.rayA[r + 0], lb = pb.rayB[r + 0];
			const f3 oDir = mk3(la.w, lb.x, lb.y);
			x[0] = mk3(la.x, la.y, la.z);
			tri_shade_at(sc, __float_as_int(pb.st1[slot].w), norm[0], mat[0]);
			invProb[0] = sc.totalLightArea;
			{
				int32_t id; float t;
				hit_with_miss_rule(lb, id, t);
				tri_shade_at(sc, id, norm[1], mat[1]);
				x[1] = radd3(x[0], rmul3(oDir, t));
				const float G = rdiv(fabsf(rdot(norm[1], oDir)), fmaxf(0.001f, rmul(t, t)));
				invProb[1] = (float)(2 * 3.14159 / (double)G);
			}
			// eye vertex 4 (camera), 3 (primary hit, kernel.cu:306-328) and 2 (bounce, kernel.cu:329-348)
			const float4 ca = pb.rayA[r + 1], ba = pb.rayA[r + 2], bb = pb.rayB[r + 2];
			x[4] = mk3(ca.x, ca.y, ca.z);
			x[3] = mk3(ba.x, ba.y, ba.z);
			{
				const float4 cb = pb.rayB[r + 1];
				int32_t id; float t;
				hit_with_miss_rule(cb, id, t);
				tri_shade_at(sc, id, norm[3], mat[3]);
				invProb[3] = 1.0f;
			}
			{
				const f3 dir = mk3(ba.w, bb.x, bb.y);
				const float t = (float)((double)bb.z - 0.001);                      // no miss check (kernel.cu:334)
				tri_shade_at(sc, __float_as_int(bb.w), norm[2], mat[2]);
				float G = rdiv(fabsf(rmul(rdot(norm[3], dir), rdot(norm[2], dir))), rmul(t, t));
				if (G == 0.0f) G = 1.0f;
				if (G != G) G = 1.0f;
				x[2] = radd3(x[3], rmul3(dir, t));
				invProb[2] = (float)(3.14159 / (double)G);
			}
			const float4 me0 = sc.mats[2 * mat[0] + 1];
			const dcol L_e = {(double)me0.x, (double)me0.y, (double)me0.z};
#pragma unroll
			for (int i = 0; i < 2; ++i)
#pragma unroll
				for (int j = 2; j < 4; ++j) {                                       // kernel.cu:353-412
					dcol w = {L_e.r * (double)invProb[0], L_e.g * (double)invProb[0], L_e.b * (double)invProb[0]};
#pragma unroll
					for (int k = 1; k <= 3; ++k) {
						if (!((k <= i) || (k >= j + 1))) continue;                  // light sub-path k = 1..i, eye sub-path k = j+1..3
						const f3 seg = rsub3(x[k], x[k - 1]);
						const f3 ray = rnormalized(seg);
						float G = rdiv(fabsf(rmul(rdot(ray, norm[k]), rdot(ray, norm[k - 1]))), rdot(seg, seg));
						if (G != G) G = 0.0f;
						const double* fk = mat_fs + 3 * mat[k];   // f_s = albedo / 3.14159f (kernel.cu:367,379)
						w.r = ((w.r * fk[0]) * (double)G) * (double)invProb[k];
						w.g = ((w.g * fk[1]) * (double)G) * (double)invProb[k];
						w.b = ((w.b * fk[2]) * (double)G) * (double)invProb[k];
					}
					const f3 seg = rsub3(x[j], x[i]);
					const float len = rlength(seg);
					const f3 ray = rnormalized(seg);
					float G = rdiv(fmaxf(0.0f, rmul(rdot(ray, norm[j]), rdot(rmul3(ray, -1.0f), norm[i]))), rdot(seg, seg));
					if (G != G) G = 0.0f;
					const double* fj = mat_fs + 3 * mat[j];   // kernel.cu:391
					w.r = ((w.r * fj[0]) * (double)G) * (double)invProb[j];
					w.g = ((w.g * fj[1]) * (double)G) * (double)invProb[j];
					w.b = ((w.b * fj[2]) * (double)G) * (double)invProb[j];
					const float m = (float)fmax(w.r, fmax(w.g, w.b));
					const int c = i * 2 + (j - 2);
					want[c] = (double)m > 0.01;
					// the visibility ray overwrites this path's ray record c: every record was read above
					if (want[c]) {
						pb.st0[r + c] = make_float4((float)w.r, (float)w.g, (float)w.b, len);
					} else {
						pb.st0[r + c] = make_float4(0.f, 0.f, 0.f, -1.f);
					}
					if (want[c]) {
						// (written after the loop would need the values kept; x[i] and ray are still live here, but record c
						// may be one of the inputs of a LATER (i, j): all inputs were copied into x/norm/mat above)
						// The visibility test only asks whether the CLOSEST hit lies within 0.01 of `len` (kernel.cu:396-405),
						// so the wide traversal may stop looking beyond len + 0.02: if nothing is hit up to there the extend
						// reports t = that limit, which fails |t - len| <= 0.01 exactly like the farther hit (or the miss)
						// the reference would have found; and a hit nearer than len - 0.02 already decides "occluded" whatever
						// else lies in front of it, so the traversal stops there (last component).  The reference-exact mode
						// ignores both.
						pb.rayA[r + c] = make_float4(x[i].x, x[i].y, x[i].z, ray.x);
						pb.rayB[r + c] = make_float4(ray.y, ray.z, fminf(__fadd_ru(len, 0.02f), CUPT_MAX_FLOAT), __fadd_rd(len, -0.02f));
					}
				}
			const float4 me3 = sc.mats[2 * mat[3] + 1];   // accum += emission of the primary hit, four times (kernel.cu:409-410)
			pb.st1[slot] = make_float4(4.0f * me3.x, 4.0f * me3.y, 4.0f * me3.z, 0.f);
		}
		uint32_t qpos[4];
		block_queue_reserve4(want, &pb.counters[2], s_warp, qpos);
#pragma unroll
		for (int c = 0; c < 4; ++c)
			if (want[c]) pb.queue[0][qpos[c]] = slot * 4u + (uint32_t)c;
	}
}

// The last stage: visibility of the connections (visible iff |t - len| <= 0.01, kernel.cu:396-405: an occluded connection
// contributes nothing), then the per-pixel sum.  Until r2bc a kernel of its own (k_bdpt_resolve) walked the queue of visibility
// rays and cleared the weights of the occluded ones — queue entry, hit and weight record read and the weight written back, about
// 68 bytes per visibility ray at 70 % of the DRAM peak, 0.96 ms per pass — before this one read the weights again.  Now the hit
// distances are read here, next to the weights (64 more bytes per sample, contiguous), and the test is made on the way to the sum.
__global__ void __launch_bounds__(256) k_bdpt_accumulate(PassParams pp, PathBuffers pb, float* __restrict__ accum)
{
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= (uint32_t)pp.p_pad) return;
	int x, y, sl;
	slot_to_pixel(pp, r, x, y, sl);
	if (x >= pp.cam.pxlWidth || y >= pp.cam.pxlHeight) return;
	float sr = 0.f, sg = 0.f, sb = 0.f;
	for (int s = 0; s < pp.samples; ++s) {
		const size_t slot = (size_t)s * pp.p_pad + r;
		float cr = 0.f, cg = 0.f, cb = 0.f;
		for (int c = 0; c < 4; ++c) {                   // accum = ((w0 + e) + (w1 + e)) ... in the reference's order
			float4 w = pb.st0[slot * 4u + c];           // (w.w = connection length; -1: no visibility ray was wanted, weight 0)
			if (w.w >= 0.f) {
				const float t = pb.rayB[slot * 4u + c].z;
				if (!((double)fabsf(rsub(t, w.w)) <= 0.01)) { w.x = 0.f; w.y = 0.f; w.z = 0.f; }
			}
			cr += w.x; cg += w.y; cb += w.z;
		}
		const float4 e4 = pb.st1[slot];
		sr += cr + e4.x; sg += cg + e4.y; sb += cb + e4.z;
	}
	float* a = accum + 3 * ((size_t)y * pp.cam.pxlWidth + x);
	a[0] += sr; a[1] += sg; a[2] += sb;
}

}  // namespace cupt
