// k_shade_regroup — the shade stage (one iteration of radianceAlongSingleStep2, kernel.cu:424-512) with the block's work
// re-grouped in shared memory so that the divergent parts run in full, homogeneous warps.
//
// k_shade (kernels.cuh) gives every queue entry to the thread that loaded it.  At depth >= 1 half of the paths are on their
// last segment (they only collect emission and stop), and the rest splits 50/50 between the cosine bounce (kernel.cu:462-464)
// and the light-point sample (kernel.cu:468-507): the Philox blocks ran at 14.7 of 32 lanes and either branch at 7.4
// (profiles/r1n_shade_lines.txt), the kernel spent 53 % of its issue slots at 13 active threads per instruction.  Here a block
// of 256 entries goes through four phases:
//   A  every thread: its entry's records, the hit's material, emission (exactly k_shade's code); a path that CONTINUES puts
//      what the next vertex needs (hit point, normal, direction, weight, albedo, counters) into a shared-memory item
//   B  the continuing items, compacted: thread t takes the t-th of them — full warps — and draws the vertex's four uniforms
//      (one or two Philox blocks), which also decides the item's class: mirror / cosine bounce / light sample
//   C  the items ordered by class: thread t takes the t-th in that order, so a warp runs ONE side of the 50/50 branch (two
//      warps of a block at most straddle a class boundary); then the common tail (normalise, weight, early-stop distance)
//   D  every thread again: the results of ITS entry back from shared memory, records written and the queue entry pushed in the
//      original thread order — the next queue is in slot order exactly as before, which k_extend's and the next k_shade's
//      coalescing rely on (a class-sorted QUEUE was measured in round 1 and lost: DESIGN.md section 5)
// The arithmetic per path is k_shade's, operation for operation, so the image is bit-identical.
#pragma once
#include "kernels.cuh"

namespace cupt {

enum { SG_POSX = 0, SG_POSY, SG_POSZ, SG_NX, SG_NY, SG_NZ, SG_DX, SG_DY, SG_DZ, SG_WR, SG_WG, SG_WB, SG_AR, SG_AG, SG_AB,
       SG_PACKED, SG_SLOT, SG_UA, SG_UB, SG_UC, SG_STOP, SG_FIELDS };
// SG_PACKED in: iter | draw << 8 | mirror << 16 ; out: iter | draw << 8 | push << 31

__global__ void __launch_bounds__(256, CUPT_SHADE_MIN_BLOCKS) k_shade_regroup(DevScene sc, PassParams pp, PathBuffers pb, int depth, int qsel)
{
	__shared__ uint32_t s_warp[9];
	__shared__ float s_item[SG_FIELDS][256];
	__shared__ uint16_t s_cont[256];      // phase B: the continuing items, in thread order
	__shared__ uint16_t s_order[256];     // phase C: the same items ordered by class
	__shared__ uint32_t s_wcnt[3][8];     // per warp and class: items of phase B's warp w in class c
	__shared__ uint32_t s_n[4];           // [0] continuing items, [1..3] class starts
	const uint32_t count = pb.counters[depth];
	const uint32_t* __restrict__ qin = pb.queue[qsel];
	uint32_t* __restrict__ qout = pb.queue[qsel ^ 1];
	const float kBrdf = (float)(1 / 3.14159);   // BRDF(): mul(albedo, 1 / 3.14159) with a float parameter (kernel.cu:101-104)
	const float kPi = (float)3.14159;            // color::operator*(const float&) (kernel.cu:463)
	const uint32_t tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;
	uint32_t slot_next = 0;
	{
		const uint32_t q0 = blockIdx.x * blockDim.x + tid;
		if (q0 < count) slot_next = qin[q0];
	}
	for (uint32_t base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
		const uint32_t qi = base + tid;
		const uint32_t slot = slot_next;
		{
			const uint32_t qn = qi + gridDim.x * blockDim.x;
			if (qn < count) slot_next = qin[qn];
		}
		// ---- A: the hit (k_shade's code up to the decision to continue) -------------------------------------------------
		bool cont = false;
		if (qi < count) {
			const float4 ra = pb.rayA[slot], rb = pb.rayB[slot], s0 = pb.st0[slot];
			const f3 o = mk3(ra.x, ra.y, ra.z), d = mk3(ra.w, rb.x, rb.y);
			float wr = s0.x, wg = s0.y, wb = s0.z;
			const uint32_t packed = __float_as_uint(s0.w);
			const int iter = (int)(packed & 0xffu);
			int32_t id = __float_as_int(rb.w);
			float closestT = (float)((double)rb.z - 0.001);                        // kernel.cu:431
			if ((double)closestT < 0.001) { wr = wg = wb = 0.f; }                   // kernel.cu:432-435
			if (closestT > CUPT_MAX_FLOAT - 1) { wr = wg = wb = 0.f; id = 0; closestT = 0.f; }  // kernel.cu:436-441
			const f3 pos = radd3(o, rmul3(d, closestT));                            // kernel.cu:449
			f3 normal;
			float4 m0, m1;
			if (id >= 0) {
				const float4 tsh = sc.trishade[id];
				normal = mk3(tsh.x, tsh.y, tsh.z);
				const int mat = __float_as_int(tsh.w);
				m0 = sc.mats[2 * mat]; m1 = sc.mats[2 * mat + 1];
			} else {
				const uint32_t si = (uint32_t)id & 0x7fffffffu;
				const float4 sp0 = sc.spheres[3 * si];
				const f3 hp = radd3(o, rmul3(d, rb.z));
				normal = rdiv3(rsub3(hp, mk3(sp0.x, sp0.y, sp0.z)), sp0.w);
				m0 = sc.spheres[3 * si + 1]; m1 = sc.spheres[3 * si + 2];
			}
			if (m0.w != 0.f) {                                                      // kernel.cu:453-457
				float4 s1 = pb.st1[slot];
				s1.x += wr * m1.x; s1.y += wg * m1.y; s1.z += wb * m1.z;
				pb.st1[slot] = s1;
				wr = wg = wb = 0.f;
			}
			const bool dead = !(wr != 0.f || wg != 0.f || wb != 0.f);
			cont = !(dead || iter >= pp.max_depth - 1);
			if (cont) {
				s_item[SG_POSX][tid] = pos.x; s_item[SG_POSY][tid] = pos.y; s_item[SG_POSZ][tid] = pos.z;
				s_item[SG_NX][tid] = normal.x; s_item[SG_NY][tid] = normal.y; s_item[SG_NZ][tid] = normal.z;
				s_item[SG_DX][tid] = d.x; s_item[SG_DY][tid] = d.y; s_item[SG_DZ][tid] = d.z;
				s_item[SG_WR][tid] = wr; s_item[SG_WG][tid] = wg; s_item[SG_WB][tid] = wb;
				s_item[SG_AR][tid] = m0.x; s_item[SG_AG][tid] = m0.y; s_item[SG_AB][tid] = m0.z;
				s_item[SG_PACKED][tid] = __uint_as_float((packed & 0xffffu) | ((m1.w != 0.f && id >= 0) ? 0x10000u : 0u));
				s_item[SG_SLOT][tid] = __uint_as_float(slot);
			}
		}
		// compaction of the continuing items, thread order kept
		{
			const uint32_t m = __ballot_sync(0xffffffffu, cont);
			if (lane == 0u) s_warp[w] = (uint32_t)__popc(m);
			__syncthreads();
			if (tid == 0u) {
				uint32_t run = 0u;
				for (uint32_t i = 0; i < 8u; ++i) { const uint32_t c = s_warp[i]; s_warp[i] = run; run += c; }
				s_n[0] = run;
			}
			__syncthreads();
			if (cont) s_cont[s_warp[w] + (uint32_t)__popc(m & ((1u << lane) - 1u))] = (uint16_t)tid;
			__syncthreads();
		}
		const uint32_t ncont = s_n[0];
		// ---- B: the vertex's uniforms, in full warps; class of the item -----------------------------------------------------
		uint32_t cls = 3u, item = 0u;
		if (tid < ncont) {
			item = s_cont[tid];
			const uint32_t packed = __float_as_uint(s_item[SG_PACKED][item]);
			const uint32_t draw = (packed >> 8) & 0xffu;
			int x, y, sl;
			slot_to_pixel(pp, __float_as_uint(s_item[SG_SLOT][item]), x, y, sl);
			PhiloxStream rng;
			rng.init(pp.seed, (uint32_t)x, (uint32_t)y, (uint32_t)(pp.first_sample + sl));
			const float ud = rng.draw(draw), ua = rng.draw(draw + 1u), ub = rng.draw(draw + 2u), uc = rng.draw(draw + 3u);
			s_item[SG_UA][item] = ua; s_item[SG_UB][item] = ub; s_item[SG_UC][item] = uc;
			cls = (packed & 0x10000u) ? 0u : (ud < 0.5f ? 1u : 2u);                 // mirror / kernel.cu:459
		}
		// order by class: per-warp counts, a 3 x 8 prefix, the item's place
		{
			const uint32_t m0b = __ballot_sync(0xffffffffu, cls == 0u), m1b = __ballot_sync(0xffffffffu, cls == 1u), m2b = __ballot_sync(0xffffffffu, cls == 2u);
			if (lane == 0u) { s_wcnt[0][w] = (uint32_t)__popc(m0b); s_wcnt[1][w] = (uint32_t)__popc(m1b); s_wcnt[2][w] = (uint32_t)__popc(m2b); }
			__syncthreads();
			if (tid == 0u) {
				uint32_t run = 0u;
				for (uint32_t c = 0; c < 3u; ++c) {
					s_n[1 + c] = run;
					for (uint32_t i = 0; i < 8u; ++i) { const uint32_t v = s_wcnt[c][i]; s_wcnt[c][i] = run; run += v; }
				}
			}
			__syncthreads();
			if (cls < 3u) {
				const uint32_t mm = cls == 0u ? m0b : (cls == 1u ? m1b : m2b);
				s_order[s_wcnt[cls][w] + (uint32_t)__popc(mm & ((1u << lane) - 1u))] = (uint16_t)item;
			}
			__syncthreads();
		}
		// ---- C: the next vertex, one class per warp ------------------------------------------------------------------------------
		if (tid < ncont) {
			const uint32_t it = s_order[tid];
			const uint32_t kind = tid < s_n[2] ? 0u : (tid < s_n[3] ? 1u : 2u);
			const f3 pos = mk3(s_item[SG_POSX][it], s_item[SG_POSY][it], s_item[SG_POSZ][it]);
			const f3 normal = mk3(s_item[SG_NX][it], s_item[SG_NY][it], s_item[SG_NZ][it]);
			const uint32_t packed = __float_as_uint(s_item[SG_PACKED][it]);
			int iter = (int)(packed & 0xffu);
			uint32_t draw = (packed >> 8) & 0xffu;
			float wr = s_item[SG_WR][it], wg = s_item[SG_WG][it], wb = s_item[SG_WB][it];
			const float ar = s_item[SG_AR][it], ag = s_item[SG_AG][it], ab = s_item[SG_AB][it];
			f3 vdir;
			float k = 1.0f;
			if (kind == 0u) {
				// perfect mirror (extension, config 5; no reference code): d - 2 (d.n) n, weight *= albedo, no random numbers
				const f3 d = mk3(s_item[SG_DX][it], s_item[SG_DY][it], s_item[SG_DZ][it]);
				vdir = rsub3(d, rmul3(normal, rmul(2.0f, rdot(d, normal))));
			} else if (kind == 1u) {
				// cosineWeightedRay, kernel.cu:78-99
				draw += 3u;
				const float u1 = s_item[SG_UA][it], u2 = s_item[SG_UB][it];
				const float r = __fsqrt_rn(u1);
				const float theta = (float)(2 * 3.14159 * (double)u2);
				const float cx = rmul(r, cosf(theta)), cz = rmul(r, sinf(theta));
				const float cy = __fsqrt_rn(fmaxf(0.0f, rsub(1.0f, u1)));
				const f3 tangent = get_tangent(normal), bitangent = rcross(normal, tangent);
				vdir = radd3(radd3(rmul3(normal, cy), rmul3(tangent, cx)), rmul3(bitangent, cz));
				k = kBrdf * kPi;                                                 // kernel.cu:463-464
			} else {
				// light-point sample, kernel.cu:468-507
				draw += 4u;
				float randArea = rmul(sc.totalLightArea, s_item[SG_UA][it]);
				int sel = sc.numLights;   // fallback entry = triangle 0 (`selectedTri = 0`, kernel.cu:469)
				if (sc.numLights <= CUPT_LIGHT_SCAN_MAX) {
					for (int j = 0; j < sc.numLights; ++j) {   // the reference's linear scan, kernel.cu:470-481, op for op
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
				float u = s_item[SG_UB][it], v = s_item[SG_UC][it];
				const float4 l0 = sc.lights[3 * sel], l1 = sc.lights[3 * sel + 1], l2 = sc.lights[3 * sel + 2];
				if (radd(u, v) > 1.0f) {                                          // kernel.cu:491-495
					u = (float)((double)u + 2 * (0.5 - (double)u));
					v = (float)((double)v + 2 * (0.5 - (double)v));
				}
				const f3 pos1 = radd3(radd3(mk3(l0.x, l0.y, l0.z), rmul3(mk3(l1.x, l1.y, l1.z), u)), rmul3(mk3(l2.x, l2.y, l2.z), v));
				vdir = rsub3(pos1, pos);
			}
			const f3 lDir = rnormalized(vdir);
			if (kind == 2u) {
				const float cosL = fmaxf(0.0f, rdot(lDir, normal));
				const float cosO1 = fmaxf(0.0f, rdot(mk3(0.f, -1.f, 0.f), rmul3(lDir, -1.f)));   // hard-coded light normal, kernel.cu:503
				const float G = rdiv(rmul(cosL, cosO1), rdot(vdir, vdir));
				k = kBrdf * G * sc.totalLightArea;                               // kernel.cu:506
				iter = max(iter, pp.max_depth - 2);                              // kernel.cu:507
			}
			if (kind == 0u) { wr *= ar; wg *= ag; wb *= ab; }
			else { wr *= ar * k; wg *= ag * k; wb *= ab * k; }
			++iter;
			const bool push = (wr != 0.f || wg != 0.f || wb != 0.f);
			float stop = __int_as_float(-1);   // NaN bits = no early stop (`best < NaN` is never true)
			if (push && pp.term_stop && iter >= pp.max_depth - 1) stop = emitter_entry_distance(pp, pos, lDir);
			s_item[SG_DX][it] = lDir.x; s_item[SG_DY][it] = lDir.y; s_item[SG_DZ][it] = lDir.z;
			s_item[SG_WR][it] = wr; s_item[SG_WG][it] = wg; s_item[SG_WB][it] = wb;
			s_item[SG_STOP][it] = stop;
			s_item[SG_PACKED][it] = __uint_as_float((uint32_t)iter | (draw << 8) | (push ? 0x80000000u : 0u));
		}
		__syncthreads();
		// ---- D: records and queue entry, by the thread that owns the entry (slot order) ------------------------------------------
		bool push = false;
		if (cont) {
			const uint32_t packed = __float_as_uint(s_item[SG_PACKED][tid]);
			push = (packed >> 31) != 0u;
			if (push) {
				pb.rayA[slot] = make_float4(s_item[SG_POSX][tid], s_item[SG_POSY][tid], s_item[SG_POSZ][tid], s_item[SG_DX][tid]);
				pb.rayB[slot] = make_float4(s_item[SG_DY][tid], s_item[SG_DZ][tid], CUPT_MAX_FLOAT, s_item[SG_STOP][tid]);
				pb.st0[slot] = make_float4(s_item[SG_WR][tid], s_item[SG_WG][tid], s_item[SG_WB][tid], __uint_as_float(packed & 0xffffu));
			}
		}
		const uint32_t pos = block_queue_reserve(push, &pb.counters[depth + 1], s_warp);
		if (push) qout[pos] = slot;
	}
}

}  // namespace cupt
