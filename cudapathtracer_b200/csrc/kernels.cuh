// Wavefront pipeline: raygen -> [extend -> shade]* -> accumulate, replacing the megakernel drawPixel
// (reference kernel.cu:535-553) and the integrator radianceAlongSingleStep2 (kernel.cu:417-515).
//
// A pass renders S samples of every pixel: path slot i = s_local * P_pad + tile * 32 + lane, tiles of 8x4 pixels
// so that one warp of primary rays is a compact screen tile (the reference gets the same effect from Morton
// indexing, camera.h:57-65, which only covers square power-of-two images).  Pixel, sample and RNG stream are
// functions of the slot index, so no per-path RNG state exists (the reference carries 48 bytes of XORWOW state
// per pixel, kernel.cu:527-533).
//
// Per-path records (all 16-byte aligned, indexed by slot):
//   rayA  float4  o.xyz, d.x
//   rayB  float4  d.y, d.z, t (extend output), triangle id (extend output, int bits)
//   st0   float4  weight rgb, packed {iter (bits 0-7), next Philox draw index (bits 8-15)}
//   st1   float4  accumulated radiance rgb, unused
#pragma once
#include "traverse.cuh"

namespace cupt {

struct DevScene {
	const uint4* nodes;            // Node8 x5 uint4
	const float4* prims;           // Prim48 x3 float4
	const cupt_bvh2_node* bvh2;    // reference layout (parity mode)
	const float* verts;            // reference layout
	const cupt_triangle* tris;     // reference layout
	const float4* trishade;        // TriShade
	const float4* mats;            // MatF x2 float4
	const float4* lights;          // LightTri x3 float4, numLights + 1 entries (last = triangle 0 fallback)
	const cupt_bvh2_node* bvh2_up;  // strict mode: copy of the reference nodes with `left` = PARENT index (root: 0xffffffff): box and
	                                // next link of the ancestor walk in one 32-byte sector
	const uint32_t* bvh2_leaf_parent;  // strict mode: the inner node that holds each triangle's leaf link
	const float* light_cdf;        // inclusive prefix sums of the light areas (only used above CUPT_LIGHT_SCAN_MAX lights)
	const float4* spheres;         // SphereShade x3 float4: {centre, radius}, {diffuse, emissive flag}, {emission, 0}
	int32_t numLights;
	float totalLightArea;
	int32_t stack_cap;             // shared-memory stack entries per thread
};

struct PassParams {
	cupt_camera cam;
	uint64_t seed;
	int32_t first_sample;      // global index of s_local = 0
	int32_t samples;           // S
	int32_t tiles_x, tiles_y;
	int32_t tile0;             // first tile of the window this pass renders (cupt_set_tile_window; 0 = from the image's first tile)
	int32_t p_pad;             // tiles in the window * 32
	int32_t max_depth;         // NUM_BOUNCES made run-time
	int32_t fixed_lens;        // 1: lens draws fixed at (0.5, 0.5) (primary-hit hook)
	uint64_t inv_p_pad, inv_tiles_x;   // 2^64 / d + 1: n / d == umul64hi(inv, n) for every 32-bit n and d > 1 (Lemire & Kaser 2019)
	// (kept out of DevScene: k_extend takes that struct by value, and its register allocation moved with the struct's size —
	// 28 more bytes cost 1 ms of extend per frame)
	int32_t term_stop;             // 1: rays of a path's LAST segment carry an early-stop distance (k_shade, k_extend<.., VIS>)
	float emis_lo[3], emis_hi[3];  // bounds of every emissive primitive (empty box: lo > hi)
};

struct PathBuffers {
	float4* rayA;
	float4* rayB;
	float4* st0;
	float4* st1;
	uint32_t* queue[2];
	uint32_t* retrace;         // strict mode: slots whose hit the reference would not have reached (counter: counters[126])
	uint32_t* counters;        // [0..63] queue sizes per depth, [64..127] extend fetch cursors per depth
	unsigned long long* stats; // [0] rays, [1] nodes, [2] tris, [3] samples, [4] strict-mode re-traces
};

__device__ __forceinline__ void slot_to_pixel(const PassParams& pp, uint32_t slot, int& x, int& y, int& s_local)
{
	// two integer divisions by run-time constants: a multiply-high each instead of ~20 instructions (4 % of k_shade)
	s_local = (int)__umul64hi(pp.inv_p_pad, (uint64_t)slot);
	const uint32_t r = slot - (uint32_t)s_local * (uint32_t)pp.p_pad;
	const uint32_t tile = (r >> 5) + (uint32_t)pp.tile0, lane = r & 31u;
	const uint32_t ty = pp.tiles_x == 1 ? tile : (uint32_t)__umul64hi(pp.inv_tiles_x, (uint64_t)tile), tx = tile - ty * (uint32_t)pp.tiles_x;
	x = (int)(tx * 8u + (lane & 7u));
	y = (int)(ty * 4u + (lane >> 3));
}

// camera::cameraRay, camera.h:77-97 (with pxlToFilm camera.h:36-43); exact-op arithmetic
__device__ __forceinline__ void camera_ray(const cupt_camera& c, int px, int py, float u1, float u2, f3& o_out, f3& d_out)
{
	f3 prf;
	prf.x = rsub(rdiv((float)px, (float)c.pxlWidth), 0.5f);
	prf.y = rsub(rdiv((float)py, (float)c.pxlHeight), 0.5f);
	const float r = rmul(c.radius, __fsqrt_rn(u1));
	const float theta = (float)(2 * 3.14159 * (double)u2);
	const f3 o = mk3(rmul(r, cosf(theta)), rmul(r, sinf(theta)), 0.0f);
	prf.z = c.distFromFilm;
	prf = rdiv3(rmul3(prf, -c.focalLength), c.distFromFilm);
	o_out = radd3(o, mk3(c.pos.x, c.pos.y, c.pos.z));
	d_out = rnormalized(rsub3(prf, o));
}

// Queue compaction for a 256-thread block inside a block-uniform loop: warp ballots, one shared-memory exchange and
// ONE global atomic per block (a single L2 address retires about one atomic per clock, so one per warp — 5 M per
// frame on config 2 — cost milliseconds).  Returns the queue position of this thread's entry (valid only if `push`).
__device__ __forceinline__ uint32_t block_queue_reserve(bool push, uint32_t* counter, uint32_t* s_warp /* [9] */)
{
	const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
	const uint32_t m = __ballot_sync(0xffffffffu, push);
	if (lane == 0u) s_warp[w] = (uint32_t)__popc(m);
	__syncthreads();
	if (threadIdx.x == 0u) {
		uint32_t run = 0u;
		for (uint32_t i = 0; i < (blockDim.x >> 5); ++i) { const uint32_t c = s_warp[i]; s_warp[i] = run; run += c; }
		s_warp[8] = run ? atomicAdd(counter, run) : 0u;
	}
	__syncthreads();
	const uint32_t pos = s_warp[8] + s_warp[w] + (uint32_t)__popc(m & ((1u << lane) - 1u));
	__syncthreads();   // s_warp is reused by the next loop iteration
	return pos;
}

// The same for up to four entries per thread (k_bdpt_connect: one visibility ray per connection): ONE reservation — three
// barriers and one atomic — instead of four.  The block's entries are laid out connection by connection, as four separate
// reservations would (a chunk of 64 consecutive queue entries then holds rays of one kind: laid out warp by warp instead, the
// visibility extend was 0.3 ms per frame slower).  s_cnt: 4 x 8 + 1 words.
__device__ __forceinline__ void block_queue_reserve4(const bool push[4], uint32_t* counter, uint32_t* s_cnt /* [33] */, uint32_t pos[4])
{
	const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
	const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
	for (int c = 0; c < 4; ++c) {
		const uint32_t m = __ballot_sync(0xffffffffu, push[c]);
		pos[c] = (uint32_t)__popc(m & lt);
		if (lane == 0u) s_cnt[c * 8 + w] = (uint32_t)__popc(m);
	}
	__syncthreads();
	if (threadIdx.x == 0u) {
		uint32_t tot = 0u;
		for (uint32_t i = 0; i < 4u * (blockDim.x >> 5); ++i) { const uint32_t c = s_cnt[i]; s_cnt[i] = tot; tot += c; }   // (blockDim.x = 256: 8 warps, entries in (connection, warp) order)
		s_cnt[32] = tot ? atomicAdd(counter, tot) : 0u;
	}
	__syncthreads();
	const uint32_t base = s_cnt[32];
#pragma unroll
	for (int c = 0; c < 4; ++c) pos[c] += base + s_cnt[c * 8 + w];
	__syncthreads();   // s_cnt is reused by the next loop iteration
}

// ---- raygen --------------------------------------------------------------------------------------------------
// share_primary: additionally list the slots of the pass's first sample in queue[1] / counters[63] (cupt_render_opts.share_primary_hits)
__global__ void __launch_bounds__(256) k_raygen(PassParams pp, PathBuffers pb, const bool share_primary = false)
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
				float u1 = 0.5f, u2 = 0.5f;
				if (!pp.fixed_lens) {
					PhiloxStream rng;
					rng.init(pp.seed, (uint32_t)x, (uint32_t)y, (uint32_t)(pp.first_sample + sl));
					u1 = rng.draw(0); u2 = rng.draw(1);   // kernel.cu:547
				}
				f3 o, d;
				camera_ray(pp.cam, x, y, u1, u2, o, d);
				pb.rayA[slot] = make_float4(o.x, o.y, o.z, d.x);
				pb.rayB[slot] = make_float4(d.y, d.z, CUPT_MAX_FLOAT, __int_as_float(-1));
				pb.st0[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(0u | (2u << 8)));
				pb.st1[slot] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
			}
		}
		const uint32_t pos = block_queue_reserve(valid, &pb.counters[0], s_warp);
		if (valid) pb.queue[0][pos] = slot;
		if (share_primary && base < (uint32_t)pp.p_pad) {   // (the condition is block-uniform; a block that straddles p_pad sorts its threads out with `first`)
			const bool first = valid && slot < (uint32_t)pp.p_pad;
			const uint32_t p1 = block_queue_reserve(first, &pb.counters[63], s_warp);
			if (first) pb.queue[1][p1] = slot;
		}
	}
	if (blockIdx.x == 0 && threadIdx.x == 0)
		atomicAdd(&pb.stats[3], (unsigned long long)pp.cam.pxlWidth * pp.cam.pxlHeight * pp.samples);
}

// the hit of the pass's first sample of every pixel, copied to the other samples' records (share_primary_hits)
__global__ void __launch_bounds__(256) k_broadcast_primary(PassParams pp, PathBuffers pb)
{
	const uint32_t total = (uint32_t)pp.p_pad * (uint32_t)pp.samples;
	for (uint32_t slot = (uint32_t)pp.p_pad + blockIdx.x * blockDim.x + threadIdx.x; slot < total; slot += gridDim.x * blockDim.x) {
		const uint32_t src = slot % (uint32_t)pp.p_pad;
		reinterpret_cast<float2*>(&pb.rayB[slot])[1] = reinterpret_cast<const float2*>(&pb.rayB[src])[1];
	}
}

// ---- extend: closest hit for every queued path ---------------------------------------------------------------
// Persistent warps over a global queue cursor.  MODE 0 = wide BVH, 1 = reference BVH2 (parity mode).
//
// MODE 0 is warp-synchronous: every lane carries its own ray, the warp runs node step / primitive steps / pop in
// lock step and re-forms after every iteration, and
//   * a lane whose ray has ended is refilled from the queue as soon as fewer than `fetch_thresh` lanes are busy
//     (the first version traced 32 rays to the end per fetch and ran at 7 of 32 threads per instruction on
//     diffuse bounces, profiles/r1a_extend_metrics.md),
//   * primitive tests are shared: the (ray, primitive) pairs of the whole warp are listed in shared memory and
//     every lane tests one pair, whoever owns the ray (see the primitive phase below).
// Queue entries are handed out from warp-private chunks of CHUNK entries (one atomic per chunk).
struct ExtendTuning {
	int32_t fetch_thresh;   // refill idle lanes when active lanes <= this
	int32_t defer_pairs;    // hold primitive tests one iteration while the warp has fewer (ray, primitive) pairs than this
	uint32_t qbias_bits;    // CUPT_QBIAS_BITS, passed as a parameter on purpose (see u8f_biased)
};
#define CUPT_LIGHT_SCAN_MAX 64   // up to this many lights k_shade repeats the reference's linear scan exactly
#ifndef CUPT_QUEUE_CHUNK
#define CUPT_QUEUE_CHUNK 64u
#endif
#ifndef CUPT_EXTEND_MIN_BLOCKS
#define CUPT_EXTEND_MIN_BLOCKS 7
#endif
#ifndef CUPT_WAIT_LANES
#define CUPT_WAIT_LANES 32   // primitive phase of extend_wide: a round is forced once this many lanes wait for one (1 = the round-1 rule: any)
#endif
#ifndef CUPT_NODE_STEPS
#define CUPT_NODE_STEPS 2   // node steps (and the pops between them) per look at the primitive groups and the queue; 1 = one, the loop up to r2ba
#endif
#ifndef CUPT_NODE_FAST
#define CUPT_NODE_FAST 1   // 0: the hot kernel uses wide_node_step (the definition) instead of wide_node_step_fast
#endif

// VIS: the rays are BDPT visibility rays whose record carries an early-stop distance (bdpt.cuh); a separate
// instantiation because the extra register and compare cost the ordinary launches 3 % (measured)
template <bool STATS, bool VIS>
__device__ __forceinline__ void extend_wide(const DevScene& sc, const PathBuffers& pb, const uint32_t count, uint32_t* cursor,
                                            const uint32_t* __restrict__ queue, uint2* stack, const int stride,
                                            uint32_t* wlist, unsigned long long* wkeys, uint32_t* wids, const ExtendTuning tune, const uint8_t* perm, const uint32_t* nib,
                                            uint32_t& n_nodes, uint32_t& n_tris, uint32_t& n_rays)
{
	const uint32_t FULL = 0xffffffffu;
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t lt_mask = (1u << lane) - 1u;
	bool active = false, exhausted = false;
	uint32_t chunk_next = 0u, chunk_end = 0u;
	WideRay r;
	uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u), pgroup = make_uint2(0u, 0u);
	int sp = 0;
	uint32_t slot = 0u;
	float tstop = 0.0f;
	for (;;) {
		uint32_t act = __ballot_sync(FULL, active);
		if (!exhausted && __popc(act) <= tune.fetch_thresh) {
			uint32_t need = ~act;
			while (need != 0u) {
				if (chunk_next >= chunk_end) {
					uint32_t b = 0u;
					if (lane == 0u) b = atomicAdd(cursor, CUPT_QUEUE_CHUNK);
					b = __shfl_sync(FULL, b, 0);
					if (b >= count) { exhausted = true; break; }
					chunk_next = b; chunk_end = min(b + CUPT_QUEUE_CHUNK, count);
				}
				const uint32_t avail = chunk_end - chunk_next;
				const uint32_t my = __popc(need & lt_mask);
				if (((need >> lane) & 1u) && my < avail) {
					slot = queue[chunk_next + my];
					const float4 a = pb.rayA[slot];
					const float4 b = pb.rayB[slot];
					wide_ray_init(r, mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y));
					r.best = b.z;   // the producer's t limit: CUPT_MAX_FLOAT (kernel.cu:117) except for BDPT visibility rays
					if (VIS) tstop = b.w;   // stop as soon as a hit nearer than this is known
					ngroup = make_uint2(0u, 0x80000000u); tgroup = make_uint2(0u, 0u);
					pgroup = make_uint2(0u, 0u);
					sp = 0;
					active = true;
					++n_rays;
				}
				chunk_next += min((uint32_t)__popc(need), avail);
				need = ~__ballot_sync(FULL, active);
			}
			act = __ballot_sync(FULL, active);
		}
		if (act == 0u) break;

		// ---- node step ------------------------------------------------------------------------------------------
#if CUPT_NODE_STEPS > 1
		// CUPT_NODE_STEPS node steps (with the pops between them) per look at the primitive groups and the queue
#pragma unroll 1
		for (int rep = 0; rep < CUPT_NODE_STEPS; ++rep) {
			if (rep > 0 && active && !group_has_node(ngroup) && sp > 0) { --sp; ngroup = stack[sp * stride]; }
			if (active && group_has_node(ngroup) && !(tgroup.y != 0u && pgroup.y != 0u)) {
				if (tgroup.y != 0u) { pgroup = tgroup; }   // (pgroup is free here; the step overwrites tgroup)
				uint2 rest;
				wide_node_step_fast(sc.nodes, r, ngroup, rest, tgroup, tune.qbias_bits, perm + (r.oct_inv << 8), nib);
				if (group_has_node(rest)) { stack[sp * stride] = rest; ++sp; }
				if (STATS) ++n_nodes;
			}
		}
#else
#if CUPT_WAIT_LANES > 1
		if (active && group_has_node(ngroup) && !(tgroup.y != 0u && pgroup.y != 0u)) {   // (a lane holding two primitive groups waits for the flush)
#else
		if (active && group_has_node(ngroup)) {   // (only node groups are ever pushed, so a popped group always has one)
#endif
			uint2 rest;
#if CUPT_NODE_FAST
			wide_node_step_fast(sc.nodes, r, ngroup, rest, tgroup, tune.qbias_bits, perm + (r.oct_inv << 8), nib);
#else
			wide_node_step(sc.nodes, r, ngroup, rest, tgroup, tune.qbias_bits);
#endif
			if (group_has_node(rest)) { stack[sp * stride] = rest; ++sp; }
			if (STATS) ++n_nodes;
		}
#endif
		// ---- primitive tests, shared across the warp ----------------------------------------------------------------
		// Lanes with pending primitives are few (3.2 of 32 in profiles/r1c, where this phase took 40 % of the issue
		// slots), so the (ray, primitive) pairs of the whole warp are listed in shared memory and every lane tests
		// one pair: lane i fetches the owner's ray by shuffle, runs the exact test and merges its hit into the
		// owner's 64-bit key (t bits << 32 | rank) with a shared-memory atomicMin, which is the acceptance rule
		// `t < best || (t == best && rank < best_rank)` because positive floats order like their bit patterns.
		// A warp that holds fewer than `defer_pairs` pairs keeps them (tgroup -> pgroup).  A lane that cannot go on without its
		// tests — its ray would end, or it would need a third group — waits for the round instead of forcing one (round 1 forced
		// it, which was the case in most iterations: rounds at 15 lanes, now 23; profiles/r2ba_extend_lines.txt).
		{
			const uint32_t cnt0 = active ? (uint32_t)(__popc(tgroup.y) + __popc(pgroup.y)) : 0u;
			const uint32_t total0 = __reduce_add_sync(FULL, cnt0);
			if (total0 != 0u) {
				const bool last = !group_has_node(ngroup) && sp == 0;        // this ray ends after the iteration
				const bool clash = tgroup.y != 0u && pgroup.y != 0u;         // both slots taken
#if CUPT_WAIT_LANES > 1
				// a lane that cannot go on without its tests (its ray would end, or it holds two groups) WAITS instead of forcing a
				// round for the whole warp, until CUPT_WAIT_LANES lanes wait, the warp holds defer_pairs pairs or nobody else moves
				const bool waiting = cnt0 != 0u && (last || clash);
#if CUPT_WAIT_LANES >= 32
				const bool flush = total0 >= (uint32_t)tune.defer_pairs || !__any_sync(FULL, active && !waiting);
#else
				const uint32_t wmask = __ballot_sync(FULL, waiting);
				const bool flush = total0 >= (uint32_t)tune.defer_pairs || __popc(wmask) >= CUPT_WAIT_LANES || !__any_sync(FULL, active && !waiting);
#endif
#else
				const bool flush = total0 >= (uint32_t)tune.defer_pairs || __any_sync(FULL, cnt0 != 0u && (last || clash));
#endif
				if (!flush) {
					if (active && tgroup.y != 0u && pgroup.y == 0u) { pgroup = tgroup; tgroup.y = 0u; }
				} else {
					for (;;) {
						const uint32_t cnt = active ? (uint32_t)(__popc(tgroup.y) + __popc(pgroup.y)) : 0u;
						uint32_t incl = cnt;
#pragma unroll
						for (int ofs = 1; ofs < 32; ofs <<= 1) {
							const uint32_t v = __shfl_up_sync(FULL, incl, ofs);
							if (lane >= (uint32_t)ofs) incl += v;
						}
						const uint32_t total = __shfl_sync(FULL, incl, 31);
						uint32_t pos = incl - cnt;
						if (cnt != 0u) {
#if CUPT_NODE_FAST
							// groups are {node index, sparse record mask} (wide_node_step_fast): the entry names the node and the
							// bit; the lane that TESTS the pair turns them into a record index (sparse_record), so this serial
							// low-occupancy loop stays at one store per pair.  (The 64-bit entries overlay wlist + wids: wids
							// is written only after every lane has read its entry.)
							uint2* wl2 = reinterpret_cast<uint2*>(wlist);
							while (pgroup.y != 0u && pos < 32u) {
								const uint32_t k = __ffs(pgroup.y) - 1u;
								pgroup.y &= pgroup.y - 1u;
								wl2[pos++] = make_uint2(pgroup.x, (k << 5) | lane);
							}
							while (tgroup.y != 0u && pos < 32u) {
								const uint32_t k = __ffs(tgroup.y) - 1u;
								tgroup.y &= tgroup.y - 1u;
								wl2[pos++] = make_uint2(tgroup.x, (k << 5) | lane);
							}
#else
							while (pgroup.y != 0u && pos < 32u) {
								const uint32_t k = __ffs(pgroup.y) - 1u;
								pgroup.y &= pgroup.y - 1u;
								wlist[pos++] = ((pgroup.x + k) << 5) | lane;
							}
							while (tgroup.y != 0u && pos < 32u) {
								const uint32_t k = __ffs(tgroup.y) - 1u;
								tgroup.y &= tgroup.y - 1u;
								wlist[pos++] = ((tgroup.x + k) << 5) | lane;
							}
#endif
						}
						const unsigned long long mykey = ((unsigned long long)__float_as_uint(r.best) << 32) | r.best_rank;
						wkeys[lane] = mykey;
						__syncwarp();
						const bool work = lane < total;
#if CUPT_NODE_FAST
						const uint2 e2 = work ? reinterpret_cast<const uint2*>(wlist)[lane] : make_uint2(0u, 0u);
						const uint32_t owner = e2.y & 31u;
#else
						const uint32_t e = work ? wlist[lane] : 0u;
						const uint32_t owner = e & 31u;
#endif
						f3 po, pd;
						po.x = __shfl_sync(FULL, r.o.x, owner); po.y = __shfl_sync(FULL, r.o.y, owner); po.z = __shfl_sync(FULL, r.o.z, owner);
						pd.x = __shfl_sync(FULL, r.d.x, owner); pd.y = __shfl_sync(FULL, r.d.y, owner); pd.z = __shfl_sync(FULL, r.d.z, owner);
						unsigned long long hkey = ~0ull;
						uint32_t hid = 0u;
						if (work) {
#if CUPT_NODE_FAST
							const float4* pp = sc.prims + (size_t)sparse_record(node_records(sc.nodes, e2.x), e2.y >> 5) * 3u;   // (the node is an L1 hit)
#else
							const float4* pp = sc.prims + (size_t)(e >> 5) * 3u;
#endif
							const float4 r0 = ldg_nc_f4(pp + 0), r1 = ldg_nc_f4(pp + 1), r2 = ldg_nc_f4(pp + 2);
							const uint32_t kind = __float_as_uint(r2.w);
							float t;
							if (kind == 0u) t = tri_refexact(po, pd, r0, r1, r2);
							else if (kind == 2u) t = tri_watertight(po, pd, r0, r1, r2);
							else t = sphere_nearest(po, pd, r0, r1);
							if (0.0f < t) {
								hkey = ((unsigned long long)__float_as_uint(t) << 32) | __float_as_uint(r2.z);
								hid = __float_as_uint(r2.y) | ((kind & 1u) << 31);
								atomicMin(&wkeys[owner], hkey);
							}
							if (STATS) ++n_tris;
						}
						__syncwarp();
						if (hkey != ~0ull && wkeys[owner] == hkey) wids[owner] = hid;
						__syncwarp();
						const unsigned long long nk = wkeys[lane];
						if (nk != mykey) {
							r.best = __uint_as_float((uint32_t)(nk >> 32)); r.best_rank = (uint32_t)nk; r.best_id = (int32_t)wids[lane];
						}
						if (total <= 32u) break;   // everything was listed in this round
						__syncwarp();
					}
				}
			}
		}
		// ---- pop / finish -------------------------------------------------------------------------------------------
		if (VIS && active && r.best < tstop) { ngroup.y = 0u; tgroup.y = 0u; pgroup.y = 0u; sp = 0; }   // an occluder is known: whatever is nearer changes nothing
		if (active && !group_has_node(ngroup)) {
			if (sp == 0) {     // (a ray that ends has had its pending primitives flushed above)
#if CUPT_WAIT_LANES > 1
				if ((tgroup.y | pgroup.y) == 0u)
#endif
				{
					reinterpret_cast<float2*>(&pb.rayB[slot])[1] = make_float2(r.best, __int_as_float(r.best_id));
					active = false;
				}
			} else {
				--sp;
				ngroup = stack[sp * stride];
			}
		}
	}
}

template <int MODE, bool STATS, bool VIS>
__global__ void __launch_bounds__(128, CUPT_EXTEND_MIN_BLOCKS) k_extend(DevScene sc, PathBuffers pb, int depth, int qsel, ExtendTuning tune)
{
	extern __shared__ uint2 smem_stack[];
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t count = pb.counters[depth];
	uint32_t* cursor = &pb.counters[64 + depth];
	const uint32_t* __restrict__ queue = pb.queue[qsel];
	uint32_t n_nodes = 0, n_tris = 0, n_rays = 0;
	if (MODE == 0) {
		// front-to-back order of the hit inner children (wide_node_step_fast): perm[oct][mask] has bit (s ^ oct) set for every slot s in mask
		__shared__ uint8_t s_perm[8 * 256];
		for (uint32_t i = threadIdx.x; i < 8u * 256u; i += 128u) {   // (128 = the block size this kernel is launched with; a run-time stride costs a division)
			uint32_t v = 0u;
			for (uint32_t sl = 0; sl < 8u; ++sl)
				if ((i >> sl) & 1u) v |= 1u << (sl ^ (i >> 8));
			s_perm[i] = (uint8_t)v;
		}
		// nib[mask]: the nibbles of the slots in mask (layout.h: one nibble of valid-record bits per slot)
		__shared__ uint32_t s_nib[256];
		for (uint32_t i = threadIdx.x; i < 256u; i += 128u) {
			uint32_t v = 0u;
			for (uint32_t sl = 0; sl < 8u; ++sl)
				if ((i >> sl) & 1u) v |= 7u << (4u * sl);
			s_nib[i] = v;
		}
		__syncthreads();
		// shared memory: per-thread stacks [stack_cap][128] uint2, then per warp: 32 keys (u64), 32 list entries, 32 ids
		unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_stack + (size_t)sc.stack_cap * blockDim.x);
		uint32_t* lists = reinterpret_cast<uint32_t*>(keys + blockDim.x);
		const uint32_t w = threadIdx.x >> 5;
		extend_wide<STATS, VIS>(sc, pb, count, cursor, queue, smem_stack + threadIdx.x, blockDim.x, lists + 64u * w, keys + 32u * w,
		                   lists + 64u * w + 32u, tune, s_perm, s_nib, n_nodes, n_tris, n_rays);
	} else {
		for (;;) {
			uint32_t base = 0;
			if (lane == 0) base = atomicAdd(cursor, 32u);
			base = __shfl_sync(0xffffffffu, base, 0);
			if (base >= count) break;
			const uint32_t qi = base + lane;
			if (qi < count) {
				const uint32_t slot = queue[qi];
				const float4 a = pb.rayA[slot];
				const float4 b = pb.rayB[slot];
				Hit hit;
				traverse_bvh2_ref<STATS>(sc.bvh2, sc.verts, sc.tris, mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), hit, n_nodes, n_tris);
				reinterpret_cast<float2*>(&pb.rayB[slot])[1] = make_float2(hit.t, __int_as_float(hit.id));
				++n_rays;
			}
			__syncwarp();
		}
	}
	// one atomic per warp for the ray counter (rays are counted, not nominal: SURVEY.md §8d)
	for (int ofs = 16; ofs > 0; ofs >>= 1) {
		n_rays += __shfl_xor_sync(0xffffffffu, n_rays, ofs);
		if (STATS) { n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, ofs); n_tris += __shfl_xor_sync(0xffffffffu, n_tris, ofs); }
	}
	if (lane == 0) {
		if (n_rays) atomicAdd(&pb.stats[0], (unsigned long long)n_rays);
		if (STATS) { atomicAdd(&pb.stats[1], (unsigned long long)n_nodes); atomicAdd(&pb.stats[2], (unsigned long long)n_tris); }
	}
}

// ---- strict mode: make the wide result equal the reference's on EVERY ray -----------------------------------------
// The reference returns the argmin over the triangles its own traversal REACHES, and it reaches a triangle iff
// every ancestor box passes rayAABBIntersect (kernel.cu:139-151, BVH.h:51-83), which is not conservative.  The wide
// traversal returns the argmin over ALL triangles (ties by the reference's visiting order).  If that triangle is
// reached by the reference, it is also the reference's answer; so this kernel walks the hit triangle's ancestor chain
// in the reference tree with the reference's slab arithmetic, and only when an ancestor rejects the ray (about one
// ray in 5 400 on config 2) re-traces that ray with the reference's algorithm.  A miss needs no check (reached is a
// subset of all).  One thread per queued ray.
// Two kernels: k_verify_strict lists the rays whose hit the reference would not have reached (pb.retrace, counter
// counters[126]); k_retrace_strict re-traces those few with the reference's algorithm.  Keeping the re-trace (and its
// 256-byte stack) out of the checking kernel leaves that one small enough for full occupancy: it is bound by three
// dependent loads per ray (queue -> ray record -> leaf parent -> node).
__device__ __forceinline__ bool strict_reached(const DevScene& sc, const f3 o, const f3 d, const int32_t id)
{
	// Every comparison of rayAABBIntersect is monotone in the box, and IEEE rounding is monotone, so with no NaN in
	// play (no zero direction component) a box that passes makes every box that contains it pass.  The reference's
	// node boxes are nested (checked per node at upload: `right` = 1), so the test of the deepest ancestor decides;
	// only otherwise the chain is walked to the root, where a division-free test with an explicit error margin proves
	// "the reference accepts" for the ancestors that contain the hit point comfortably (quotients by multiplication
	// with 1/d are within 2^-22 relative of the reference's correctly rounded divisions).
	const bool no_nan = d.x != 0.0f && d.y != 0.0f && d.z != 0.0f;
	const float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
	for (uint32_t n = sc.bvh2_leaf_parent[id]; n != 0xffffffffu;) {
		const uint4 w0 = ldg_nc_u4(reinterpret_cast<const uint4*>(sc.bvh2_up + n)), w1 = ldg_nc_u4(reinterpret_cast<const uint4*>(sc.bvh2_up + n) + 1);
		cupt_bvh2_node nd;   // two 128-bit fetches (the struct itself is only 4-byte aligned and would load as eight words)
		nd.lo.x = __uint_as_float(w0.x); nd.lo.y = __uint_as_float(w0.y); nd.lo.z = __uint_as_float(w0.z);
		nd.hi.x = __uint_as_float(w0.w); nd.hi.y = __uint_as_float(w1.x); nd.hi.z = __uint_as_float(w1.y);
		nd.left = w1.z; nd.right = w1.w;
		if (no_nan && nd.right != 0u) return ref_slab_hit(o, d, nd);
		n = nd.left;
		const float x1 = rsub(nd.lo.x, o.x) * ix, x2 = rsub(nd.hi.x, o.x) * ix;
		const float y1 = rsub(nd.lo.y, o.y) * iy, y2 = rsub(nd.hi.y, o.y) * iy;
		const float z1 = rsub(nd.lo.z, o.z) * iz, z2 = rsub(nd.hi.z, o.z) * iz;
		const float txn = fminf(x1, x2), txf = fmaxf(x1, x2), tyn = fminf(y1, y2), tyf = fmaxf(y1, y2);
		const float tzn = fminf(z1, z2), tzf = fmaxf(z1, z2);
		const float m = fmaxf(fmaxf(fmaxf(fabsf(x1), fabsf(x2)), fmaxf(fabsf(y1), fabsf(y2))), fmaxf(fabsf(z1), fabsf(z2))) * 9.5367431640625e-7f;
		const bool sure = m < 1.0e30f && txn <= tyf - m && tyn <= txf - m && fmaxf(txn, tyn) <= tzf - m && tzn <= fminf(txf, tyf) - m;
		if (!sure && !ref_slab_hit(o, d, nd)) return false;
	}
	return true;
}

__global__ void __launch_bounds__(256) k_verify_strict(DevScene sc, PathBuffers pb, int depth, int qsel)
{
	const uint32_t count = pb.counters[depth];
	const uint32_t* __restrict__ queue = pb.queue[qsel];
	const uint32_t stride = gridDim.x * blockDim.x;
	// four rays per iteration, every link of the load chain (queue -> hit -> leaf parent -> node -> ray) issued for all
	// four before the next one is used (one ray at a time: 7.8 ms per config 2 frame; two: 6.2; four, phased: 5.7)
	constexpr int U = 4;
	for (uint32_t q0 = blockIdx.x * blockDim.x + threadIdx.x; q0 < count; q0 += (uint32_t)U * stride) {
		uint32_t sl[U], lp[U]; float4 rb[U], ra[U]; int32_t id[U]; uint4 w0[U], w1[U];
#pragma unroll
		for (int u = 0; u < U; ++u) { const uint32_t q = q0 + (uint32_t)u * stride; sl[u] = q < count ? queue[q] : 0xffffffffu; }
#pragma unroll
		for (int u = 0; u < U; ++u) { rb[u] = sl[u] != 0xffffffffu ? pb.rayB[sl[u]] : make_float4(0.f, 0.f, 0.f, __int_as_float(-1)); id[u] = __float_as_int(rb[u].w); }
#pragma unroll
		for (int u = 0; u < U; ++u) {
			lp[u] = id[u] >= 0 ? sc.bvh2_leaf_parent[id[u]] : 0xffffffffu;
			ra[u] = id[u] >= 0 ? pb.rayA[sl[u]] : make_float4(0.f, 0.f, 0.f, 0.f);
		}
#pragma unroll
		for (int u = 0; u < U; ++u) {
			w0[u] = w1[u] = make_uint4(0u, 0u, 0u, 0u);
			if (lp[u] != 0xffffffffu) { w0[u] = ldg_nc_u4(reinterpret_cast<const uint4*>(sc.bvh2_up + lp[u])); w1[u] = ldg_nc_u4(reinterpret_cast<const uint4*>(sc.bvh2_up + lp[u]) + 1); }
		}
#pragma unroll
		for (int u = 0; u < U; ++u) {
			if (id[u] < 0) continue;
			const f3 o = mk3(ra[u].x, ra[u].y, ra[u].z), d = mk3(ra[u].w, rb[u].x, rb[u].y);
			bool reached;
			if (w1[u].w != 0u && d.x != 0.0f && d.y != 0.0f && d.z != 0.0f) {   // nested chain, no NaN: this one box decides
				cupt_bvh2_node nd;
				nd.lo.x = __uint_as_float(w0[u].x); nd.lo.y = __uint_as_float(w0[u].y); nd.lo.z = __uint_as_float(w0[u].z);
				nd.hi.x = __uint_as_float(w0[u].w); nd.hi.y = __uint_as_float(w1[u].x); nd.hi.z = __uint_as_float(w1[u].y);
				nd.left = nd.right = 0u;
				reached = ref_slab_hit(o, d, nd);
			} else {
				reached = strict_reached(sc, o, d, id[u]);
			}
			if (!reached) pb.retrace[atomicAdd(&pb.counters[126], 1u)] = sl[u];
		}
	}
}

// The reference's answer for a listed ray = argmin over the triangles its traversal REACHES (it has no t range, so it
// tests every triangle of every node whose box passes; equal t: the first visited, i.e. the lowest rank).  Instead of
// repeating that traversal (hundreds of dependent node fetches per ray: 0.5 ms per launch for a few thousand rays),
// walk the wide BVH and accept a hit only if the reference reaches its triangle (strict_reached).
__global__ void __launch_bounds__(64) k_retrace_strict(DevScene sc, PathBuffers pb)
{
	extern __shared__ uint2 rt_stack[];
	uint2* stack = rt_stack + threadIdx.x;
	const int stride = (int)blockDim.x;
	const uint32_t count = pb.counters[126];
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
		const uint32_t slot = pb.retrace[i];
		const float4 a = pb.rayA[slot], b = pb.rayB[slot];
		const f3 o = mk3(a.x, a.y, a.z), d = mk3(a.w, b.x, b.y);
		WideRay r;
		wide_ray_init(r, o, d);
		uint2 ngroup = make_uint2(0u, 0x80000000u), tgroup = make_uint2(0u, 0u);
		int sp = 0;
		for (;;) {
			if (group_has_node(ngroup)) {
				uint2 rest;
				wide_node_step(sc.nodes, r, ngroup, rest, tgroup);
				if (group_has_node(rest)) { if (sp < sc.stack_cap) stack[sp * stride] = rest; ++sp; }
			} else {
				tgroup = ngroup;
				ngroup = make_uint2(0u, 0u);
			}
			while (tgroup.y != 0u) {
				const uint32_t k = __ffs(tgroup.y) - 1u;
				tgroup.y &= tgroup.y - 1u;
				const float4* pp = sc.prims + (size_t)(tgroup.x + k) * 3u;
				const float4 r0 = ldg_nc_f4(pp + 0), r1 = ldg_nc_f4(pp + 1), r2 = ldg_nc_f4(pp + 2);
				const float t = tri_refexact(o, d, r0, r1, r2);   // (strict mode: triangles only, the reference's own test)
				const uint32_t rank = __float_as_uint(r2.z);
				const int32_t id = (int32_t)__float_as_uint(r2.y);
				if (0.0f < t && (t < r.best || (t == r.best && rank < r.best_rank)) && strict_reached(sc, o, d, id)) {
					r.best = t; r.best_rank = rank; r.best_id = id;
				}
			}
			if (!group_has_node(ngroup)) {
				if (sp == 0) break;
				--sp;
				ngroup = stack[sp * stride];
			}
		}
		reinterpret_cast<float2*>(&pb.rayB[slot])[1] = make_float2(r.best, __int_as_float(r.best_id));
	}
	if (blockIdx.x == 0 && threadIdx.x == 0 && count) atomicAdd(&pb.stats[4], (unsigned long long)count);
}

// Early stop for the LAST segment of a path.  All the next k_shade does with such a ray's hit is add the emission of the
// primitive it meets (kernel.cu:453-457), so a hit in front of every emitter ends the search: whatever is nearer is not an
// emitter either.  Returns a distance d such that no emissive primitive is met before d along the ray: the entry distance
// of the emitters' bounding box, shortened by 0.1 % + 0.001 against the rounding of the box test and of triIntersect
// (<= 0: no early stop possible), or CUPT_MAX_FLOAT — any hit ends the search — when the ray passes the box by at least that
// margin or the scene has no emitter.  k_extend<.., VIS> stops once best < d.
__device__ __forceinline__ float emitter_entry_distance(const PassParams& pp, const f3 o, const f3 d)
{
	if (pp.emis_lo[0] > pp.emis_hi[0]) return CUPT_MAX_FLOAT;
	float tin = 0.0f, tout = CUPT_MAX_FLOAT;
	// (an axis the ray runs along is left out: that only widens [tin, tout])
	if (fabsf(d.x) > 1.0e-20f) {
		const float inv = box_rcp(d.x), t1 = (pp.emis_lo[0] - o.x) * inv, t2 = (pp.emis_hi[0] - o.x) * inv;
		tin = fmaxf(tin, fminf(t1, t2)); tout = fminf(tout, fmaxf(t1, t2));
	}
	if (fabsf(d.y) > 1.0e-20f) {
		const float inv = box_rcp(d.y), t1 = (pp.emis_lo[1] - o.y) * inv, t2 = (pp.emis_hi[1] - o.y) * inv;
		tin = fmaxf(tin, fminf(t1, t2)); tout = fminf(tout, fmaxf(t1, t2));
	}
	if (fabsf(d.z) > 1.0e-20f) {
		const float inv = box_rcp(d.z), t1 = (pp.emis_lo[2] - o.z) * inv, t2 = (pp.emis_hi[2] - o.z) * inv;
		tin = fmaxf(tin, fminf(t1, t2)); tout = fminf(tout, fmaxf(t1, t2));
	}
	const float near = tin * 0.999f - 0.001f;
	if (fmaf(fabsf(tout), 0.001f, tout) + 0.001f < near) return CUPT_MAX_FLOAT;   // the box is behind the ray or beside it
	return fminf(near, CUPT_MAX_FLOAT);
}

// ---- shade: one iteration of radianceAlongSingleStep2 (kernel.cu:424-512) per queued path ----------------------
// Consumes the hit, adds emission, draws the 50/50 choice (kernel.cu:459-460) between a cosine-weighted bounce
// (kernel.cu:462-464) and a light-point sample (kernel.cu:468-507), writes the next segment and pushes the path
// into the next queue (warp ballot + one atomic per warp).  A path whose weight is exactly zero can never
// contribute again (the weight is only ever multiplied, kernel.cu:464,506), so it is dropped instead of being
// traced to the end as the reference does; the light-sample branch jumps to the last iteration exactly like
// `i = max(i, NUM_BOUNCES - 2)` (kernel.cu:507), which makes the next extend the shadow test.
__device__ __forceinline__ f3 get_tangent(const f3 n)  // kernel.cu:44-54
{
	const f3 c1 = rcross(n, mk3(0.f, 0.f, 1.f)), c2 = rcross(n, mk3(0.f, 1.f, 0.f));
	return (rdot(c1, c1) > rdot(c2, c2)) ? c1 : c2;
}

#ifndef CUPT_SHADE_MIN_BLOCKS
#define CUPT_SHADE_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(256, CUPT_SHADE_MIN_BLOCKS) k_shade(DevScene sc, PassParams pp, PathBuffers pb, int depth, int qsel)
{
	__shared__ uint32_t s_warp[9];
	const uint32_t count = pb.counters[depth];
	const uint32_t* __restrict__ qin = pb.queue[qsel];
	uint32_t* __restrict__ qout = pb.queue[qsel ^ 1];
	const float kBrdf = (float)(1 / 3.14159);   // BRDF(): mul(albedo, 1 / 3.14159) with a float parameter (kernel.cu:101-104)
	const float kPi = (float)3.14159;            // color::operator*(const float&) (kernel.cu:463)
	// the queue entry of the NEXT iteration is fetched one iteration ahead: the kernel is bound by the chain of
	// dependent loads queue -> path records -> triangle -> material (profiles/r1h_shade_metrics.md: 36 % long-scoreboard
	// stalls at 34 % of the DRAM peak), and this takes the first link out of it (7.03 -> 6.86 ms per frame; prefetching the
	// path records as well costs registers and gains nothing: 79 registers / 3 blocks per SM 7.55 ms, 64 with spills 6.92 ms;
	// forcing 5 / 6 / 8 blocks per SM by spilling: 7.66 / 8.20 / 8.83 ms)
	uint32_t slot_next = 0;
	{
		const uint32_t q0 = blockIdx.x * blockDim.x + threadIdx.x;
		if (q0 < count) slot_next = qin[q0];
	}
	for (uint32_t base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
		const uint32_t qi = base + threadIdx.x;
		bool push = false;
		uint32_t slot = slot_next;
		{
			const uint32_t qn = qi + gridDim.x * blockDim.x;
			if (qn < count) slot_next = qin[qn];
		}

		if (qi < count) {
			const float4 ra = pb.rayA[slot], rb = pb.rayB[slot], s0 = pb.st0[slot];
			const f3 o = mk3(ra.x, ra.y, ra.z), d = mk3(ra.w, rb.x, rb.y);
			float wr = s0.x, wg = s0.y, wb = s0.z;
			const uint32_t packed = __float_as_uint(s0.w);
			int iter = (int)(packed & 0xffu);
			uint32_t draw = (packed >> 8) & 0xffu;
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
				// sphere (sphere.h:7-13 is a POD without an intersector or shading code: parity unpinned).  Normal =
				// (hit point - centre) / radius at the UNSHORTENED hit distance; diffuse / emm play albedo / emission.
				const uint32_t si = (uint32_t)id & 0x7fffffffu;
				const float4 sp0 = sc.spheres[3 * si];
				const f3 hp = radd3(o, rmul3(d, rb.z));
				normal = rdiv3(rsub3(hp, mk3(sp0.x, sp0.y, sp0.z)), sp0.w);
				m0 = sc.spheres[3 * si + 1]; m1 = sc.spheres[3 * si + 2];
			}
			if (m0.w != 0.f) {                                                      // kernel.cu:453-457
				// The radiance record is touched only here, and at most once per path: its weight is zero after (the path is dropped
				// below), and k_raygen zeroed the record.  So this is a plain store of 0 + w * emission, not a read-modify-write: the
				// read was the fifth dependent access of the chain queue -> records -> triangle -> material -> radiance record and 10 %
				// of the kernel's stall samples (ncu capture before this change; profiles/r2bc_shade_lines.txt is after).  (0 + x: a zero product keeps the sign the sum had.)
				pb.st1[slot] = make_float4(fmaf(wr, m1.x, 0.0f), fmaf(wg, m1.y, 0.0f), fmaf(wb, m1.z, 0.0f), 0.0f);
				wr = wg = wb = 0.f;
			}
			const bool dead = !(wr != 0.f || wg != 0.f || wb != 0.f);
			if (!(dead || iter >= pp.max_depth - 1)) {
				int x, y, sl;
				slot_to_pixel(pp, slot, x, y, sl);
				PhiloxStream rng;
				rng.init(pp.seed, (uint32_t)x, (uint32_t)y, (uint32_t)(pp.first_sample + sl));
				// the (at most) four uniforms of this vertex, drawn BEFORE the 50/50 branch: they span at most two Philox
				// blocks, and evaluating the second one inside both divergent sides ran its ~100 instructions twice per warp
				const float ud = rng.draw(draw), ua = rng.draw(draw + 1u), ub = rng.draw(draw + 2u), uc = rng.draw(draw + 3u);
				// every branch ends in `lDir = normalised(vdir)` and `weight *= albedo * k`: the branches only make vdir (and k)
				// and the square root, three IEEE divisions and three products run once, after they have rejoined
				f3 vdir;
				float k = 1.0f;
				bool light = false;
				if (m1.w != 0.f && id >= 0) {
					// perfect mirror (extension, config 5; no reference code): d - 2 (d.n) n, weight *= albedo, no random numbers
					vdir = rsub3(d, rmul3(normal, rmul(2.0f, rdot(d, normal))));
				} else if (ud < 0.5f) {                                              // kernel.cu:459
					// cosineWeightedRay, kernel.cu:78-99
					draw += 3u;
					const float u1 = ua, u2 = ub;
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
					float randArea = rmul(sc.totalLightArea, ua);
					int sel = sc.numLights;   // fallback entry = triangle 0 (`selectedTri = 0`, kernel.cu:469)
					if (sc.numLights <= CUPT_LIGHT_SCAN_MAX) {
						for (int j = 0; j < sc.numLights; ++j) {   // the reference's linear scan, kernel.cu:470-481, op for op
							const float area = sc.lights[3 * j].w;
							if (randArea < area && randArea > 0.f) sel = j;
							randArea = rsub(randArea, area);
						}
					} else {
						// The scan is O(numLights) per sample (58 k lights on the 30 M-triangle soup: 95 % of the frame time).
						// Above CUPT_LIGHT_SCAN_MAX lights the same selection is made by binary search in the prefix sums:
						// first j with randArea < cdf[j].  It differs from the sequential f32 subtraction only when randArea
						// lies within rounding of a light boundary (probability ~ numLights * 1e-7 per draw).
						int lo = 0, hi = sc.numLights;
						while (lo < hi) {
							const int mid = (lo + hi) >> 1;
							if (randArea < sc.light_cdf[mid]) hi = mid; else lo = mid + 1;
						}
						sel = lo;   // == numLights (fallback) when randArea >= cdf[last]
					}
					float u = ub, v = uc;
					const float4 l0 = sc.lights[3 * sel], l1 = sc.lights[3 * sel + 1], l2 = sc.lights[3 * sel + 2];
					if (radd(u, v) > 1.0f) {                                          // kernel.cu:491-495
						u = (float)((double)u + 2 * (0.5 - (double)u));
						v = (float)((double)v + 2 * (0.5 - (double)v));
					}
					const f3 pos1 = radd3(radd3(mk3(l0.x, l0.y, l0.z), rmul3(mk3(l1.x, l1.y, l1.z), u)), rmul3(mk3(l2.x, l2.y, l2.z), v));
					vdir = rsub3(pos1, pos);
					light = true;
				}
				const f3 lDir = rnormalized(vdir);
				if (light) {
					const float cosL = fmaxf(0.0f, rdot(lDir, normal));
					const float cosO1 = fmaxf(0.0f, rdot(mk3(0.f, -1.f, 0.f), rmul3(lDir, -1.f)));   // hard-coded light normal, kernel.cu:503
					const float G = rdiv(rmul(cosL, cosO1), rdot(vdir, vdir));
					k = kBrdf * G * sc.totalLightArea;                               // kernel.cu:506
					iter = max(iter, pp.max_depth - 2);                              // kernel.cu:507
				}
				if (m1.w != 0.f && id >= 0) { wr *= m0.x; wg *= m0.y; wb *= m0.z; }   // (mirror: the albedo alone, not albedo * 1.0f: same value, but keep the operation the test's restatement has)
				else { wr *= m0.x * k; wg *= m0.y * k; wb *= m0.z * k; }
				++iter;
				push = (wr != 0.f || wg != 0.f || wb != 0.f);
				if (push) {
					pb.rayA[slot] = make_float4(pos.x, pos.y, pos.z, lDir.x);
					// .w: early-stop distance of a last segment (NaN bits = none: `best < NaN` is never true)
					const float stop = (pp.term_stop && iter >= pp.max_depth - 1) ? emitter_entry_distance(pp, pos, lDir) : __int_as_float(-1);
					pb.rayB[slot] = make_float4(lDir.y, lDir.z, CUPT_MAX_FLOAT, stop);
					pb.st0[slot] = make_float4(wr, wg, wb, __uint_as_float((uint32_t)iter | (draw << 8)));
				}
			}
		}
		const uint32_t pos = block_queue_reserve(push, &pb.counters[depth + 1], s_warp);
		if (push) qout[pos] = slot;
	}
}

// ---- accumulate: sum the pass's samples per pixel in a fixed order, add to the f32 sum image ------------------
// (the reference keeps an f64 running mean per sample launch, kernel.cu:551-552)
__global__ void __launch_bounds__(256) k_accumulate(PassParams pp, PathBuffers pb, float* __restrict__ accum)
{
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= (uint32_t)pp.p_pad) return;
	int x, y, sl;
	slot_to_pixel(pp, r, x, y, sl);
	if (x >= pp.cam.pxlWidth || y >= pp.cam.pxlHeight) return;
	float sr = 0.f, sg = 0.f, sb = 0.f;
	for (int s = 0; s < pp.samples; ++s) {
		const float4 v = pb.st1[(size_t)s * pp.p_pad + r];
		sr += v.x; sg += v.y; sb += v.z;
	}
	float* a = accum + 3 * ((size_t)y * pp.cam.pxlWidth + x);
	a[0] += sr; a[1] += sg; a[2] += sb;
}

// CUPT_TRI_WATERTIGHT: a second primitive array in the same order whose triangle records hold the three exact
// vertices (kind 2) instead of v0, e1, e2
__global__ void __launch_bounds__(256) k_vertex_prims(const float4* __restrict__ prims, const int64_t n, const float* __restrict__ verts,
                                                      const cupt_triangle* __restrict__ tris, float4* __restrict__ out)
{
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 r0 = prims[3 * i], r1 = prims[3 * i + 1], r2 = prims[3 * i + 2];
	if (__float_as_uint(r2.w) == 0u) {
		const cupt_triangle t = tris[__float_as_uint(r2.y)];
		const float* a = verts + 3 * (size_t)t.v0; const float* b = verts + 3 * (size_t)t.v1; const float* c = verts + 3 * (size_t)t.v2;
		r0 = make_float4(a[0], a[1], a[2], b[0]);
		r1 = make_float4(b[1], b[2], c[0], c[1]);
		r2 = make_float4(c[2], r2.y, r2.z, __uint_as_float(2u));
	}
	out[3 * i] = r0; out[3 * i + 1] = r1; out[3 * i + 2] = r2;
}

// ---- parity hooks ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gather_primary(PassParams pp, PathBuffers pb, int32_t* __restrict__ ids, float* __restrict__ ts,
                                                         float* __restrict__ rays6)
{
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= (uint32_t)pp.p_pad) return;
	int x, y, sl;
	slot_to_pixel(pp, r, x, y, sl);
	if (x >= pp.cam.pxlWidth || y >= pp.cam.pxlHeight) return;
	const size_t p = (size_t)y * pp.cam.pxlWidth + x;
	const float4 a = pb.rayA[r], b = pb.rayB[r];
	if (ids) { ids[p] = __float_as_int(b.w); ts[p] = b.z; }
	if (rays6) { float* o = rays6 + 6 * p; o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; }
}

__global__ void __launch_bounds__(256) k_load_rays(const float* __restrict__ rays6, uint32_t n, uint32_t slot0, PathBuffers pb)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float* r = rays6 + 6 * (size_t)i;
	pb.rayA[slot0 + i] = make_float4(r[0], r[1], r[2], r[3]);
	pb.rayB[slot0 + i] = make_float4(r[4], r[5], CUPT_MAX_FLOAT, __int_as_float(-1));
	pb.queue[0][i] = slot0 + i;
}

__global__ void __launch_bounds__(256) k_store_hits(uint32_t n, uint32_t slot0, PathBuffers pb, int32_t* __restrict__ ids, float* __restrict__ ts)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float4 b = pb.rayB[slot0 + i];
	ids[i] = __float_as_int(b.w); ts[i] = b.z;
}

}  // namespace cupt
