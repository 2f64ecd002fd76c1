// GPU-side construction of the wide quantised BVH: replaces buildBVH (BVH.h:133-474), the step
// immediately before the hot path, with kernels that read the scene arrays cupt_set_scene already put in HBM and
// write the Node8 / Prim48 arrays k_extend traverses, without a host round trip.
//
//   1. k_leaf_boxes     padded box per primitive (same 4-ulp padding as the host builder), centroid bounds
//   2. k_morton         63-bit Morton code of the centroid, radix sort of (code, primitive)      [radix_sort.cuh]
//   3. PLOC rounds      parallel locally-ordered clustering (Meister & Bittner, "Parallel Locally-Ordered Clustering
//                       for Bounding Volume Hierarchy Construction", TVCG 2018): every cluster finds the neighbour
//                       within +-R positions of the Morton order (R = 4 by default: measured best on config 2, DESIGN.md
//                       §4) whose union box has the smallest area; mutual pairs
//                       merge; an order-preserving compaction closes the gaps.  ONE launch per round (k_ploc_round:
//                       neighbour search in shared memory, decoupled look-back for the compaction, merge), ~50 rounds
//                       at 100 M primitives, the last ~14 inside one block (k_ploc_tail).
//   4. collapse + emit  level by level from the root: the cost-optimal cut of every wide node (dynamic programme over
//                       slot budgets, filled in as nodes are created; same rule as the host builder),
//                       octant-slot assignment, 8-bit quantisation, breadth-first emission (k_collapse_count,
//                       k_scan_tiles, k_collapse_emit per level), then the 48-byte records in tree order (k_emit_prims).
//
// The tree differs from the host builder's (binned SAH), the traversal result does not: k_extend returns the argmin
// over all primitives whatever the tree (tests/test_gpu_parity.py compares with the brute force bit for bit).
// Node numbering is deterministic (positions come from prefix sums, not from atomics).
#include "host_common.h"
#include "layout.h"
#include "bvh_device.h"
#include "radix_sort.cuh"
#include <cuda_runtime.h>
#include <algorithm>
#include <cstring>

namespace cupt {
namespace devbuild {

constexpr int kTile = 256;        // threads per block = elements per scan tile

// binary node: leaves are 0..N-1 in Morton order (left = -1 - primitive), inner nodes N..2N-2
struct alignas(16) N2 {
	float lo[3]; int32_t left;
	float hi[3]; int32_t right;
};

__device__ __forceinline__ float half_area(const float lx, const float ly, const float lz, const float hx, const float hy, const float hz)
{
	const float dx = hx - lx, dy = hy - ly, dz = hz - lz;
	return dx * dy + dx * dz + dy * dz;
}

// order-preserving float <-> uint for atomicMin/atomicMax
__device__ __forceinline__ uint32_t f2ord(const float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float ord2f(const uint32_t o) { const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o; float f; memcpy(&f, &u, 4); return f; }

// ---- 1. primitive boxes ---------------------------------------------------------------------------------------
// Grid-stride over the primitives with the centroid bounds kept in registers: one atomic pair per BLOCK and axis at the end
// (one pair per warp — 5.6 M atomics on six addresses for 30 M primitives — made this kernel 4.7 ms where its 3 GB of
// traffic need 0.5 ms).
// Also written here, while the triangle and its vertices are in registers anyway: the primitive's 48-byte traversal record, in
// primitive order.  k_emit_prims then moves ONE 48-byte piece per record into tree order; round 1 gathered the triangle and its
// three vertices a second time there (2.3 cache lines per record on the soups, more on a mesh with shared vertices: 30 GB of DRAM
// traffic and 6.8 ms for 100 M triangles).
__global__ void __launch_bounds__(kTile) k_leaf_boxes(const float* __restrict__ verts, const cupt_triangle* __restrict__ tris, const int64_t nt,
                                                      const float4* __restrict__ spheres, const uint32_t* __restrict__ rank, const int64_t N,
                                                      N2* __restrict__ leaf, float4* __restrict__ records, uint32_t* __restrict__ cbounds)
{
	uint32_t mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0u, 0u, 0u};
	for (int64_t i = (int64_t)blockIdx.x * kTile + threadIdx.x; i < N; i += (int64_t)gridDim.x * kTile) {
		float lo[3], hi[3];
		float4 r0, r1, r2;
		if (i < nt) {
			const cupt_triangle t = tris[i];
			const int32_t vi[3] = {t.v0, t.v1, t.v2};
			float v[3][3];
			for (int a = 0; a < 3; ++a) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; }
			for (int k = 0; k < 3; ++k)
				for (int a = 0; a < 3; ++a) { v[k][a] = verts[3 * (size_t)vi[k] + a]; lo[a] = fminf(lo[a], v[k][a]); hi[a] = fmaxf(hi[a], v[k][a]); }
			// the record: v0, e1 = v1 - v0, e2 = v2 - v0 (modelLoader.h:57-58), primitive, rank, kind 0
			r0 = make_float4(v[0][0], v[0][1], v[0][2], __fsub_rn(v[1][0], v[0][0]));
			r1 = make_float4(__fsub_rn(v[1][1], v[0][1]), __fsub_rn(v[1][2], v[0][2]), __fsub_rn(v[2][0], v[0][0]), __fsub_rn(v[2][1], v[0][1]));
			r2 = make_float4(__fsub_rn(v[2][2], v[0][2]), __uint_as_float((uint32_t)i), __uint_as_float(rank ? rank[i] : (uint32_t)i), __uint_as_float(0u));
		} else {
			const float4 s = spheres[3 * (i - nt)];   // SphereShade: centre, radius
			const float cc[3] = {s.x, s.y, s.z};
			for (int a = 0; a < 3; ++a) { lo[a] = __fsub_rn(cc[a], s.w); hi[a] = __fadd_rn(cc[a], s.w); }
			r0 = s;
			r1 = make_float4(0.f, 0.f, 0.f, 0.f);
			r2 = make_float4(0.f, __uint_as_float((uint32_t)(i - nt)), __uint_as_float((uint32_t)i), __uint_as_float(1u));
		}
		records[3 * i] = r0; records[3 * i + 1] = r1; records[3 * i + 2] = r2;
		N2 n;
		for (int a = 0; a < 3; ++a) {
			const float c = __fadd_rn(__fmul_rn(0.5f, lo[a]), __fmul_rn(0.5f, hi[a]));
			const uint32_t co = f2ord(c);
			mn[a] = min(mn[a], co); mx[a] = max(mx[a], co);
			const float m = fmaxf(fmaxf(fabsf(lo[a]), fabsf(hi[a])), 1e-30f);
			const float pad = __fmul_rn(m, 4.8e-7f);   // 4 ulp, as bvh8_build.cpp: the reference's triangle test is plain f32
			n.lo[a] = __fsub_rn(lo[a], pad); n.hi[a] = __fadd_rn(hi[a], pad);
		}
		n.left = -1 - (int32_t)i; n.right = -1;
		leaf[i] = n;
	}
	// centroid bounds: warp reduce, block reduce through shared memory, one atomic pair per block and axis
	__shared__ uint32_t s_mn[3][kTile / 32], s_mx[3][kTile / 32];
	for (int a = 0; a < 3; ++a) {
		const uint32_t wn = __reduce_min_sync(0xffffffffu, mn[a]), wx = __reduce_max_sync(0xffffffffu, mx[a]);
		if ((threadIdx.x & 31) == 0) { s_mn[a][threadIdx.x >> 5] = wn; s_mx[a][threadIdx.x >> 5] = wx; }
	}
	__syncthreads();
	if (threadIdx.x < 3) {
		uint32_t bn = 0xffffffffu, bx = 0u;
		for (int w = 0; w < kTile / 32; ++w) { bn = min(bn, s_mn[threadIdx.x][w]); bx = max(bx, s_mx[threadIdx.x][w]); }
		atomicMin(&cbounds[threadIdx.x], bn); atomicMax(&cbounds[3 + threadIdx.x], bx);
	}
}

// ---- 2. Morton codes -------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread21(uint64_t x)
{
	x &= 0x1fffffull;
	x = (x | x << 32) & 0x1f00000000ffffull;
	x = (x | x << 16) & 0x1f0000ff0000ffull;
	x = (x | x << 8) & 0x100f00f00f00f00full;
	x = (x | x << 4) & 0x10c30c30c30c30c3ull;
	x = (x | x << 2) & 0x1249249249249249ull;
	return x;
}

__global__ void __launch_bounds__(kTile) k_morton(const N2* __restrict__ leaf, const int64_t N, const uint32_t* __restrict__ cbounds,
                                                  uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
	const int64_t i = (int64_t)blockIdx.x * kTile + threadIdx.x;
	if (i >= N) return;
	const N2 n = leaf[i];
	uint64_t code = 0;
	for (int a = 0; a < 3; ++a) {
		const float lo = ord2f(cbounds[a]), hi = ord2f(cbounds[3 + a]);
		const float c = 0.5f * n.lo[a] + 0.5f * n.hi[a];
		const float ext = hi - lo;
		float u = ext > 0.f ? (c - lo) / ext : 0.f;
		u = fminf(fmaxf(u, 0.f), 1.f);
		const uint64_t q = (uint64_t)fminf(u * 2097152.0f, 2097151.0f);
		code |= spread21(q) << (2 - a);
	}
	keys[i] = code; vals[i] = (uint32_t)i;
}

// ---- the binary tree ---------------------------------------------------------------------------------------------------------
// Leaves are nodes 0..N-1 in Morton order (32 bytes: box, left = -1 - primitive); inner node id >= N is inner[id - N], ONE 64-byte
// record: box, children, the collapse costs of its subtree (Dp below) and, in bits 24..26 of `splits`, its primitive count
// saturated at 7 (all anybody asks is "at most max_leaf <= 3?" and, if so, how many).  Round 1 kept boxes + children, costs and
// counts in three arrays: every node visited by the clustering's merge or by the collapse's cut walk was three random 128-byte
// line fetches instead of one (profiles/r2w_builder_metrics.md: k_collapse_count at 66 % of the DRAM peak for 740 bytes per item).
struct alignas(16) Inner { float lo[3]; int32_t left; float hi[3]; int32_t right; float c[7]; uint32_t splits; };
static_assert(sizeof(Inner) == 64, "one inner node = half a cache line");
struct Tree {
	N2* leaf; Inner* inner; uint32_t N;
	__device__ __forceinline__ const float4* rec(const uint32_t id) const
	{
		return id < N ? reinterpret_cast<const float4*>(leaf + id) : reinterpret_cast<const float4*>(inner + (id - N));
	}
	__device__ __forceinline__ int32_t left(const uint32_t id) const { return id < N ? leaf[id].left : inner[id - N].left; }
	__device__ __forceinline__ int32_t right(const uint32_t id) const { return id < N ? leaf[id].right : inner[id - N].right; }
	__device__ __forceinline__ uint32_t count(const uint32_t id) const { return id < N ? 1u : (inner[id - N].splits >> 24) & 7u; }
};

// a cluster's box in the per-round cluster list: 24 bytes, read and written as three 8-byte words
struct alignas(8) CBox { float lo[3], hi[3]; };
__device__ __forceinline__ void cbox_load(const CBox* p, float4& a, float4& b)
{
	const float2* q = reinterpret_cast<const float2*>(p);
	const float2 x = q[0], y = q[1], z = q[2];
	a = make_float4(x.x, x.y, y.x, 0.f); b = make_float4(y.y, z.x, z.y, 0.f);
}
__device__ __forceinline__ void cbox_store(CBox* p, const float4 a, const float4 b)
{
	float2* q = reinterpret_cast<float2*>(p);
	q[0] = make_float2(a.x, a.y); q[1] = make_float2(a.z, b.x); q[2] = make_float2(b.y, b.z);
}

// the sorted leaves, and the first round's cluster list: cluster i = leaf i, its box in cbox (every round reads the clusters' boxes
// from such a list in cluster order — a stream — instead of through the node indices)
__global__ void __launch_bounds__(kTile) k_sorted_leaves(const N2* __restrict__ leaf, const uint32_t* __restrict__ order, const int64_t N,
                                                         N2* __restrict__ sorted, CBox* __restrict__ cbox, uint32_t* __restrict__ cid)
{
	const int64_t i = (int64_t)blockIdx.x * kTile + threadIdx.x;
	if (i >= N) return;
	const float4* src = reinterpret_cast<const float4*>(&leaf[order[i]]);
	const float4 a = __ldg(src), b = __ldg(src + 1);
	float4* d0 = reinterpret_cast<float4*>(&sorted[i]);
	d0[0] = a; d0[1] = b;
	cbox_store(&cbox[i], a, b);
	cid[i] = (uint32_t)i;
}

// ---- exclusive scan of per-tile 64-bit sums (two 32-bit counters packed in one word), one block -------------------
__device__ __forceinline__ uint64_t warp_scan_incl(uint64_t v)
{
	for (int d = 1; d < 32; d <<= 1) {
		const uint64_t o = __shfl_up_sync(0xffffffffu, v, d);
		if ((threadIdx.x & 31) >= d) v += o;
	}
	return v;
}
// exclusive scan over the block's kTile values; returns the exclusive prefix, *total = the block sum
__device__ __forceinline__ uint64_t block_scan_excl(const uint64_t v, uint64_t* s_warp /* kTile/32 + 1 */, uint64_t* total)
{
	const uint64_t incl = warp_scan_incl(v);
	const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
	__syncthreads();   // s_warp may still be read from a previous call
	if (l == 31) s_warp[w] = incl;
	__syncthreads();
	if (w == 0) {
		uint64_t x = l < kTile / 32 ? s_warp[l] : 0ull;
		const uint64_t xi = warp_scan_incl(x);
		if (l < kTile / 32) s_warp[l] = xi - x;
		if (l == kTile / 32 - 1) s_warp[kTile / 32] = xi;
	}
	__syncthreads();
	*total = s_warp[kTile / 32];
	return s_warp[w] + incl - v;
}

// exclusive scan of the per-tile sums by ONE block, in place.  Every thread takes kScanRun consecutive entries per sweep
// (one entry per thread and sweep was 458 dependent sweeps of three barriers each at 30 M primitives: 0.39 ms per
// clustering round, 3.7 ms per build).
constexpr int kScanRun = 16;
__global__ void __launch_bounds__(kTile) k_scan_tiles(uint64_t* __restrict__ tile_sums, const int ntiles, uint64_t* __restrict__ total_out)
{
	__shared__ uint64_t s_warp[kTile / 32 + 1];
	uint64_t carry = 0;
	for (int base = 0; base < ntiles; base += kTile * kScanRun) {
		const int first = base + (int)threadIdx.x * kScanRun;
		uint64_t v[kScanRun];
		uint64_t sum = 0;
#pragma unroll
		for (int k = 0; k < kScanRun; ++k) { v[k] = first + k < ntiles ? tile_sums[first + k] : 0ull; sum += v[k]; }
		uint64_t tot;
		uint64_t run = carry + block_scan_excl(sum, s_warp, &tot);
#pragma unroll
		for (int k = 0; k < kScanRun; ++k) {
			if (first + k < ntiles) tile_sums[first + k] = run;
			run += v[k];
		}
		carry += tot;
	}
	if (threadIdx.x == 0) *total_out = carry;
}

// ---- 3. PLOC ---------------------------------------------------------------------------------------------------
// Nearest neighbour of cluster i among positions i-R..i+R by the half-area of the union box.  Ties are broken by a
// strict total order on UNORDERED pairs — (area, |i-j|, min(i,j) odd, min(i,j)) — so the globally smallest pair is
// always mutual (progress every round) and a run of identical boxes pairs up as (0,1)(2,3)... instead of chaining.
// (k_ploc_tail's version: one block, plain loads of what it wrote itself a round earlier)
template <int kRadius>
__device__ __forceinline__ void ploc_nearest_block(const CBox* cbox, const int n, uint32_t* nn, float (*s_box)[kTile + 2 * kRadius])
{
	for (int k = threadIdx.x; k < kTile + 2 * kRadius; k += kTile) {
		const int g = k - kRadius;
		if (g >= 0 && g < n) {
			float4 a, b;
			cbox_load(&cbox[g], a, b);
			s_box[0][k] = a.x; s_box[1][k] = a.y; s_box[2][k] = a.z;
			s_box[3][k] = b.x; s_box[4][k] = b.y; s_box[5][k] = b.z;
		}
	}
	__syncthreads();
	const int i = threadIdx.x;
	if (i >= n) return;
	const int li = threadIdx.x + kRadius;
	const float lx = s_box[0][li], ly = s_box[1][li], lz = s_box[2][li], hx = s_box[3][li], hy = s_box[4][li], hz = s_box[5][li];
	float best = 3.0e38f;
	int bj = -1;
#pragma unroll
	for (int d = 1; d <= kRadius; ++d) {
		// within one distance: the candidate whose pair minimum is even first, then the lower minimum (= the left one)
		const bool right_first = (((i - d) & 1) == 1) && ((i & 1) == 0);
#pragma unroll
		for (int side = 0; side < 2; ++side) {
			const int off = ((side == 0) == right_first) ? d : -d;
			const int j = i + off;
			if (j < 0 || j >= n) continue;
			const int lj = li + off;
			const float a = half_area(fminf(lx, s_box[0][lj]), fminf(ly, s_box[1][lj]), fminf(lz, s_box[2][lj]),
			                          fmaxf(hx, s_box[3][lj]), fmaxf(hy, s_box[4][lj]), fmaxf(hz, s_box[5][lj]));
			if (a < best || bj < 0) { best = a; bj = j; }
		}
	}
	nn[i] = (uint32_t)bj;
}

// valid = the slot survives the round (everything but the higher half of a mutual pair); leader = lower half
__device__ __forceinline__ uint64_t ploc_flags(const uint32_t* nn, const int i, const int n, int* partner)
{
	if (i >= n) { *partner = -1; return 0ull; }
	const int j = (int)nn[i];
	const bool mutual = (int)nn[j] == i;
	*partner = mutual ? j : -1;
	const uint64_t valid = !(mutual && i > j), leader = mutual && i < j;
	return valid | (leader << 32);
}

// Cost-optimal collapse (Ylitie, Karras, Laine 2017, section 3.1; the host builder does the same, bvh8_build.cpp):
// c[i-1] = cheapest way to hang the subtree under a wide node using at most i of its slots (i = 1..7), splits = for
// j = 2..8 the number of slots that go to the left child, 3 bits each.  A subtree of at most max_leaf primitives is a
// leaf slot: A * P * c_prim for every i.  Filled in the moment a node is created (its children are complete subtrees by
// then), read by gather_kids top-down.
constexpr float kCostNode = 1.0f, kCostPrim = 0.3f;

// costs c[1..7] and primitive count (saturated at 7) of the subtree `id` whose box is (a, b)
__device__ __forceinline__ uint32_t subtree_costs(const Tree& t, const uint32_t id, const float4 a, const float4 b, const uint32_t max_leaf, float c[8])
{
	uint32_t n = 1u;
	float4 c0 = make_float4(0.f, 0.f, 0.f, 0.f), c1 = c0;
	if (id >= t.N) {
		const float4* p = reinterpret_cast<const float4*>(t.inner + (id - t.N));
		c0 = p[2]; c1 = p[3];
		n = (__float_as_uint(c1.w) >> 24) & 7u;
	}
	if (n <= max_leaf) {
		const float v = half_area(a.x, a.y, a.z, b.x, b.y, b.z) * (float)n * kCostPrim;
#pragma unroll
		for (int i = 1; i <= 7; ++i) c[i] = v;
	} else {
		c[1] = c0.x; c[2] = c0.y; c[3] = c0.z; c[4] = c0.w; c[5] = c1.x; c[6] = c1.y; c[7] = c1.z;
	}
	return n;
}

// the merged node `id` of clusters l and r (boxes la/lb and ra/rb = {lo, .} / {hi, .}): its 64-byte record; returns its box
__device__ __forceinline__ void ploc_create_node(const Tree& t, const uint32_t max_leaf, const uint32_t id, const uint32_t l, const uint32_t r,
                                                 const float4 la, const float4 lb, const float4 ra, const float4 rb, float4* box_lo, float4* box_hi)
{
	const float4 oa = make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), __int_as_float((int32_t)l));
	const float4 ob = make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), __int_as_float((int32_t)r));
	*box_lo = oa; *box_hi = ob;
	float cl[8], cr[8];
	const uint32_t nl = subtree_costs(t, l, la, lb, max_leaf, cl), nr = subtree_costs(t, r, ra, rb, max_leaf, cr);
	const uint32_t total = min(nl + nr, 7u);
	float c[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
	uint32_t splits = 0u;
	if (total > max_leaf) {
		float D8 = 0.f;
#pragma unroll
		for (int j = 2; j <= 8; ++j) {
			float best = 3.0e38f; uint32_t bk = 1u;
#pragma unroll
			for (int k = 1; k < j; ++k) {
				const float v = cl[k < 7 ? k : 7] + cr[(j - k) < 7 ? (j - k) : 7];
				if (v < best) { best = v; bk = (uint32_t)k; }
			}
			splits |= bk << (3 * (j - 2));
			if (j <= 7) c[j - 1] = best; else D8 = best;
		}
		c[0] = half_area(oa.x, oa.y, oa.z, ob.x, ob.y, ob.z) * kCostNode + D8;
#pragma unroll
		for (int i = 1; i < 7; ++i) c[i] = fminf(c[i], c[i - 1]);
	}
	float4* po = reinterpret_cast<float4*>(t.inner + (id - t.N));
	po[0] = oa; po[1] = ob;
	po[2] = make_float4(c[0], c[1], c[2], c[3]);
	po[3] = make_float4(c[4], c[5], c[6], __uint_as_float(splits | (total << 24)));
}

// one tile of a PLOC round (k_ploc_tail): order-preserving compaction of the surviving clusters, creation of the merged nodes and
// their collapse costs.  *tile_total = (survivors, leaders << 32) of this tile.
__device__ __forceinline__ void ploc_merge_block(const Tree& t, const uint32_t* cid_in, const CBox* cbox_in, uint32_t* cid_out, CBox* cbox_out, const uint32_t* nn,
                                                 const int n, const uint32_t node_base, const uint32_t max_leaf, uint64_t* s_warp, uint64_t* tile_total)
{
	const int i = threadIdx.x;
	int partner;
	const uint64_t f = ploc_flags(nn, i, n, &partner);
	uint64_t tot;
	const uint64_t pos = block_scan_excl(f, s_warp, &tot);
	*tile_total = tot;
	if (!(f & 1ull)) return;
	const uint32_t out = (uint32_t)pos;
	float4 a, b;
	cbox_load(&cbox_in[i], a, b);
	uint32_t id = cid_in[i];
	if (f >> 32) {
		float4 ra, rb;
		cbox_load(&cbox_in[partner], ra, rb);
		const uint32_t l = id, r = cid_in[partner];
		id = node_base + (uint32_t)(pos >> 32);
		ploc_create_node(t, max_leaf, id, l, r, a, b, ra, rb, &a, &b);
	}
	cid_out[out] = id;
	cbox_store(&cbox_out[out], a, b);
}

// ---- one clustering round = ONE launch (round 2) -------------------------------------------------------------------------------
// Round 1 ran four kernels per round (nearest, count, scan of the tile sums by one block, merge) and read the survivor count back
// after every round.  Now a block does the whole round for its tile:
//   * it loads the boxes of its kTile clusters and of 2R neighbours on either side into shared memory ONCE (the separate kernels
//     read every box three times: nearest, and both halves of a pair in the merge), finds the nearest neighbour of its clusters
//     and of R on either side (a partner lies within R, and mutual needs the partner's choice too) — the neighbour array never
//     leaves the block;
//   * the tile's position in the compacted output comes from a decoupled look-back over the per-tile (survivors, leaders) sums
//     (Merrill & Garland, "Single-pass Parallel Prefix Scan with Decoupled Look-back", 2016) instead of a count kernel and a
//     one-block scan: tiles are handed out by a ticket counter so that every lower tile is already running, a tile publishes
//     its sum, then adds up its predecessors' sums until it meets one that carries an inclusive prefix.  Sums are exact
//     integers, so node numbering is as deterministic as before;
//   * the state of the round (cluster count, next node index, rounds so far = ping-pong parity) lives on the device in
//     state[launch & 1]; the block of the LAST tile writes state[(launch + 1) & 1], so the host can queue several rounds back to
//     back with grids sized for the count it last saw and read the state once per batch.  A launch that finds at most kTile
//     clusters left only copies the state forward (k_ploc_tail finishes the build).
struct PlocState { int32_t n; uint32_t node_base; uint32_t rounds; uint32_t stalled; };
static_assert(sizeof(PlocState) == 16, "PlocState travels through the arena's 16-byte pinned word");
constexpr int kMaxRoundLaunches = 4096;   // ticket counters, one per launch (51 rounds at 100 M primitives)

// look-back word: flag (2 bits: 1 = this tile's sum, 2 = inclusive prefix) | launch tag (6 bits) | leaders (28) | survivors (28)
__device__ __forceinline__ uint64_t lb_word(const uint64_t flag, const uint32_t launch, const uint64_t sums /* survivors | leaders << 32 */)
{
	return (flag << 62) | ((uint64_t)(launch & 63u) << 56) | ((sums >> 32) << 28) | (sums & 0xfffffffull);
}
__device__ __forceinline__ uint64_t lb_sums(const uint64_t w) { return (w & 0xfffffffull) | (((w >> 28) & 0xfffffffull) << 32); }

// kRoundItems clusters per thread: large tiles keep the look-back short at 10^7+ clusters (with 256-cluster tiles the blocks arrive
// faster than a look-back of 32 tiles per L2 round trip can follow: 30.7 ms of clustering at 100 M primitives against 19.8 with
// 1024-cluster tiles), small tiles keep all SMs busy at 10^5 (0.47 against 0.54 ms at 262 k)
#ifndef CUPT_PLOC_ITEMS
#define CUPT_PLOC_ITEMS 4
#endif
template <int kRadius, int kRoundItems>
__global__ void __launch_bounds__(kTile, 6) k_ploc_round(const Tree t, uint32_t* cid0, uint32_t* cid1, CBox* cbox0, CBox* cbox1, PlocState* state, const uint32_t launch,
                                                      uint32_t* tickets, uint64_t* tile_state, const uint32_t max_leaf)
{
	constexpr int kRoundTile = kTile * kRoundItems;
	constexpr int W = kRoundTile + 4 * kRadius;
	__shared__ float s_box[6][W];
	__shared__ uint32_t s_cid[W];
	__shared__ int8_t s_nn[kRoundTile + 2 * kRadius];            // neighbour's position minus the cluster's own (within +-R)
	__shared__ __align__(4) uint8_t s_flag[kRoundTile];     // bit 0: survives the round, bit 4: leader (lower half of a mutual pair)
	__shared__ uint64_t s_pref[kRoundTile / 4];                    // exclusive (survivors | leaders << 32) before every group of four positions
	__shared__ uint64_t s_warp[kTile / 32 + 1];
	__shared__ uint32_t s_tile;
	__shared__ uint64_t s_off;
	const PlocState cur = state[launch & 1u];
	const int n = cur.n;
	if (n <= kTile) {
		if (blockIdx.x == 0 && threadIdx.x == 0) state[(launch + 1u) & 1u] = cur;
		return;
	}
	if (threadIdx.x == 0) s_tile = atomicAdd(&tickets[launch], 1u);
	__syncthreads();
	const int tile = (int)s_tile;
	const int ntiles = (n + kRoundTile - 1) / kRoundTile;
	if (tile >= ntiles) return;
	const uint32_t* cin = (cur.rounds & 1u) ? cid1 : cid0;
	uint32_t* cout = (cur.rounds & 1u) ? cid0 : cid1;
	const CBox* cbin = (cur.rounds & 1u) ? cbox1 : cbox0;
	CBox* cbout = (cur.rounds & 1u) ? cbox0 : cbox1;
	// boxes of positions [tile * kRoundTile - 2R, (tile + 1) * kRoundTile + 2R)
	const int base = tile * kRoundTile - 2 * kRadius;
	for (int k = threadIdx.x; k < W; k += kTile) {
		const int g = base + k;
		if (g >= 0 && g < n) {
			float4 a, b;
			cbox_load(&cbin[g], a, b);
			s_cid[k] = cin[g];
			s_box[0][k] = a.x; s_box[1][k] = a.y; s_box[2][k] = a.z;
			s_box[3][k] = b.x; s_box[4][k] = b.y; s_box[5][k] = b.z;
		}
	}
	__syncthreads();
	// nearest neighbour of positions [tile * kRoundTile - R, (tile + 1) * kRoundTile + R): same rule as ploc_nearest_block
	const int nn0 = tile * kRoundTile - kRadius;   // position of s_nn[0]
	for (int k2 = threadIdx.x; k2 < kRoundTile + 2 * kRadius; k2 += kTile) {
		const int i = nn0 + k2;
		int bj = i;
		bool found = false;
		if (i >= 0 && i < n) {
			const int li = k2 + kRadius;
			const float lx = s_box[0][li], ly = s_box[1][li], lz = s_box[2][li], hx = s_box[3][li], hy = s_box[4][li], hz = s_box[5][li];
			float best = 3.0e38f;
#pragma unroll
			for (int d = 1; d <= kRadius; ++d) {
				const bool right_first = (((i - d) & 1) == 1) && ((i & 1) == 0);
#pragma unroll
				for (int side = 0; side < 2; ++side) {
					const int off = ((side == 0) == right_first) ? d : -d;
					const int j = i + off;
					if (j < 0 || j >= n) continue;
					const int lj = li + off;
					const float a = half_area(fminf(lx, s_box[0][lj]), fminf(ly, s_box[1][lj]), fminf(lz, s_box[2][lj]),
					                          fmaxf(hx, s_box[3][lj]), fmaxf(hy, s_box[4][lj]), fmaxf(hz, s_box[5][lj]));
					if (a < best || !found) { best = a; bj = j; found = true; }
				}
			}
		}
		s_nn[k2] = (int8_t)(bj - i);
	}
	__syncthreads();
	// survivors and leaders: thread t owns positions t, t + kTile, ... of the tile
	int partner[kRoundItems];
#pragma unroll
	for (int k = 0; k < kRoundItems; ++k) {
		const int p = k * kTile + (int)threadIdx.x;
		const int i = tile * kRoundTile + p;
		partner[k] = -1;
		uint32_t fl = 0u;
		if (i < n) {
			const int j = i + (int)s_nn[i - nn0];
			const bool mutual = j + (int)s_nn[j - nn0] == i;
			if (mutual) partner[k] = j;
			fl = (uint32_t)(!(mutual && i > j)) | ((uint32_t)(mutual && i < j) << 4);
		}
		s_flag[p] = (uint8_t)fl;
	}
	__syncthreads();
	const uint32_t fw = threadIdx.x < kRoundTile / 4 ? reinterpret_cast<const uint32_t*>(s_flag)[threadIdx.x] : 0u;   // positions 4 t .. 4 t + 3
	uint64_t tot;
	const uint64_t ex = block_scan_excl((uint64_t)__popc(fw & 0x01010101u) | ((uint64_t)__popc(fw & 0x10101010u) << 32), s_warp, &tot);
	if (threadIdx.x < kRoundTile / 4) s_pref[threadIdx.x] = ex;
	// decoupled look-back (warp 0): exclusive prefix of the tile sums
	if (threadIdx.x < 32) {
		volatile uint64_t* ts = tile_state;
		uint64_t excl = 0ull;
		if (tile == 0) {
			if (threadIdx.x == 0) ts[0] = lb_word(2ull, launch, tot);
		} else {
			if (threadIdx.x == 0) ts[tile] = lb_word(1ull, launch, tot);
			int look = tile - 1;
			for (;;) {
				const int t = look - (int)threadIdx.x;
				uint64_t w = 0ull;
				bool prefix = false;
				if (t >= 0) {
					do { w = ts[t]; } while (((w >> 56) & 63ull) != (uint64_t)(launch & 63u) || (w >> 62) == 0ull);
					prefix = (w >> 62) == 2ull;
				}
				const uint32_t pm = __ballot_sync(0xffffffffu, prefix);
				const int first = pm ? __ffs(pm) - 1 : 32;                 // nearest predecessor that carries an inclusive prefix
				uint64_t v = (t >= 0 && (int)threadIdx.x <= first) ? lb_sums(w) : 0ull;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
				excl += v;
				if (pm != 0u || look < 32) break;
				look -= 32;
			}
			if (threadIdx.x == 0) ts[tile] = lb_word(2ull, launch, excl + tot);
		}
		if (threadIdx.x == 0) {
			s_off = excl;
			if (tile == ntiles - 1) {   // the round's totals: the state of the next launch
				const uint64_t all = excl + tot;
				const uint32_t valid = (uint32_t)all, leaders = (uint32_t)(all >> 32);
				PlocState nx = cur;
				if (leaders == 0u || (int)(valid + leaders) != n) { nx.stalled = 1u; nx.n = 0; }   // (cannot happen: the smallest pair is always mutual)
				else { nx.n = (int32_t)valid; nx.node_base = cur.node_base + leaders; nx.rounds = cur.rounds + 1u; }
				state[(launch + 1u) & 1u] = nx;
			}
		}
	}
	__syncthreads();
	const uint64_t off = s_off;
#pragma unroll
	for (int k = 0; k < kRoundItems; ++k) {
		const int p = k * kTile + (int)threadIdx.x;
		const uint32_t fl = s_flag[p];
		if (!(fl & 1u)) continue;
		const uint32_t w4 = reinterpret_cast<const uint32_t*>(s_flag)[p >> 2];
		const uint32_t below = (1u << (8 * (p & 3))) - 1u;
		const uint64_t pos = off + s_pref[p >> 2] + ((uint64_t)__popc(w4 & 0x01010101u & below) | ((uint64_t)__popc(w4 & 0x10101010u & below) << 32));
		const uint32_t out = (uint32_t)pos;
		const int li = p + 2 * kRadius;
		float4 a = make_float4(s_box[0][li], s_box[1][li], s_box[2][li], 0.f), b = make_float4(s_box[3][li], s_box[4][li], s_box[5][li], 0.f);
		uint32_t id = s_cid[li];
		if (fl & 16u) {
			const int lp = li + (partner[k] - (tile * kRoundTile + p));
			const uint32_t l = id, r = s_cid[lp];
			id = cur.node_base + (uint32_t)(pos >> 32);
			ploc_create_node(t, max_leaf, id, l, r, a, b, make_float4(s_box[0][lp], s_box[1][lp], s_box[2][lp], 0.f),
			                 make_float4(s_box[3][lp], s_box[4][lp], s_box[5][lp], 0.f), &a, &b);
		}
		cout[out] = id;
		cbox_store(&cbout[out], a, b);
	}
}

// The last rounds in ONE launch: once at most kTile clusters are left, a single block runs nearest / merge round after round
// with block barriers in between (no launches, no host read-back): about log_1.5(256) = 14 of the ~30 rounds of a 262 k-triangle
// build.  state[0] = node_base afterwards, state[1] = rounds run here, state[2] = which cid buffer holds the root.
template <int kRadius>
__global__ void __launch_bounds__(kTile) k_ploc_tail(const Tree t, uint32_t* cid0, uint32_t* cid1, CBox* cbox0, CBox* cbox1, uint32_t* nn, int n, uint32_t node_base,
                                                     const uint32_t max_leaf, int cur, uint32_t* state)
{
	__shared__ float s_box[6][kTile + 2 * kRadius];
	__shared__ uint64_t s_warp[kTile / 32 + 1];
	uint32_t rounds = 0u;
	while (n > 1) {
		uint32_t* cin = cur ? cid1 : cid0;
		uint32_t* cout = cur ? cid0 : cid1;
		CBox* cbin = cur ? cbox1 : cbox0;
		CBox* cbout = cur ? cbox0 : cbox1;
		ploc_nearest_block<kRadius>(cbin, n, nn, s_box);
		__syncthreads();
		uint64_t tot;
		ploc_merge_block(t, cin, cbin, cout, cbout, nn, n, node_base, max_leaf, s_warp, &tot);
		__syncthreads();
		n = (int)(uint32_t)tot; node_base += (uint32_t)(tot >> 32); cur ^= 1; ++rounds;
		if ((tot >> 32) == 0ull) break;   // (no mutual pair: cannot happen, see ploc_nearest_block; the host checks n)
	}
	if (threadIdx.x == 0) { state[0] = node_base; state[1] = rounds; state[2] = (uint32_t)cur; state[3] = (uint32_t)n; }
}

// ---- 4. collapse to 8-wide + emission ----------------------------------------------------------------------------
struct Kids { int32_t id[8]; uint32_t prims[8]; int n; };   // prims: primitives of a leaf slot, 0 for a child that becomes a wide node

// the children of one wide node: the cost-optimal cut below the binary node `root`; a subtree with at most max_leaf primitives
// is a leaf slot.  Children come out in left-first order.  One 64-byte record per inner node visited.
__device__ __forceinline__ void gather_kids(const Tree& t, const int32_t root, const uint32_t max_leaf, Kids& k)
{
	k.n = 0;
	if ((uint32_t)root < t.N) { k.id[0] = root; k.prims[0] = 1u; k.n = 1; return; }
	int32_t st_n[16]; int st_b[16]; int sp = 0;
	{
		const float4* p = reinterpret_cast<const float4*>(t.inner + ((uint32_t)root - t.N));
		const float4 a = __ldg(p), b = __ldg(p + 1), d = __ldg(p + 3);
		const uint32_t sw = __float_as_uint(d.w), cn = (sw >> 24) & 7u;
		if (cn <= max_leaf) { k.id[0] = root; k.prims[0] = cn; k.n = 1; return; }
		const int k8 = (int)((sw >> 18) & 7u);
		st_n[sp] = __float_as_int(b.w); st_b[sp++] = 8 - k8;
		st_n[sp] = __float_as_int(a.w); st_b[sp++] = k8;
	}
	while (sp > 0) {
		--sp;
		const int32_t n = st_n[sp];
		int i = min(st_b[sp], 7);
		if ((uint32_t)n < t.N) { k.prims[k.n] = 1u; k.id[k.n++] = n; continue; }
		const float4* p = reinterpret_cast<const float4*>(t.inner + ((uint32_t)n - t.N));
		const float4 a = __ldg(p), b = __ldg(p + 1), c0 = __ldg(p + 2), c1 = __ldg(p + 3);
		const uint32_t sw = __float_as_uint(c1.w), cn = (sw >> 24) & 7u;
		if (cn <= max_leaf) { k.prims[k.n] = cn; k.id[k.n++] = n; continue; }
		const float c[8] = {0.f, c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z};
		while (i > 1 && c[i] == c[i - 1]) --i;
		if (i == 1) { k.prims[k.n] = 0u; k.id[k.n++] = n; continue; }   // a wide node of its own
		const int kl = (int)((sw >> (3 * (i - 2))) & 7u);
		st_n[sp] = __float_as_int(b.w); st_b[sp++] = i - kl;
		st_n[sp] = __float_as_int(a.w); st_b[sp++] = kl;
	}
}

// per item of the level: (inner children | leaf primitives << 32), reduced to one sum per tile.  The item's children (the
// cost-optimal cut, a serial walk of up to 15 dependent node reads) are kept for k_collapse_emit, and so is the item's
// exclusive prefix inside its tile: the emission then needs neither the walk nor a block-wide scan.
__global__ void __launch_bounds__(kTile) k_collapse_count(const Tree t, const int32_t* __restrict__ items,
                                                          const int n_items, const uint32_t max_leaf, uint64_t* __restrict__ item_prefix, int32_t* __restrict__ item_kids,
                                                          uint64_t* __restrict__ tile_sums)
{
	__shared__ uint64_t s_warp[kTile / 32 + 1];
	const int i = (int)blockIdx.x * kTile + threadIdx.x;
	uint64_t f = 0;
	if (i < n_items) {
		Kids k;
		gather_kids(t, items[i], max_leaf, k);
		uint32_t ni = 0, np = 0;
		int32_t ids[8];
#pragma unroll
		for (int q = 0; q < 8; ++q) {
			ids[q] = q < k.n ? k.id[q] : -1;
			if (q < k.n) {
				if (k.prims[q] == 0u) ni++; else np += k.prims[q];
			}
		}
		f = (uint64_t)ni | ((uint64_t)np << 32);
		int4* dst = reinterpret_cast<int4*>(item_kids + 8 * (size_t)i);
		dst[0] = make_int4(ids[0], ids[1], ids[2], ids[3]);
		dst[1] = make_int4(ids[4], ids[5], ids[6], ids[7]);
	}
	uint64_t tot;
	const uint64_t ex = block_scan_excl(f, s_warp, &tot);
	if (i < n_items) item_prefix[i] = ex;
	if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__device__ __forceinline__ double pow2_double(const int e) { return __hiloint2double((1023 + e) << 20, 0); }   // 2^e, -1022 <= e <= 1023
__device__ __forceinline__ int dev_float_exponent(const float v) { return (int)((__float_as_uint(v) >> 23) & 0xffu) - 127; }

struct EmitArgs {
	Tree t; const int32_t* items; int n_items; uint32_t max_leaf;
	const uint64_t* item_prefix; const int32_t* item_kids; const uint64_t* tile_offs;
	uint32_t level_base, next_base, prim_base;
	int32_t* next_items; Node8* out_nodes; float4* out_prims;
	int32_t* rec_prim;   // primitive of every record, in record order (k_emit_prims)
	double inv_root_area; double* sah;
};

// The 48-byte records in tree order, one thread per record: one 48-byte piece from the primitive-order array k_leaf_boxes wrote.
__global__ void __launch_bounds__(kTile) k_emit_prims(const int32_t* __restrict__ rec_prim, const int64_t N, const float4* __restrict__ records,
                                                      float4* __restrict__ out_prims)
{
	const int64_t r = (int64_t)blockIdx.x * kTile + threadIdx.x;
	if (r >= N) return;
	const float4* src = records + 3 * (size_t)rec_prim[r];
	const float4 r0 = __ldg(src), r1 = __ldg(src + 1), r2 = __ldg(src + 2);
	float4* dst = out_prims + 3 * (size_t)r;
	dst[0] = r0; dst[1] = r1; dst[2] = r2;
}

// One wide node per EIGHT lanes, lane q of the group = child q of the cut (round 1: one thread per wide node with the 8 x 8
// slot assignment, the boxes and the node image in local memory — 23 of the 57 ms of a 100 M-triangle build).  The serial
// rules of the host builder (bvh8_build.cpp) are kept exactly: greedy slot assignment = eight rounds of "largest cost over
// (unassigned child, free slot), first in (child, slot) order wins"; children's inner-node and record offsets in SLOT order.
__global__ void __launch_bounds__(kTile) k_collapse_emit(const EmitArgs A)
{
	__shared__ __align__(16) uint8_t s_q[kTile / 8][48];   // per group: qlo[3][8], qhi[3][8] of the node being assembled
	__shared__ double s_sah[kTile / 32];
	const uint32_t lane = threadIdx.x & 31u, q = lane & 7u, grp = threadIdx.x >> 3;
	const uint32_t gmask = 0xffu << (lane & 24u);     // the eight lanes of this group
	const int i = (int)blockIdx.x * (kTile / 8) + (int)grp;
	double sah = 0;
	if (i < A.n_items) {   // (uniform over the group)
		const int32_t self = A.items[i];
		const int32_t kid = A.item_kids[8 * (size_t)i + q];
		const uint64_t ex = A.tile_offs[i / kTile] + A.item_prefix[i];
		float blo[3], bhi[3], clo[3] = {0.f, 0.f, 0.f}, chi[3] = {0.f, 0.f, 0.f};
		{
			const float4* p = A.t.rec((uint32_t)self);
			const float4 a = __ldg(p), b = __ldg(p + 1);
			blo[0] = a.x; blo[1] = a.y; blo[2] = a.z; bhi[0] = b.x; bhi[1] = b.y; bhi[2] = b.z;
		}
		float cost[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
		if (kid >= 0) {
			const float4* p = A.t.rec((uint32_t)kid);
			const float4 a = __ldg(p), b = __ldg(p + 1);
			clo[0] = a.x; clo[1] = a.y; clo[2] = a.z; chi[0] = b.x; chi[1] = b.y; chi[2] = b.z;
			float d[3];
#pragma unroll
			for (int a3 = 0; a3 < 3; ++a3) d[a3] = (0.5f * clo[a3] + 0.5f * chi[a3]) - (0.5f * blo[a3] + 0.5f * bhi[a3]);
#pragma unroll
			for (int sl = 0; sl < 8; ++sl) cost[sl] = (((sl & 4) ? d[0] : -d[0]) + ((sl & 2) ? d[1] : -d[1]) + ((sl & 1) ? d[2] : -d[2])) + 0.0f;   // (+ 0: no -0, so that
		}                                                                                                                                   //  the integer order below is the float order)
		// greedy octant slots
		const uint32_t nk = (uint32_t)__popc(__ballot_sync(gmask, kid >= 0) & gmask);
		// Per round: every unassigned child's best free slot (largest cost, lowest slot on a tie; a taken slot's cost is -inf in every
		// lane), then the group's winner — largest cost, lowest child on a tie: one integer max-reduction over the order-preserving
		// image of the cost and a ballot — takes its slot.  (The first version of this kernel spent 47 % of its instructions here on a
		// slot-by-slot scan with a free mask and a three-value butterfly: profiles/r2w_builder_metrics.md.)
		int slot = -1;
		for (uint32_t it = 0; it < nk; ++it) {
			const bool cand = kid >= 0 && slot < 0;
			const float m = fmaxf(fmaxf(fmaxf(cost[0], cost[1]), fmaxf(cost[2], cost[3])), fmaxf(fmaxf(cost[4], cost[5]), fmaxf(cost[6], cost[7])));
			int bs = 7;
#pragma unroll
			for (int sl = 6; sl >= 0; --sl)
				if (cost[sl] == m) bs = sl;
			const uint32_t key = cand ? f2ord(m) : 0u;   // (f2ord of a finite float is never 0)
			const uint32_t top = __reduce_max_sync(gmask, key);
			const uint32_t winners = __ballot_sync(gmask, cand && key == top) & gmask;
			const int wl = __ffs(winners) - 1;             // lowest lane = lowest child index
			const int ws = __shfl_sync(gmask, bs, wl);
			if ((int)lane == wl) slot = ws;
#pragma unroll
			for (int sl = 0; sl < 8; ++sl)
				if (sl == ws) cost[sl] = -INFINITY;
		}
		// quantisation grid: power-of-two cell per axis, at least 2^-24 of the coordinate magnitude (every lane: the same values)
		double scale[3], inv_scale[3];
		uint32_t ebits = 0u;
#pragma unroll
		for (int a = 0; a < 3; ++a) {
			const double ext = (double)bhi[a] - (double)blo[a];
			int e = -126;
			if (ext > 0) {
				e = ((__double2hiint(ext) >> 20) & 0x7ff) - 1023 - 8;   // 2^k <= ext < 2^(k+1): 255 * 2^(k-8) < ext, at most two steps up
				while (255.0 * pow2_double(e) < ext) ++e;
			}
			const float m = fmaxf(fabsf(blo[a]), fabsf(bhi[a]));
			e = max(e, dev_float_exponent(m) - 24);
			e = max(-126, min(127, e));
			ebits |= (uint32_t)(e + 127) << (8 * a);
			scale[a] = pow2_double(e);
			inv_scale[a] = pow2_double(-e);
		}
		// this child's record count / kind, and what lies before it in SLOT order
		const uint32_t cc = kid >= 0 ? A.t.count((uint32_t)kid) : 0u;
		const bool inner = kid >= 0 && cc > A.max_leaf, leaf = kid >= 0 && !inner;
		uint32_t inner_before = 0u, prims_before = 0u;
#pragma unroll
		for (int j = 0; j < 8; ++j) {
			const int os = __shfl_sync(gmask, slot, (int)(lane & 24u) + j);
			const uint32_t occ = __shfl_sync(gmask, cc, (int)(lane & 24u) + j);
			const int oin = __shfl_sync(gmask, (int)inner, (int)(lane & 24u) + j);
			if (os >= 0 && slot >= 0 && os < slot) { if (oin) ++inner_before; else prims_before += occ; }
		}
		const uint32_t child_off = (uint32_t)ex, prim_off = (uint32_t)(ex >> 32);
		const uint32_t child_base = A.next_base + child_off, tri_base = A.prim_base + prim_off;
		for (uint32_t k = q; k < 48u; k += 8u) s_q[grp][k] = 0u;
		__syncwarp(gmask);
		uint32_t my_imask = 0u, my_pvalid = 0u;
		if (kid >= 0) {
#pragma unroll
			for (int a = 0; a < 3; ++a) {
				const double p = (double)blo[a];
				double ql = floor(((double)clo[a] - p) * inv_scale[a]);
				double qh = ceil(((double)chi[a] - p) * inv_scale[a]);
				while (ql > 0 && p + ql * scale[a] > (double)clo[a]) ql -= 1;
				while (p + qh * scale[a] < (double)chi[a]) qh += 1;
				ql = fmax(0.0, fmin(255.0, ql));
				qh = fmin(255.0, fmax(qh, ql));   // hi <= node hi <= p + 255 * scale by construction of e
				s_q[grp][8 * a + slot] = (uint8_t)ql; s_q[grp][24 + 8 * a + slot] = (uint8_t)qh;
			}
			const double area2 = 2.0 * (double)half_area(clo[0], clo[1], clo[2], chi[0], chi[1], chi[2]);
			if (inner) {
				my_imask = 1u << slot;
				A.next_items[child_off + inner_before] = kid;
				sah += area2 * A.inv_root_area;
			} else {
				my_pvalid = (cc == 1 ? 1u : (cc == 2 ? 3u : 7u)) << (4 * slot);
				// <= 3 primitives: the tiny subtree is at most two levels deep; left first (the records themselves are written by k_emit_prims)
				int32_t* dst = A.rec_prim + tri_base + prims_before;
				const int32_t l0 = A.t.left((uint32_t)kid);
				if (l0 < 0) {
					dst[0] = -1 - l0;
				} else {
					const int32_t r0 = A.t.right((uint32_t)kid);
					uint32_t w = 0;
					const int32_t ll = A.t.left((uint32_t)l0);
					if (ll < 0) dst[w++] = -1 - ll;
					else { dst[w++] = -1 - A.t.leaf[ll].left; dst[w++] = -1 - A.t.leaf[A.t.right((uint32_t)l0)].left; }
					const int32_t rl = A.t.left((uint32_t)r0);
					if (rl < 0) dst[w++] = -1 - rl;
					else { dst[w++] = -1 - A.t.leaf[rl].left; dst[w++] = -1 - A.t.leaf[A.t.right((uint32_t)r0)].left; }
				}
				sah += area2 * A.inv_root_area * cc;
			}
		}
		(void)leaf;
		const uint32_t imask = __reduce_or_sync(gmask, my_imask), pvalid = __reduce_or_sync(gmask, my_pvalid);
		__syncwarp(gmask);
		if (q < 5u) {   // the 80-byte node: five 16-byte words, one per lane
			uint4 w;
			if (q == 0u) w = make_uint4(__float_as_uint(blo[0]), __float_as_uint(blo[1]), __float_as_uint(blo[2]), ebits | (imask << 24));
			else if (q == 1u) w = make_uint4(child_base, 0u, tri_base, pvalid);
			else w = *reinterpret_cast<const uint4*>(&s_q[grp][16u * (q - 2u)]);
			reinterpret_cast<uint4*>(&A.out_nodes[A.level_base + i])[q] = w;
		}
	}
	// one atomic per block: one per thread on the single address serialised the kernel (2.9 M wide nodes at 30 M triangles)
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) sah += __shfl_down_sync(0xffffffffu, sah, o);
	if ((threadIdx.x & 31) == 0) s_sah[threadIdx.x >> 5] = sah;
	__syncthreads();
	if (threadIdx.x == 0) {
		double t = 0;
		for (int w = 0; w < kTile / 32; ++w) t += s_sah[w];
		atomicAdd(A.sah, t);
	}
}

}  // namespace devbuild

using namespace devbuild;

// ---- workspace arena: one allocation that only grows, kept by the context between builds --------------------------
// (cudaMalloc / cudaFree per build cost more than the build itself and cudaFree synchronises the device)
int BuildArena::reserve(size_t bytes)
{
	if (!h_total) {
		if (cudaMallocHost((void**)&h_total, 2 * sizeof(uint64_t)) != cudaSuccess) return set_error(CUPT_ERR_CUDA, "cudaMallocHost failed");
		for (auto& e : ev)
			if (cudaEventCreate(&e) != cudaSuccess) return set_error(CUPT_ERR_CUDA, "cudaEventCreate failed");
	}
	off = 0;
	if (bytes <= cap) return CUPT_OK;
	if (base) { cudaFree(base); base = nullptr; cap = 0; }
	const cudaError_t e = cudaMalloc(&base, bytes);
	if (e != cudaSuccess) return set_error(CUPT_ERR_NOMEM, "BVH build workspace of %zu bytes: %s", bytes, cudaGetErrorString(e));
	cap = bytes;
	return CUPT_OK;
}

void BuildArena::destroy()
{
	if (base) cudaFree(base);
	if (h_total) { cudaFreeHost(h_total); for (auto& e : ev) cudaEventDestroy(e); }
	base = nullptr; h_total = nullptr; cap = off = 0;
}

namespace {
inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }
template <class T>
T* take(BuildArena& a, size_t n)
{
	T* p = reinterpret_cast<T*>(static_cast<char*>(a.base) + a.off);
	a.off += up256(std::max<size_t>(n, 1) * sizeof(T));
	return p;
}
}  // namespace

#define CKB(call)                                                                                          \
	do {                                                                                                   \
		cudaError_t e_ = (call);                                                                           \
		if (e_ != cudaSuccess)                                                                             \
			return set_error(CUPT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
	} while (0)

int build_wide_bvh_device(cudaStream_t st, BuildArena& arena, const float* verts, const cupt_triangle* tris, int64_t nt, const float4* spheres, int64_t ns,
                          const uint32_t* rank, int max_leaf, int radius, DeviceBvh* out)
{
	const int64_t N = nt + ns;
	if (N < 1) return set_error(CUPT_ERR_INVALID, "scene has no primitives");
	if (N > (1ll << 27)) return set_error(CUPT_ERR_LIMIT, "more than 2^27 primitives");
	max_leaf = std::max(1, std::min(3, max_leaf));
	const int tilesN = (int)((N + kTile - 1) / kTile);
	const size_t n = (size_t)N;

	const size_t temp_bytes = rsort::scratch_bytes((uint32_t)N);
	// sort on the top log2(N) + 9 bits of the code, rounded up to whole 8-bit passes (radix_sort.cuh)
	int log2n = 0;
	while ((1ll << log2n) < N) ++log2n;
	const int sort_passes = std::min(8, (log2n + 9 + 7) / 8);
	// everything the build touches except its two outputs, in one arena; the sort buffers are dead once the sorted
	// leaves are written, the collapse reuses the cluster arrays for its level queues
	const size_t need = up256(n * sizeof(N2)) + up256(n * 48) + up256(64) + 2 * up256(n * 8) + 2 * up256(n * 4) + up256(temp_bytes + 256)
	                    + up256(n * sizeof(N2)) + up256(n * sizeof(Inner)) + 2 * up256(n * sizeof(CBox)) + 3 * up256(n * 4) + up256(((size_t)tilesN + 1) * 8) + up256(64)
	                    + up256((n / 2 + 2) * 8) + up256((n / 2 + 2) * 32) + up256(64) + up256(kMaxRoundLaunches * 4) + 4096;
	int rc;
	if ((rc = arena.reserve(need))) return rc;
	cudaEvent_t* ev = arena.ev;
	uint64_t* h_total = arena.h_total;

	// 1-2: boxes, Morton codes, sort
	N2* leaf = take<N2>(arena, n);
	float4* records = take<float4>(arena, 3 * n);
	uint32_t* cbounds = take<uint32_t>(arena, 6);
	uint64_t* keys[2]; uint32_t* vals[2];
	for (int b = 0; b < 2; ++b) keys[b] = take<uint64_t>(arena, n);
	for (int b = 0; b < 2; ++b) vals[b] = take<uint32_t>(arena, n);
	uint8_t* temp = take<uint8_t>(arena, temp_bytes + 256);
	static const uint32_t cb_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
	CKB(cudaEventRecord(ev[0], st));
	CKB(cudaMemcpyAsync(cbounds, cb_init, sizeof(cb_init), cudaMemcpyHostToDevice, st));
	k_leaf_boxes<<<std::min(tilesN, 148 * 16), kTile, 0, st>>>(verts, tris, nt, spheres, rank, N, leaf, records, cbounds);
	k_morton<<<tilesN, kTile, 0, st>>>(leaf, N, cbounds, keys[0], vals[0]);
	const int sorted = rsort::sort_pairs(st, keys, vals, (uint32_t)N, sort_passes, reinterpret_cast<uint32_t*>(temp));
	CKB(cudaEventRecord(ev[1], st));

	// 3: PLOC
	Tree tree;
	tree.leaf = take<N2>(arena, n); tree.inner = take<Inner>(arena, n); tree.N = (uint32_t)N;
	CBox* cbox[2] = {take<CBox>(arena, n), take<CBox>(arena, n)};
	uint32_t* cid[2] = {take<uint32_t>(arena, n), take<uint32_t>(arena, n)};
	uint32_t* nn = take<uint32_t>(arena, n);
	uint64_t* tile_sums = take<uint64_t>(arena, (size_t)tilesN + 1);
	uint64_t* d_total = take<uint64_t>(arena, 1);
	uint64_t* item_counts = take<uint64_t>(arena, n / 2 + 2);
	int32_t* item_kids = take<int32_t>(arena, (n / 2 + 2) * 8);
	double* d_sah = take<double>(arena, 1);
	uint32_t* tickets = take<uint32_t>(arena, kMaxRoundLaunches);
	if (arena.off > arena.cap) return set_error(CUPT_ERR_CUDA, "internal: BVH build workspace undersized (%zu > %zu)", arena.off, arena.cap);
	k_sorted_leaves<<<tilesN, kTile, 0, st>>>(leaf, vals[sorted], N, tree.leaf, cbox[0], cid[0]);
	int cn = (int)N, cur = 0, rounds = 0;
	uint32_t node_base = (uint32_t)N;
	if (cn > kTile) {
		PlocState* d_state = reinterpret_cast<PlocState*>(item_counts);   // state[2] (the collapse's per-item counts are not in use yet)
		PlocState* h_state = reinterpret_cast<PlocState*>(h_total);
		h_state->n = cn; h_state->node_base = node_base; h_state->rounds = 0u; h_state->stalled = 0u;
		CKB(cudaMemcpyAsync(d_state, h_state, sizeof(PlocState), cudaMemcpyHostToDevice, st));
		CKB(cudaMemsetAsync(tickets, 0, kMaxRoundLaunches * sizeof(uint32_t), st));
		CKB(cudaMemsetAsync(tile_sums, 0, ((size_t)tilesN + 1) * sizeof(uint64_t), st));
		int bound = cn;
		uint32_t launch = 0;
		while (bound > kTile) {
			const int tiles = (bound + kTile - 1) / kTile;
			// rounds per read-back: clusters shrink by a quarter to a half per round, so a batch's later rounds launch up to ~3 x the
			// blocks they need; cheap when the grids are small (where the synchronisation is what costs), not at 10^5 tiles
			const int batch = tiles >= 32768 ? 1 : (tiles >= 4096 ? 2 : (tiles >= 512 ? 4 : 8));
			if (launch + (uint32_t)batch > (uint32_t)kMaxRoundLaunches) return set_error(CUPT_ERR_LIMIT, "PLOC needs more than %d rounds", kMaxRoundLaunches);
			const int items = bound >= (1 << 20) ? CUPT_PLOC_ITEMS : 1;
			const int grid = (bound + kTile * items - 1) / (kTile * items);
			for (int b = 0; b < batch; ++b, ++launch) {
#define CUPT_ROUND(R)                                                                                                                                        \
	if (items == CUPT_PLOC_ITEMS) k_ploc_round<R, CUPT_PLOC_ITEMS><<<grid, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], d_state, launch, tickets, tile_sums, (uint32_t)max_leaf); \
	else k_ploc_round<R, 1><<<grid, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], d_state, launch, tickets, tile_sums, (uint32_t)max_leaf);
				switch (radius) {
				case 32: CUPT_ROUND(32) break;
				case 16: CUPT_ROUND(16) break;
				case 8: CUPT_ROUND(8) break;
				case 2: CUPT_ROUND(2) break;
				case 1: CUPT_ROUND(1) break;
				default: CUPT_ROUND(4) break;
				}
#undef CUPT_ROUND
			}
			CKB(cudaMemcpyAsync(h_state, d_state + (launch & 1u), sizeof(PlocState), cudaMemcpyDeviceToHost, st));
			CKB(cudaStreamSynchronize(st));
			if (h_state->stalled || h_state->n >= bound || h_state->n < 1)
				return set_error(CUPT_ERR_CUDA, "internal: PLOC made no progress (clusters %d -> %d after %u rounds)", bound, h_state->n, h_state->rounds);
			bound = h_state->n;
		}
		cn = h_state->n; node_base = h_state->node_base; rounds = (int)h_state->rounds; cur = rounds & 1;
	}
	if (cn > 1) {   // the remaining rounds in one launch (k_ploc_tail)
		uint32_t* state = reinterpret_cast<uint32_t*>(tile_sums);
		switch (radius) {
		case 32: k_ploc_tail<32><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		case 16: k_ploc_tail<16><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		case 8: k_ploc_tail<8><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		case 2: k_ploc_tail<2><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		case 1: k_ploc_tail<1><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		default: k_ploc_tail<4><<<1, kTile, 0, st>>>(tree, cid[0], cid[1], cbox[0], cbox[1], nn, cn, node_base, (uint32_t)max_leaf, cur, state); break;
		}
		CKB(cudaMemcpyAsync(h_total, state, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		CKB(cudaStreamSynchronize(st));
		const uint32_t* hs = reinterpret_cast<const uint32_t*>(h_total);
		if (hs[3] != 1u || hs[0] != node_base + (uint32_t)cn - 1u)
			return set_error(CUPT_ERR_CUDA, "internal: PLOC tail did not reach one root (n %d -> %u, nodes %u -> %u)", cn, hs[3], node_base, hs[0]);
		node_base = hs[0]; rounds += (int)hs[1]; cur = (int)hs[2]; cn = 1;
	}
	CKB(cudaEventRecord(ev[2], st));
	const int32_t root = N == 1 ? 0 : (int32_t)node_base - 1;

	// 4: collapse + emit, level by level, into the caller's buffers (reused when large enough)
	N2 root_box;
	CKB(cudaMemcpyAsync(&root_box, N == 1 ? static_cast<const void*>(tree.leaf) : static_cast<const void*>(tree.inner + (root - (int32_t)N)), sizeof(N2), cudaMemcpyDeviceToHost, st));
	CKB(cudaStreamSynchronize(st));
	const double rdx = (double)root_box.hi[0] - root_box.lo[0], rdy = (double)root_box.hi[1] - root_box.lo[1], rdz = (double)root_box.hi[2] - root_box.lo[2];
	const double root_area = std::max(2.0 * (rdx * rdy + rdx * rdz + rdy * rdz), 1e-300);
	int32_t* items[2] = {reinterpret_cast<int32_t*>(cid[0]), reinterpret_cast<int32_t*>(cid[1])};
	CKB(cudaMemsetAsync(d_sah, 0, sizeof(double), st));
	if (out->prims_capacity_bytes < n * 48) {
		if (out->prims) { CKB(cudaFree(out->prims)); out->prims = nullptr; out->prims_capacity_bytes = 0; }
		CKB(cudaMalloc(&out->prims, n * 48));
		out->prims_capacity_bytes = n * 48;
	}
	if (out->nodes_capacity_bytes < (n / 4 + 64) * sizeof(Node8)) {
		if (out->nodes) { CKB(cudaFree(out->nodes)); out->nodes = nullptr; out->nodes_capacity_bytes = 0; }
		CKB(cudaMalloc(&out->nodes, (n / 4 + 64) * sizeof(Node8)));
		out->nodes_capacity_bytes = (n / 4 + 64) * sizeof(Node8);
	}
	CKB(cudaMemcpyAsync(items[0], &root, sizeof(int32_t), cudaMemcpyHostToDevice, st));
	int n_items = 1, lvl = 0, depth = 0;
	uint32_t level_base = 0, prim_base = 0;
	while (n_items > 0) {
		++depth;
		const size_t want = ((size_t)level_base + (size_t)n_items) * sizeof(Node8);
		if (want > out->nodes_capacity_bytes) {   // rare (the first guess is N/4 nodes; the optimal cut needs about N/9 at three primitives per leaf): grow the node array
			const size_t ncap = std::max(out->nodes_capacity_bytes * 2, want);
			void* bigger = nullptr;
			CKB(cudaMalloc(&bigger, ncap));
			CKB(cudaMemcpyAsync(bigger, out->nodes, (size_t)level_base * sizeof(Node8), cudaMemcpyDeviceToDevice, st));
			CKB(cudaStreamSynchronize(st));
			CKB(cudaFree(out->nodes));
			out->nodes = bigger; out->nodes_capacity_bytes = ncap;
		}
		const int tiles = (n_items + kTile - 1) / kTile;
		k_collapse_count<<<tiles, kTile, 0, st>>>(tree, items[lvl], n_items, (uint32_t)max_leaf, item_counts, item_kids, tile_sums);
		k_scan_tiles<<<1, kTile, 0, st>>>(tile_sums, tiles, d_total);
		EmitArgs A;
		A.t = tree; A.items = items[lvl]; A.n_items = n_items; A.max_leaf = (uint32_t)max_leaf;
		A.item_prefix = item_counts; A.item_kids = item_kids; A.tile_offs = tile_sums;
		A.level_base = level_base; A.next_base = level_base + (uint32_t)n_items; A.prim_base = prim_base;
		A.next_items = items[lvl ^ 1]; A.out_nodes = static_cast<Node8*>(out->nodes); A.out_prims = static_cast<float4*>(out->prims);
		A.rec_prim = reinterpret_cast<int32_t*>(nn);   // (the PLOC neighbour array is free by now)
		A.inv_root_area = 1.0 / root_area; A.sah = d_sah;
		k_collapse_emit<<<(n_items + kTile / 8 - 1) / (kTile / 8), kTile, 0, st>>>(A);
		CKB(cudaMemcpyAsync(h_total, d_total, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		CKB(cudaStreamSynchronize(st));
		level_base += (uint32_t)n_items;
		prim_base += (uint32_t)(*h_total >> 32);
		n_items = (int)(uint32_t)*h_total;
		lvl ^= 1;
		// the same bound cupt_set_bvh applies: the extend kernel's shared-memory stack holds `depth` entries per thread
		if (depth > 64) return set_error(CUPT_ERR_LIMIT, "wide BVH deeper than 64 levels (the traversal stack's limit, as for cupt_set_bvh)");
	}
	k_emit_prims<<<tilesN, kTile, 0, st>>>(reinterpret_cast<const int32_t*>(nn), N, records, static_cast<float4*>(out->prims));
	CKB(cudaEventRecord(ev[3], st));
	double sah = 0;
	CKB(cudaMemcpyAsync(&sah, d_sah, sizeof(double), cudaMemcpyDeviceToHost, st));
	CKB(cudaStreamSynchronize(st));
	CKB(cudaGetLastError());
	if ((int64_t)prim_base != N) return set_error(CUPT_ERR_CUDA, "internal: emitted %u primitive records for %lld primitives", prim_base, (long long)N);
	out->num_nodes = level_base; out->num_prims = N;
	out->depth = depth; out->ploc_rounds = rounds; out->sah_cost = sah;
	cudaEventElapsedTime(&out->ms_sort, ev[0], ev[1]);
	cudaEventElapsedTime(&out->ms_ploc, ev[1], ev[2]);
	cudaEventElapsedTime(&out->ms_collapse, ev[2], ev[3]);
	cudaEventElapsedTime(&out->ms_total, ev[0], ev[3]);
	return CUPT_OK;
}

// test hook (cupt_debug_sort_pairs): the build's radix sort on caller-supplied pairs, in place
int debug_sort_pairs(cudaStream_t st, uint64_t* h_keys, uint32_t* h_vals, int64_t n, int passes)
{
	if (n < 0 || n > (1ll << 27) || passes < 1 || passes > 8) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (n == 0) return CUPT_OK;
	uint64_t* keys[2] = {nullptr, nullptr}; uint32_t* vals[2] = {nullptr, nullptr}; uint32_t* scratch = nullptr;
	cudaError_t e = cudaSuccess;
	for (int b = 0; b < 2 && e == cudaSuccess; ++b) { e = cudaMalloc((void**)&keys[b], (size_t)n * 8); if (e == cudaSuccess) e = cudaMalloc((void**)&vals[b], (size_t)n * 4); }
	if (e == cudaSuccess) e = cudaMalloc((void**)&scratch, rsort::scratch_bytes((uint32_t)n));
	if (e == cudaSuccess) e = cudaMemcpyAsync(keys[0], h_keys, (size_t)n * 8, cudaMemcpyHostToDevice, st);
	if (e == cudaSuccess) e = cudaMemcpyAsync(vals[0], h_vals, (size_t)n * 4, cudaMemcpyHostToDevice, st);
	if (e == cudaSuccess) {
		const int r = rsort::sort_pairs(st, keys, vals, (uint32_t)n, passes, scratch);
		e = cudaMemcpyAsync(h_keys, keys[r], (size_t)n * 8, cudaMemcpyDeviceToHost, st);
		if (e == cudaSuccess) e = cudaMemcpyAsync(h_vals, vals[r], (size_t)n * 4, cudaMemcpyDeviceToHost, st);
		if (e == cudaSuccess) e = cudaStreamSynchronize(st);
		if (e == cudaSuccess) e = cudaGetLastError();
	}
	for (int b = 0; b < 2; ++b) { cudaFree(keys[b]); cudaFree(vals[b]); }
	cudaFree(scratch);
	return e == cudaSuccess ? CUPT_OK : set_error(CUPT_ERR_CUDA, "debug_sort_pairs: %s", cudaGetErrorString(e));
}

}  // namespace cupt
