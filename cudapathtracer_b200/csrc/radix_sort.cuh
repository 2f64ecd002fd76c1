// Stable LSD radix sort of (64-bit key, 32-bit value) pairs on the device, 8 bits per pass, for the Morton order of
// the BVH build (bvh_device.cu).  Three kernels per pass:
//
//   k_radix_hist     one block per tile of 4096 pairs: digit histogram in shared memory -> hist[digit][tile]
//   scan             exclusive prefix sum over hist in digit-major order (k_scan32_reduce, k_scan32_tiles,
//                    k_scan32_apply), which turns it into the global start of every (digit, tile) run
//   k_radix_scatter  one block per tile, each warp owns 256 consecutive pairs: __match_any_sync ranks equal digits
//                    inside a round of 32, a warp-private counter array carries the rank across the 8 rounds, a
//                    prefix over the 16 warps per digit gives the place inside the tile's run.  Order inside a run is
//                    (warp, round, lane) = memory order, so the sort is stable.  The tile is put in sorted order in
//                    shared memory first and written out from there, so that a warp's stores are consecutive
//                    addresses inside the digits' runs.
//
// Only the top `bits` bits of the 63-bit code are sorted (the caller asks for log2(N) + 9 rounded up to a multiple of
// 8): finer Morton cells than 512 per primitive cannot change a clustering window of +-4.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace cupt {
namespace rsort {

constexpr int kThreads = 256;
constexpr int kItems = 16;                    // pairs per thread
constexpr int kTileN = kThreads * kItems;     // 4096 pairs per block
constexpr int kWarps = kThreads / 32;
constexpr int kScanTile = 2048;               // elements per block of the 32-bit scan

__global__ void __launch_bounds__(kThreads) k_radix_hist(const uint64_t* __restrict__ keys, const uint32_t n, const int shift,
                                                         uint32_t* __restrict__ hist, const uint32_t ntiles)
{
	__shared__ uint32_t s_hist[256];
	s_hist[threadIdx.x] = 0u;
	__syncthreads();
	const uint32_t base = blockIdx.x * (uint32_t)kTileN;
#pragma unroll 4
	for (int i = 0; i < kItems; ++i) {
		const uint32_t idx = base + (uint32_t)i * kThreads + threadIdx.x;
		if (idx < n) atomicAdd(&s_hist[(uint32_t)(keys[idx] >> shift) & 255u], 1u);
	}
	__syncthreads();
	hist[(size_t)threadIdx.x * ntiles + blockIdx.x] = s_hist[threadIdx.x];
}

// ---- 32-bit exclusive scan, in place: per-tile sums, scan of the sums by one block, per-tile scan + offset --------
__device__ __forceinline__ uint32_t warp_incl32(uint32_t v)
{
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
		if ((threadIdx.x & 31) >= d) v += o;
	}
	return v;
}
// exclusive scan of one value per thread over the block; *total = block sum.  s_w: kWarps + 1 words.
__device__ __forceinline__ uint32_t block_excl32(const uint32_t v, uint32_t* s_w, uint32_t* total)
{
	const uint32_t incl = warp_incl32(v);
	const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
	__syncthreads();
	if (l == 31) s_w[w] = incl;
	__syncthreads();
	if (w == 0) {
		const uint32_t x = l < kWarps ? s_w[l] : 0u;
		const uint32_t xi = warp_incl32(x);
		if (l < kWarps) s_w[l] = xi - x;
		if (l == kWarps - 1) s_w[kWarps] = xi;
	}
	__syncthreads();
	*total = s_w[kWarps];
	return s_w[w] + incl - v;
}

__global__ void __launch_bounds__(kThreads) k_scan32_reduce(const uint32_t* __restrict__ data, const uint32_t n, uint32_t* __restrict__ tile_sums)
{
	__shared__ uint32_t s_w[kWarps + 1];
	const uint32_t base = blockIdx.x * (uint32_t)kScanTile + threadIdx.x * (kScanTile / kThreads);
	uint32_t v = 0u;
#pragma unroll
	for (int i = 0; i < kScanTile / kThreads; ++i)
		if (base + i < n) v += data[base + i];
	uint32_t tot;
	block_excl32(v, s_w, &tot);
	if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kThreads) k_scan32_tiles(uint32_t* __restrict__ tile_sums, const uint32_t ntiles)
{
	__shared__ uint32_t s_w[kWarps + 1];
	uint32_t carry = 0u;
	for (uint32_t base = 0; base < ntiles; base += kThreads) {
		const uint32_t i = base + threadIdx.x;
		const uint32_t v = i < ntiles ? tile_sums[i] : 0u;
		uint32_t tot;
		const uint32_t ex = block_excl32(v, s_w, &tot);
		if (i < ntiles) tile_sums[i] = carry + ex;
		carry += tot;
	}
}

__global__ void __launch_bounds__(kThreads) k_scan32_apply(uint32_t* __restrict__ data, const uint32_t n, const uint32_t* __restrict__ tile_offs)
{
	__shared__ uint32_t s_w[kWarps + 1];
	constexpr int per = kScanTile / kThreads;
	const uint32_t base = blockIdx.x * (uint32_t)kScanTile + threadIdx.x * per;
	uint32_t x[per];
	uint32_t v = 0u;
#pragma unroll
	for (int i = 0; i < per; ++i) { x[i] = base + i < n ? data[base + i] : 0u; v += x[i]; }
	uint32_t tot;
	uint32_t run = tile_offs[blockIdx.x] + block_excl32(v, s_w, &tot);
#pragma unroll
	for (int i = 0; i < per; ++i) {
		if (base + i < n) data[base + i] = run;
		run += x[i];
	}
}

// Dynamic shared memory of k_radix_scatter: the tile's pairs in sorted order (keys, values), then per digit the offset that
// turns a position inside the sorted tile into a global position.
constexpr size_t kScatterSmem = (size_t)kTileN * (sizeof(uint64_t) + sizeof(uint32_t)) + 256 * sizeof(uint32_t);
// 512 threads x 8 pairs (round 1: 256 x 16 — 100 registers and 58 KB of shared memory per block left two blocks = 16 warps per SM,
// issue slots 11 % busy, DRAM at 19 % of its peak: profiles/r2w_builder_metrics.md)
constexpr int kScThreads = 512;
constexpr int kScItems = kTileN / kScThreads;
constexpr int kScWarps = kScThreads / 32;

__global__ void __launch_bounds__(kScThreads, 2) k_radix_scatter(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, const uint32_t n,
                                                                 const int shift, const uint32_t* __restrict__ offs, const uint32_t ntiles)
{
	extern __shared__ uint64_t s_dyn[];
	uint64_t* s_key = s_dyn;                                                   // [kTileN] the tile in sorted order
	uint32_t* s_val = reinterpret_cast<uint32_t*>(s_dyn + kTileN);             // [kTileN]
	uint32_t* s_gofs = s_val + kTileN;                                         // [256] global start of the digit's run - its start inside the sorted tile
	__shared__ uint32_t s_cnt[kScWarps][256];   // per warp: pairs seen so far per digit; later: the warp's start inside the sorted tile
	__shared__ uint32_t s_w[kScWarps + 1];
	const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
	const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
	for (int i = 0; i < kScWarps * 256 / kScThreads; ++i) (&s_cnt[0][0])[i * kScThreads + threadIdx.x] = 0u;
	__syncthreads();
	const uint32_t wbase = blockIdx.x * (uint32_t)kTileN + w * (uint32_t)(kScItems * 32);
	uint64_t key[kScItems];
	uint32_t rank[kScItems];
#pragma unroll
	for (int i = 0; i < kScItems; ++i) {
		const uint32_t idx = wbase + (uint32_t)i * 32u + lane;
		const bool valid = idx < n;
		key[i] = valid ? keys_in[idx] : ~0ull;
		const uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
		// lanes beyond n take part in the vote with an impossible digit so that they never share a group with a valid lane
		const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 256u + lane);
		const uint32_t before = s_cnt[w][d];
		rank[i] = before + (uint32_t)__popc(peers & lt);
		__syncwarp();
		if (valid && (peers & lt) == 0u) s_cnt[w][d] = before + (uint32_t)__popc(peers);   // the group's lowest lane
		__syncwarp();
	}
	__syncthreads();
	{   // thread t < 256 = digit t: the digit's start inside the sorted tile (exclusive scan of the tile's digit counts), the warps'
		// starts behind it, and the offset to the global run of this (digit, tile); threads 256.. only take part in the scan
		const bool digit = threadIdx.x < 256u;
		uint32_t tot = 0u;
		if (digit) {
#pragma unroll
			for (int i = 0; i < kScWarps; ++i) tot += s_cnt[i][threadIdx.x];
		}
		// exclusive scan over the block (kScWarps warps)
		const uint32_t incl = warp_incl32(tot);
		if (lane == 31u) s_w[w] = incl;
		__syncthreads();
		if (w == 0u) {
			const uint32_t x = lane < (uint32_t)kScWarps ? s_w[lane] : 0u;
			const uint32_t xi = warp_incl32(x);
			if (lane < (uint32_t)kScWarps) s_w[lane] = xi - x;
		}
		__syncthreads();
		if (digit) {
			uint32_t run = s_w[w] + incl - tot;
			s_gofs[threadIdx.x] = offs[(size_t)threadIdx.x * ntiles + blockIdx.x] - run;
#pragma unroll
			for (int i = 0; i < kScWarps; ++i) { const uint32_t c = s_cnt[i][threadIdx.x]; s_cnt[i][threadIdx.x] = run; run += c; }
		}
	}
	__syncthreads();
	// the tile in sorted order in shared memory: order inside a digit is (warp, round, lane) = memory order, so the sort is stable
#pragma unroll
	for (int i = 0; i < kScItems; ++i) {
		const uint32_t idx = wbase + (uint32_t)i * 32u + lane;
		if (idx < n) {
			const uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
			const uint32_t pos = s_cnt[w][d] + rank[i];
			s_key[pos] = key[i];
			s_val[pos] = vals_in[idx];
		}
	}
	__syncthreads();
	// ... and out to global memory in that order: consecutive threads write consecutive addresses inside every digit's run
	// (the first version scattered straight from registers, 32 different runs per warp store: 6.9 ms per pass for 100 M
	// pairs, 350 GB/s)
	const uint32_t tile_n = min((uint32_t)kTileN, n - blockIdx.x * (uint32_t)kTileN);
#pragma unroll 4
	for (uint32_t j = threadIdx.x; j < tile_n; j += kScThreads) {
		const uint64_t k = s_key[j];
		const uint32_t pos = s_gofs[(uint32_t)(k >> shift) & 255u] + j;
		keys_out[pos] = k;
		vals_out[pos] = s_val[j];
	}
}

// bytes of scratch the sort needs besides the two (key, value) buffer pairs
inline size_t scratch_bytes(const uint32_t n)
{
	const size_t ntiles = (n + kTileN - 1) / kTileN;
	const size_t hist = ntiles * 256;
	const size_t stiles = (hist + kScanTile - 1) / kScanTile;
	return (hist + stiles + 64) * sizeof(uint32_t);
}

// sorts on key bits [63 - 8 * passes, 63); returns the index (0 / 1) of the buffer pair that holds the result
inline int sort_pairs(cudaStream_t st, uint64_t* keys[2], uint32_t* vals[2], const uint32_t n, const int passes, uint32_t* scratch)
{
	const uint32_t ntiles = (n + kTileN - 1) / kTileN;
	const uint32_t hist_n = ntiles * 256u;
	const uint32_t stiles = (hist_n + kScanTile - 1) / kScanTile;
	uint32_t* hist = scratch;
	uint32_t* tile_sums = scratch + hist_n;
	// 50 KB of dynamic shared memory: above the 48 KB default.  Per call, not once per process: the attribute belongs to the current
	// DEVICE, and a process may build on several (cupt_group)
	cudaFuncSetAttribute(k_radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kScatterSmem);
	int cur = 0;
	for (int p = 0; p < passes; ++p) {
		const int shift = 63 - 8 * (passes - p);
		k_radix_hist<<<ntiles, kThreads, 0, st>>>(keys[cur], n, shift < 0 ? 0 : shift, hist, ntiles);
		k_scan32_reduce<<<stiles, kThreads, 0, st>>>(hist, hist_n, tile_sums);
		k_scan32_tiles<<<1, kThreads, 0, st>>>(tile_sums, stiles);
		k_scan32_apply<<<stiles, kThreads, 0, st>>>(hist, hist_n, tile_sums);
		k_radix_scatter<<<ntiles, kScThreads, kScatterSmem, st>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift < 0 ? 0 : shift, hist, ntiles);
		cur ^= 1;
	}
	return cur;
}

}  // namespace rsort
}  // namespace cupt
