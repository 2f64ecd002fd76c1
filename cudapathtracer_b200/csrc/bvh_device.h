// GPU-side wide-BVH construction (bvh_device.cu); internal to libcupt, called by cupt_build_bvh_device (context.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/cupt.h"

namespace cupt {

// workspace kept by the context between builds: one device allocation that only grows, a pinned word, four events
struct BuildArena {
	void* base = nullptr;
	size_t cap = 0, off = 0;
	uint64_t* h_total = nullptr;
	cudaEvent_t ev[4] = {};
	int reserve(size_t bytes);
	void destroy();
};

struct DeviceBvh {
	// in/out: the caller's buffers (cudaMalloc) and their capacities; reused when large enough, else replaced
	void* nodes = nullptr;      // Node8[num_nodes]
	void* prims = nullptr;      // Prim48[num_prims]
	size_t nodes_capacity_bytes = 0, prims_capacity_bytes = 0;
	// out
	int64_t num_nodes = 0, num_prims = 0;
	int depth = 0, ploc_rounds = 0;
	double sah_cost = 0;
	float ms_total = 0, ms_sort = 0, ms_ploc = 0, ms_collapse = 0;   // CUDA events on the build stream
};

// verts/tris/spheres: the device arrays of the uploaded scene (spheres: SphereShade records, centre + radius first);
// rank: optional device array of tie-break ranks per triangle (NULL = triangle index)
int build_wide_bvh_device(cudaStream_t stream, BuildArena& arena, const float* verts, const cupt_triangle* tris, int64_t num_tris,
                          const float4* spheres, int64_t num_spheres, const uint32_t* rank, int max_leaf,
                          int radius /* 1, 2, 4 (default), 8, 16, 32 */, DeviceBvh* out);

// test hook: the build's radix sort on caller-supplied host pairs, in place (sorted on key bits [63 - 8 * passes, 63))
int debug_sort_pairs(cudaStream_t stream, uint64_t* keys, uint32_t* vals, int64_t n, int passes);

}  // namespace cupt
