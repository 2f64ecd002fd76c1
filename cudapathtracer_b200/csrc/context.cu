// Device context and the render loop behind the C-ABI (include/cupt.h): the replacement of the
// cudaMalloc/cudaMemcpy block and the per-sample launch loop of the reference's main() (kernel.cu:633-761).
// One context per GPU; every launch goes to the context's stream; nothing here synchronises the device except
// cupt_sync and the read-back calls.  There is no CPU fallback: without a CUDA device cupt_create fails.
#include "host_common.h"
#include "layout.h"
#include "kernels.cuh"
#include "bdpt.cuh"
#include "bvh_device.h"
#include "output.cuh"
#include <algorithm>
#include <map>
#include <cstring>
#include <thread>

using namespace cupt;

#define CK(call)                                                                                           \
	do {                                                                                                   \
		cudaError_t e_ = (call);                                                                           \
		if (e_ != cudaSuccess)                                                                             \
			return set_error(CUPT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
	} while (0)

struct StageEvent {
	cudaEvent_t a, b;
	int stage;  // 0 raygen, 1 extend, 2 shade, 3 accumulate
};

struct cupt_context {
	int device = 0, num_sms = 0;
	cudaStream_t stream = nullptr;
	bool own_stream = false;
	// scene
	float* d_verts = nullptr; cupt_triangle* d_tris = nullptr; float4* d_trishade = nullptr; float4* d_mats = nullptr;
	float4* d_lights = nullptr; float* d_light_cdf = nullptr; float4* d_spheres = nullptr; int num_spheres = 0;
	int64_t num_tris = 0; int num_lights = 0; float total_light_area = 0.f;
	float emis_lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, emis_hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};   // bounds of the emissive primitives (cupt_set_scene)
	// acceleration structures
	uint4* d_nodes = nullptr; float4* d_prims = nullptr; int bvh_depth = 0; int64_t num_nodes = 0, num_prims = 0;
	cupt_bvh2_node* d_bvh2 = nullptr; int bvh2_size = 0, bvh2_depth = 0;
	cupt_bvh2_node* d_bvh2_up = nullptr; uint32_t* d_bvh2_leaf_parent = nullptr; int64_t bvh2_leaves = 0;
	// camera / options
	cupt_camera cam{}; bool have_cam = false;
	cupt_render_opts opts{};
	// wavefront state
	PathBuffers pb{}; size_t path_capacity = 0, retrace_capacity = 0;
	float* d_accum = nullptr; size_t accum_pixels = 0; uint64_t samples_accumulated = 0;
	// bookkeeping
	uint64_t launches = 0, extend_launches = 0;
	double ms[4] = {0, 0, 0, 0};
	std::vector<StageEvent> pending, pool;
	std::map<void**, size_t> capacity;   // bytes behind each uploaded device pointer (upload())
	std::map<void**, size_t> used;       // bytes of it that the last upload filled (what a group broadcasts, group.cuh)
	// persistent grid of the extend stage per kernel instantiation (mode, stats, vis): blocks, and the dynamic shared memory it was sized for
	int extend_blocks[2][2][2] = {}; size_t extend_smem[2][2][2] = {};
	uint32_t* d_retrace_counts = nullptr; size_t retrace_pixels = 0;   // cupt_render_opts.record_retraces
	BuildArena arena; uint32_t* d_build_rank = nullptr; uint8_t* d_rgb8 = nullptr;
	float4* d_prims_wt = nullptr; bool wt_valid = false;   // CUPT_TRI_WATERTIGHT: vertex records, same order as d_prims
	float build_ms[4] = {0, 0, 0, 0}; int build_rounds = 0;   // last cupt_build_bvh_device: total, sort, PLOC, collapse
	int tile_first = 0, tile_count = 0;   // cupt_set_tile_window (count 0 = the whole image)
	// pinned staging for host -> device uploads (stage_alloc / upload)
	uint8_t* h_stage = nullptr; size_t stage_cap = 0, stage_used = 0; cudaEvent_t stage_ev = nullptr; bool stage_pending = false;
	bool pageable_upload = false;   // an upload of this call bypassed the staging buffer
};

namespace {

// Device buffers keep their capacity: re-uploading a scene or BVH of the same (or a smaller) size neither frees nor
// allocates, so it does not synchronise the device (cudaFree does) and is ordered on the context's stream behind
// whatever render is still in flight.  The caller synchronises once after its last upload (staging temporaries).
//
// Host -> device copies go through a pinned staging buffer owned by the context: the caller's (pageable) array is
// copied into it by all host threads and DMA'd from there, so the DMA of one array overlaps the host copy of the
// next and the call returns without waiting for the device (a pageable cudaMemcpyAsync is staged by ONE driver
// thread at ~12 GB/s: 2.4 ms for the 28 MB of config 2, now 0.8 ms).  The buffer is a bump allocator that is rewound
// once the last DMA out of it has completed; uploads larger than kStageMax take the pageable path.
const size_t kStageMax = (size_t)256 << 20;

void* stage_alloc(cupt_context* c, size_t bytes)
{
	bytes = (bytes + 255) & ~(size_t)255;
	if (bytes > kStageMax) return nullptr;
	if (c->stage_pending && (c->stage_used + bytes > c->stage_cap || cudaEventQuery(c->stage_ev) == cudaSuccess)) {
		if (cudaEventSynchronize(c->stage_ev) != cudaSuccess) return nullptr;
		c->stage_pending = false; c->stage_used = 0;
	}
	if (c->stage_used + bytes > c->stage_cap) {   // (nothing is in flight here)
		if (c->h_stage) { cudaFreeHost(c->h_stage); c->h_stage = nullptr; c->stage_cap = 0; }
		const size_t want = std::min(kStageMax, std::max(bytes * 2, (size_t)32 << 20));
		if (cudaHostAlloc((void**)&c->h_stage, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
		c->stage_cap = want; c->stage_used = 0;
	}
	if (!c->stage_ev && cudaEventCreateWithFlags(&c->stage_ev, cudaEventDisableTiming) != cudaSuccess) return nullptr;
	void* p = c->h_stage + c->stage_used;
	c->stage_used += bytes;
	return p;
}

// host threads for the staging copies: the machine's cores shared between the GPUs it drives (one process per GPU),
// 2..8 each; set explicitly because launchers such as torchrun export OMP_NUM_THREADS=1
__attribute__((unused)) int copy_threads()
{
	static int n = 0;
	if (n == 0) {
		int gpus = 1;
		if (cudaGetDeviceCount(&gpus) != cudaSuccess || gpus < 1) gpus = 1;
		const int hw = (int)std::thread::hardware_concurrency();
		n = std::max(2, std::min(8, hw / gpus));
	}
	return n;
}

void parallel_copy(void* dst, const void* src, size_t bytes)
{
	const size_t chunk = (size_t)1 << 20;
	if (bytes <= 2 * chunk) { memcpy(dst, src, bytes); return; }
	const int64_t n = (int64_t)((bytes + chunk - 1) / chunk);
#pragma omp parallel for schedule(static) num_threads(copy_threads())
	for (int64_t i = 0; i < n; ++i) {
		const size_t o = (size_t)i * chunk;
		memcpy((uint8_t*)dst + o, (const uint8_t*)src + o, std::min(chunk, bytes - o));
	}
}

template <class T>
int ensure_device(cupt_context* c, T** dst, size_t bytes)
{
	size_t& cap = c->capacity[(void**)dst];
	c->used[(void**)dst] = bytes;
	if (bytes > cap || !*dst) {
		if (*dst) { CK(cudaFree(*dst)); *dst = nullptr; cap = 0; }
		if (bytes == 0) return CUPT_OK;
		CK(cudaMalloc((void**)dst, bytes));
		cap = bytes;
	}
	return CUPT_OK;
}

// DMA of `bytes` that the caller has written to `staged` (from stage_alloc)
template <class T>
int upload_staged(cupt_context* c, T** dst, const void* staged, size_t bytes)
{
	int rc;
	if ((rc = ensure_device(c, dst, bytes))) return rc;
	if (bytes == 0) return CUPT_OK;
	CK(cudaMemcpyAsync(*dst, staged, bytes, cudaMemcpyHostToDevice, c->stream));
	CK(cudaEventRecord(c->stage_ev, c->stream));
	c->stage_pending = true;
	return CUPT_OK;
}

template <class T>
int upload(cupt_context* c, T** dst, const void* src, size_t bytes)
{
	int rc;
	if ((rc = ensure_device(c, dst, bytes))) return rc;
	if (bytes == 0) return CUPT_OK;
	void* st = stage_alloc(c, bytes);
	if (st) {
		parallel_copy(st, src, bytes);
		return upload_staged(c, dst, st, bytes);
	}
	c->pageable_upload = true;
	CK(cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
	return CUPT_OK;
}

// after the uploads of one call: the caller's arrays may be freed on return, which only a pageable copy still reads
int uploads_done(cupt_context* c)
{
	if (c->pageable_upload) { c->pageable_upload = false; CK(cudaStreamSynchronize(c->stream)); }
	return CUPT_OK;
}

int ensure_paths(cupt_context* c, size_t n)
{
	if (n <= c->path_capacity) return CUPT_OK;
	cudaFree(c->pb.rayA); cudaFree(c->pb.rayB); cudaFree(c->pb.st0); cudaFree(c->pb.st1);
	cudaFree(c->pb.queue[0]); cudaFree(c->pb.queue[1]);
	c->pb.rayA = c->pb.rayB = c->pb.st0 = c->pb.st1 = nullptr; c->pb.queue[0] = c->pb.queue[1] = nullptr;
	c->path_capacity = 0;
	CK(cudaMalloc((void**)&c->pb.rayA, n * sizeof(float4)));
	CK(cudaMalloc((void**)&c->pb.rayB, n * sizeof(float4)));
	CK(cudaMalloc((void**)&c->pb.st0, n * sizeof(float4)));
	CK(cudaMalloc((void**)&c->pb.st1, n * sizeof(float4)));
	CK(cudaMalloc((void**)&c->pb.queue[0], n * sizeof(uint32_t)));
	CK(cudaMalloc((void**)&c->pb.queue[1], n * sizeof(uint32_t)));
	c->path_capacity = n;
	return CUPT_OK;
}

int ensure_accum(cupt_context* c)
{
	const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
	if (c->d_accum && c->accum_pixels == P) return CUPT_OK;
	if (c->d_accum) { CK(cudaFree(c->d_accum)); c->d_accum = nullptr; }
	CK(cudaMalloc((void**)&c->d_accum, P * 3 * sizeof(float)));
	CK(cudaMemsetAsync(c->d_accum, 0, P * 3 * sizeof(float), c->stream));
	c->accum_pixels = P;
	c->samples_accumulated = 0;
	return CUPT_OK;
}

// per-pixel re-trace counters of the strict mode (cupt_render_opts.record_retraces)
int ensure_retrace_counts(cupt_context* c)
{
	if (!c->opts.record_retraces || c->opts.traversal != CUPT_TRAVERSAL_WIDE_STRICT) return CUPT_OK;
	const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
	if (c->d_retrace_counts && c->retrace_pixels == P) return CUPT_OK;
	if (c->d_retrace_counts) { CK(cudaFree(c->d_retrace_counts)); c->d_retrace_counts = nullptr; }
	CK(cudaMalloc((void**)&c->d_retrace_counts, P * sizeof(uint32_t)));
	CK(cudaMemsetAsync(c->d_retrace_counts, 0, P * sizeof(uint32_t), c->stream));
	c->retrace_pixels = P;
	return CUPT_OK;
}

void stage_begin(cupt_context* c, int stage, StageEvent& ev)
{
	if (!c->opts.time_stages) return;
	if (!c->pool.empty()) { ev = c->pool.back(); c->pool.pop_back(); }
	else { cudaEventCreate(&ev.a); cudaEventCreate(&ev.b); }
	ev.stage = stage;
	cudaEventRecord(ev.a, c->stream);
}
void stage_end(cupt_context* c, StageEvent& ev)
{
	if (!c->opts.time_stages) return;
	cudaEventRecord(ev.b, c->stream);
	c->pending.push_back(ev);
}
void drain_events(cupt_context* c)
{
	for (StageEvent& ev : c->pending) {
		float ms = 0.f;
		if (cudaEventSynchronize(ev.b) == cudaSuccess && cudaEventElapsedTime(&ms, ev.a, ev.b) == cudaSuccess)
			c->ms[ev.stage] += ms;
		c->pool.push_back(ev);
	}
	c->pending.clear();
}

// window: restrict the pass to the context's tile window (render passes); the parity hooks always cover the whole image
PassParams make_pass(const cupt_context* c, int first_sample, int samples, int fixed_lens, bool window = false)
{
	PassParams pp;
	pp.cam = c->cam;
	pp.seed = c->opts.seed;
	pp.first_sample = first_sample;
	pp.samples = samples;
	pp.tiles_x = (c->cam.pxlWidth + 7) / 8;
	pp.tiles_y = (c->cam.pxlHeight + 3) / 4;
	const int all_tiles = pp.tiles_x * pp.tiles_y;   // <= 8192 x 16384 = 2^27 for the 65535 x 65535 limit of cupt_set_camera
	pp.tile0 = 0;
	int ntiles = all_tiles;
	if (window && c->tile_count > 0) {
		pp.tile0 = std::min(c->tile_first, all_tiles);
		ntiles = std::max(0, std::min(c->tile_count, all_tiles - pp.tile0));
	}
	pp.p_pad = std::max(1, ntiles) * 32;   // (an empty window is rejected by cupt_render)
	pp.max_depth = c->opts.max_depth;
	pp.fixed_lens = fixed_lens;
	pp.inv_p_pad = ~0ull / (uint64_t)pp.p_pad + 1ull;                      // p_pad >= 32
	pp.inv_tiles_x = pp.tiles_x > 1 ? ~0ull / (uint64_t)pp.tiles_x + 1ull : 0ull;
	// last segments stop at the first hit in front of every emitter (k_shade): wide traversal only — the strict mode promises the
	// reference's hit on EVERY ray.  Stats runs do the same, so that their node / primitive counts are those of the timed kernels.
	pp.term_stop = (c->opts.traversal == CUPT_TRAVERSAL_WIDE && !c->opts.no_terminal_stop) ? 1 : 0;
	for (int a = 0; a < 3; ++a) { pp.emis_lo[a] = c->emis_lo[a]; pp.emis_hi[a] = c->emis_hi[a]; }
	return pp;
}

DevScene make_scene(const cupt_context* c)
{
	DevScene sc;
	sc.nodes = c->d_nodes; sc.prims = (c->opts.tri_test == CUPT_TRI_WATERTIGHT && c->wt_valid) ? c->d_prims_wt : c->d_prims; sc.bvh2 = c->d_bvh2; sc.verts = c->d_verts; sc.tris = c->d_tris;
	sc.trishade = c->d_trishade; sc.mats = c->d_mats; sc.bvh2_up = c->d_bvh2_up; sc.bvh2_leaf_parent = c->d_bvh2_leaf_parent;
	sc.lights = c->d_lights; sc.light_cdf = c->d_light_cdf; sc.spheres = c->d_spheres;
	sc.numLights = c->num_lights; sc.totalLightArea = c->total_light_area;
	sc.stack_cap = std::max(c->bvh_depth, 1);
	return sc;
}

int check_ready(cupt_context* c, bool need_shading, bool need_cam = true)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	if (need_cam && !c->have_cam) return set_error(CUPT_ERR_INVALID, "no camera set (cupt_set_camera)");
	if (c->opts.traversal != CUPT_TRAVERSAL_REF_BVH2 && (!c->d_nodes || c->num_nodes == 0)) return set_error(CUPT_ERR_INVALID, "no wide BVH set (cupt_set_bvh)");
	if (c->opts.traversal != CUPT_TRAVERSAL_WIDE && (!c->d_bvh2 || !c->d_tris)) return set_error(CUPT_ERR_INVALID, "no reference BVH2 / scene set");
	if (c->opts.traversal == CUPT_TRAVERSAL_WIDE_STRICT && (c->num_spheres > 0 || c->bvh2_leaves != c->num_tris))
		return set_error(CUPT_ERR_INVALID, "strict mode needs a sphere-free scene and the reference BVH2 of the same scene");
	if (need_shading && !c->d_trishade) return set_error(CUPT_ERR_INVALID, "no scene set (cupt_set_scene)");
	if (c->opts.tri_test == CUPT_TRI_WATERTIGHT) {
		if (c->opts.traversal != CUPT_TRAVERSAL_WIDE)
			return set_error(CUPT_ERR_INVALID, "CUPT_TRI_WATERTIGHT needs CUPT_TRAVERSAL_WIDE (the other modes reproduce the reference's own test)");
		if (!c->d_tris && c->num_prims > c->num_spheres) return set_error(CUPT_ERR_INVALID, "no scene set (cupt_set_scene)");
		if (!c->wt_valid) {   // vertex records (kind 2) in the order of the wide BVH's primitive array, made once per BVH
			const size_t bytes = (size_t)c->num_prims * 48;
			if (c->capacity[(void**)&c->d_prims_wt] < bytes) {
				if (c->d_prims_wt) { CK(cudaFree(c->d_prims_wt)); c->d_prims_wt = nullptr; }
				CK(cudaMalloc((void**)&c->d_prims_wt, bytes));
				c->capacity[(void**)&c->d_prims_wt] = bytes;
			}
			k_vertex_prims<<<(unsigned)((c->num_prims + 255) / 256), 256, 0, c->stream>>>(c->d_prims, c->num_prims, c->d_verts, c->d_tris, c->d_prims_wt);
			CK(cudaGetLastError());
			c->launches++;
			c->wt_valid = true;
		}
	}
	return CUPT_OK;
}

// strict mode, record_retraces: one count per pixel and re-traced ray (slots_per_path: 4 ray records per BDPT path)
__global__ void __launch_bounds__(256) k_record_retraces(PassParams pp, PathBuffers pb, uint32_t* __restrict__ counts, uint32_t slots_per_path)
{
	const uint32_t n = pb.counters[126];
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		int x, y, sl;
		slot_to_pixel(pp, pb.retrace[i] / slots_per_path, x, y, sl);
		if (x < pp.cam.pxlWidth && y < pp.cam.pxlHeight) atomicAdd(&counts[(size_t)y * pp.cam.pxlWidth + x], 1u);
	}
}

// persistent launch of the extend stage for queue `depth`; pp (render passes only): where the strict mode records its re-traces
int launch_extend(cupt_context* c, const DevScene& sc, int depth, int qsel, bool vis = false, const PassParams* pp = nullptr, uint32_t slots_per_path = 1)
{
	const int mode = c->opts.traversal == CUPT_TRAVERSAL_REF_BVH2 ? 1 : 0;
	const bool stats = c->opts.collect_stats != 0;
	if (mode == 1) vis = false;
	const size_t smem = mode == 0 ? (size_t)sc.stack_cap * 128 * sizeof(uint2) + 128 * 16 : 0;   // stacks + per-warp keys/list/ids
	ExtendTuning tune;
	tune.fetch_thresh = c->opts.fetch_threshold > 0 ? std::min(c->opts.fetch_threshold, 31) : 24;
	tune.defer_pairs = c->opts.defer_pairs > 0 ? std::min(c->opts.defer_pairs, 64) : 20;
	tune.qbias_bits = CUPT_QBIAS_BITS;
	void (*kern)(DevScene, PathBuffers, int, int, ExtendTuning) =
		mode == 0 ? (vis ? (stats ? k_extend<0, true, true> : k_extend<0, false, true>) : (stats ? k_extend<0, true, false> : k_extend<0, false, false>))
		          : (stats ? k_extend<1, true, false> : k_extend<1, false, false>);
	int& blocks = c->extend_blocks[mode][stats ? 1 : 0][vis ? 1 : 0];
	size_t& sized_for = c->extend_smem[mode][stats ? 1 : 0][vis ? 1 : 0];
	if (blocks == 0 || sized_for != smem) {   // once per instantiation and stack size (cupt_set_bvh / cupt_build_bvh_device change the latter)
		if (smem > 48 * 1024) CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		int per_sm = 0;
		CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
		if (per_sm < 1) return set_error(CUPT_ERR_LIMIT, "the extend kernel does not fit an SM with a traversal stack of %d entries (%zu bytes of shared memory per block)", sc.stack_cap, smem);
#ifdef CUPT_EXTEND_MAX_BLOCKS   // experiment: leave room on every SM for another stream's kernels
		per_sm = std::min(per_sm, CUPT_EXTEND_MAX_BLOCKS);
#endif
		blocks = per_sm * c->num_sms;
		sized_for = smem;
	}
	StageEvent ev;
	stage_begin(c, 1, ev);
	kern<<<blocks, 128, smem, c->stream>>>(sc, c->pb, depth, qsel, tune);
	stage_end(c, ev);
	CK(cudaGetLastError());
	c->launches++; c->extend_launches++;
	if (c->opts.traversal == CUPT_TRAVERSAL_WIDE_STRICT) {
		if (c->retrace_capacity < c->path_capacity) {   // one slot index per path at worst; in practice a handful per launch
			if (c->pb.retrace) { CK(cudaFree(c->pb.retrace)); c->pb.retrace = nullptr; c->retrace_capacity = 0; }
			CK(cudaMalloc((void**)&c->pb.retrace, c->path_capacity * sizeof(uint32_t)));
			c->retrace_capacity = c->path_capacity;
		}
		StageEvent ev2;
		stage_begin(c, 1, ev2);
		CK(cudaMemsetAsync(c->pb.counters + 126, 0, sizeof(uint32_t), c->stream));
		k_verify_strict<<<c->num_sms * 8, 256, 0, c->stream>>>(sc, c->pb, depth, qsel);
		const size_t rt_smem = (size_t)sc.stack_cap * 64 * sizeof(uint2);
		if (rt_smem > 48 * 1024) CK(cudaFuncSetAttribute(k_retrace_strict, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rt_smem));
		k_retrace_strict<<<c->num_sms * 4, 64, rt_smem, c->stream>>>(sc, c->pb);
		stage_end(c, ev2);
		CK(cudaGetLastError());
		c->launches += 2;
		if (pp && c->opts.record_retraces && c->d_retrace_counts) {
			k_record_retraces<<<c->num_sms, 256, 0, c->stream>>>(*pp, c->pb, c->d_retrace_counts, slots_per_path);
			CK(cudaGetLastError());
			c->launches++;
		}
	}
	return CUPT_OK;
}

__global__ void k_add(float* __restrict__ dst, const float* __restrict__ src, size_t n)
{
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] += src[i];
}

}  // namespace

extern "C" {

int cupt_create(int device, cupt_context** out)
{
	if (!out) return set_error(CUPT_ERR_INVALID, "null output");
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n <= 0)
		return set_error(CUPT_ERR_NO_DEVICE, "no CUDA device (%s): the hot path has no CPU fallback", cudaGetErrorString(e));
	if (device < 0 || device >= n) return set_error(CUPT_ERR_INVALID, "device %d out of range (%d devices)", device, n);
	CK(cudaSetDevice(device));
	cupt_context* c = new (std::nothrow) cupt_context();
	if (!c) return set_error(CUPT_ERR_NOMEM, "out of memory");
	c->device = device;
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, device));
	c->num_sms = prop.multiProcessorCount;
	CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
	c->own_stream = true;
	CK(cudaMalloc((void**)&c->pb.counters, 128 * sizeof(uint32_t)));
	CK(cudaMalloc((void**)&c->pb.stats, 8 * sizeof(unsigned long long)));
	CK(cudaMemset(c->pb.counters, 0, 128 * sizeof(uint32_t)));
	CK(cudaMemset(c->pb.stats, 0, 8 * sizeof(unsigned long long)));
	cupt_default_render_opts(&c->opts);
	*out = c;
	return CUPT_OK;
}

void cupt_destroy(cupt_context* c)
{
	if (!c) return;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	drain_events(c);
	for (StageEvent& ev : c->pool) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
	cudaFree(c->d_verts); cudaFree(c->d_tris); cudaFree(c->d_trishade); cudaFree(c->d_mats); cudaFree(c->d_lights); cudaFree(c->d_light_cdf); cudaFree(c->d_spheres);
	cudaFree(c->d_nodes); cudaFree(c->d_prims); cudaFree(c->d_bvh2); cudaFree(c->d_bvh2_up); cudaFree(c->d_bvh2_leaf_parent);
	cudaFree(c->pb.rayA); cudaFree(c->pb.rayB); cudaFree(c->pb.st0); cudaFree(c->pb.st1);
	cudaFree(c->pb.queue[0]); cudaFree(c->pb.queue[1]); cudaFree(c->pb.counters); cudaFree(c->pb.stats); cudaFree(c->pb.retrace);
	cudaFree(c->d_accum); cudaFree(c->d_retrace_counts); cudaFree(c->d_build_rank); cudaFree(c->d_rgb8); cudaFree(c->d_prims_wt); c->arena.destroy();
	if (c->h_stage) cudaFreeHost(c->h_stage);
	if (c->stage_ev) cudaEventDestroy(c->stage_ev);
	if (c->own_stream) cudaStreamDestroy(c->stream);
	delete c;
}

int cupt_set_stream(cupt_context* c, void* cuda_stream)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	CK(cudaStreamSynchronize(c->stream));
	if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
	c->stream = (cudaStream_t)cuda_stream;
	return CUPT_OK;
}

int cupt_default_render_opts(cupt_render_opts* o)
{
	if (!o) return set_error(CUPT_ERR_INVALID, "null output");
	memset(o, 0, sizeof(*o));
	o->max_depth = 3;                  // NUM_BOUNCES, kernel.cu:33
	o->integrator = CUPT_INTEGRATOR_PT;
	o->traversal = CUPT_TRAVERSAL_WIDE;
	o->tri_test = CUPT_TRI_REFEXACT;
	o->seed = 1234;                    // kernel.cu:532
	return CUPT_OK;
}

int cupt_set_render_opts(cupt_context* c, const cupt_render_opts* o)
{
	if (!c || !o) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (o->max_depth < 1 || o->max_depth > 62) return set_error(CUPT_ERR_INVALID, "max_depth must be in 1..62");
	if (o->integrator != CUPT_INTEGRATOR_PT && o->integrator != CUPT_INTEGRATOR_BDPT)
		return set_error(CUPT_ERR_INVALID, "unknown integrator %d", o->integrator);
	if (o->traversal != CUPT_TRAVERSAL_WIDE && o->traversal != CUPT_TRAVERSAL_REF_BVH2 && o->traversal != CUPT_TRAVERSAL_WIDE_STRICT)
		return set_error(CUPT_ERR_INVALID, "bad traversal");
	if (o->tri_test != CUPT_TRI_REFEXACT && o->tri_test != CUPT_TRI_WATERTIGHT) return set_error(CUPT_ERR_INVALID, "unknown tri_test %d", o->tri_test);
	c->opts = *o;
	return CUPT_OK;
}

int cupt_set_camera(cupt_context* c, const cupt_camera* cam)
{
	if (!c || !cam) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (cam->pxlWidth < 1 || cam->pxlHeight < 1 || cam->pxlWidth > 65535 || cam->pxlHeight > 65535)
		return set_error(CUPT_ERR_INVALID, "image size %dx%d outside 1..65535 (pixel coordinates are uint16_t, camera.h:36)", cam->pxlWidth, cam->pxlHeight);
	// pixel and path-slot indices are 32-bit on the device (k_raygen, k_accumulate, k_tonemap_rgb8, k_heatmap)
	if ((int64_t)((cam->pxlWidth + 7) / 8) * ((cam->pxlHeight + 3) / 4) * 32 > 0x7fffffff)
		return set_error(CUPT_ERR_LIMIT, "image of %d x %d pixels exceeds 2^31 - 1 pixel slots (8x4-pixel tiles)", cam->pxlWidth, cam->pxlHeight);
	c->cam = *cam;
	c->have_cam = true;
	return CUPT_OK;
}

int cupt_set_scene(cupt_context* c, const cupt_scene* s)
{
	if (!c || !s) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (s->tris.empty()) return set_error(CUPT_ERR_INVALID, "scene has no triangles");
	CK(cudaSetDevice(c->device));
	const size_t nt = s->tris.size(), nm = s->mats.size(), nl = s->lights.size();
	int rc;
	if ((rc = upload(c, &c->d_verts, s->verts.data(), s->verts.size() * sizeof(cupt_vec3)))) return rc;
	if ((rc = upload(c, &c->d_tris, s->tris.data(), nt * sizeof(cupt_triangle)))) return rc;
	{
		std::vector<TriShade> tmp;
		TriShade* ts = (TriShade*)stage_alloc(c, nt * sizeof(TriShade));   // built in place in the staging buffer
		const bool staged = ts != nullptr;
		if (!staged) { tmp.resize(nt); ts = tmp.data(); }
		const cupt_triangle* tr = s->tris.data();
		// the same pass bounds the emissive triangles (k_shade's early stop for last segments, DevScene::emis_lo / emis_hi)
		float elo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, ehi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma omp parallel num_threads(copy_threads()) if (nt > 65536)
		{
			float tlo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, thi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma omp for schedule(static) nowait
			for (int64_t i = 0; i < (int64_t)nt; ++i) {
				ts[i].n[0] = tr[i].norm.x; ts[i].n[1] = tr[i].norm.y; ts[i].n[2] = tr[i].norm.z;
				ts[i].mat = tr[i].mat;
				if ((size_t)tr[i].mat < nm && s->mats[tr[i].mat].emission[0] != 0) {
					for (const int32_t vi : {tr[i].v0, tr[i].v1, tr[i].v2}) {
						const float pv[3] = {s->verts[vi].x, s->verts[vi].y, s->verts[vi].z};
						for (int a = 0; a < 3; ++a) { tlo[a] = std::min(tlo[a], pv[a]); thi[a] = std::max(thi[a], pv[a]); }
					}
				}
			}
#pragma omp critical
			for (int a = 0; a < 3; ++a) { elo[a] = std::min(elo[a], tlo[a]); ehi[a] = std::max(ehi[a], thi[a]); }
		}
		for (int a = 0; a < 3; ++a) { c->emis_lo[a] = elo[a]; c->emis_hi[a] = ehi[a]; }
		if ((rc = staged ? upload_staged(c, &c->d_trishade, ts, nt * sizeof(TriShade)) : upload(c, &c->d_trishade, ts, nt * sizeof(TriShade)))) return rc;
		if (!staged) CK(cudaStreamSynchronize(c->stream));
	}
	std::vector<MatF> mf(nm);
	for (size_t i = 0; i < nm; ++i) {
		for (int k = 0; k < 3; ++k) { mf[i].albedo[k] = (float)s->mats[i].albedo[k]; mf[i].emission[k] = (float)s->mats[i].emission[k]; }
		mf[i].emissive_flag = s->mats[i].emission[0] != 0 ? 1.f : 0.f;  // kernel.cu:453 tests the f64 value
		mf[i].specular_flag = (i < s->mat_specular.size() && s->mat_specular[i]) ? 1.f : 0.f;
	}
	// behind the table: f_s = albedo / 3.14159f per material in double, as the live integrator forms it for every vertex of every
	// connection (kernel.cu:367,379,391: `color` is f64, the division is by (double)3.14159f) — the same IEEE quotient, taken
	// once per material here instead of 24 times per sample in k_bdpt_connect.  One buffer, so that it travels with the table
	// wherever that goes (cupt_group's broadcast).
	std::vector<uint8_t> mbuf(nm * (sizeof(MatF) + 3 * sizeof(double)));
	if (nm) memcpy(mbuf.data(), mf.data(), nm * sizeof(MatF));
	for (size_t i = 0; i < nm; ++i)
		for (int k = 0; k < 3; ++k) {
			const double fs = (double)mf[i].albedo[k] / (double)3.14159f;
			memcpy(mbuf.data() + nm * sizeof(MatF) + (3 * i + k) * sizeof(double), &fs, sizeof(double));
		}
	if ((rc = upload(c, reinterpret_cast<uint8_t**>(&c->d_mats), mbuf.data(), mbuf.size()))) return rc;
	// light table: per light v0, a1 = v1 - v0, a2 = v2 - v0 and area = length(cross(a1, a2)) / 2 (kernel.cu:472-477),
	// plus the fallback entry for `selectedTri = 0` (kernel.cu:469)
	std::vector<LightTri> lt(nl + 1);
	for (size_t j = 0; j <= nl; ++j) {
		const cupt_triangle& t = s->tris[j < nl ? (size_t)s->lights[j] : 0];
		const V3 v0 = toV3(s->verts[t.v0]), a1 = toV3(s->verts[t.v1]) - v0, a2 = toV3(s->verts[t.v2]) - v0;
		memset(&lt[j], 0, sizeof(LightTri));
		lt[j].v0[0] = v0.x; lt[j].v0[1] = v0.y; lt[j].v0[2] = v0.z;
		lt[j].a1[0] = a1.x; lt[j].a1[1] = a1.y; lt[j].a1[2] = a1.z;
		lt[j].a2[0] = a2.x; lt[j].a2[1] = a2.y; lt[j].a2[2] = a2.z;
		lt[j].area = length3(cross3(a1, a2)) / 2;
		const int32_t tri_index = j < nl ? s->lights[j] : 0;   // read back by the BDPT stages (scene.lights[j], kernel.cu:242)
		memcpy(&lt[j].pad0, &tri_index, sizeof(int32_t));
	}
	if ((rc = upload(c, &c->d_lights, lt.data(), lt.size() * sizeof(LightTri)))) return rc;
	std::vector<float> cdf(nl);
	double run = 0;
	for (size_t j = 0; j < nl; ++j) { run += (double)lt[j].area; cdf[j] = (float)run; }
	if ((rc = upload(c, &c->d_light_cdf, cdf.data(), nl * sizeof(float)))) return rc;
	std::vector<SphereShade> ss(s->spheres.size());
	for (size_t i = 0; i < ss.size(); ++i) {
		const cupt_sphere& sp = s->spheres[i];
		ss[i].centre[0] = sp.pos.x; ss[i].centre[1] = sp.pos.y; ss[i].centre[2] = sp.pos.z; ss[i].radius = sp.rad;
		for (int k = 0; k < 3; ++k) { ss[i].diffuse[k] = (float)sp.diffuse[k]; ss[i].emm[k] = (float)sp.emm[k]; }
		ss[i].emissive_flag = sp.emm[0] != 0 ? 1.f : 0.f;
		ss[i].pad = 0.f;
		if (sp.emm[0] != 0) {
			const float cc[3] = {sp.pos.x, sp.pos.y, sp.pos.z};
			const float rr = std::fabs(sp.rad) * 1.001f;
			for (int a = 0; a < 3; ++a) { c->emis_lo[a] = std::min(c->emis_lo[a], cc[a] - rr); c->emis_hi[a] = std::max(c->emis_hi[a], cc[a] + rr); }
		}
	}
	if ((rc = upload(c, &c->d_spheres, ss.data(), ss.size() * sizeof(SphereShade)))) return rc;
	c->num_spheres = (int)ss.size();
	c->num_tris = (int64_t)nt; c->num_lights = (int)nl; c->total_light_area = s->totalLightArea;
	c->wt_valid = false;
	return uploads_done(c);   // (the temporaries above were copied to the staging buffer, or the pageable path synchronises here)
}

int cupt_set_bvh(cupt_context* c, const cupt_bvh* b)
{
	if (!c || !b) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (b->depth > 64) return set_error(CUPT_ERR_LIMIT, "wide BVH depth %d exceeds the traversal stack (64)", b->depth);
	if (b->prims.size() >= ((size_t)1 << 27))
		return set_error(CUPT_ERR_LIMIT, "more than 2^27 primitive records (the extend kernel packs record index and lane into 32 bits)");
	CK(cudaSetDevice(c->device));
	int rc;
	if ((rc = upload(c, &c->d_nodes, b->nodes.data(), b->nodes.size() * sizeof(Node8)))) return rc;
	if ((rc = upload(c, &c->d_prims, b->prims.data(), b->prims.size() * sizeof(Prim48)))) return rc;
	c->bvh_depth = b->depth; c->num_nodes = (int64_t)b->nodes.size(); c->num_prims = (int64_t)b->prims.size();
	c->wt_valid = false;
	return uploads_done(c);   // the caller may destroy the cupt_bvh right after this call
}

int cupt_build_bvh_device(cupt_context* c, const cupt_bvh_opts* opts, cupt_bvh_info* info)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	if (!c->d_tris && c->num_spheres == 0) return set_error(CUPT_ERR_INVALID, "cupt_set_scene has not been called");
	CK(cudaSetDevice(c->device));
	uint32_t* d_rank = nullptr;
	if (opts && opts->tie_rank && c->num_tris > 0) {
		int rcu;
		if ((rcu = upload(c, &c->d_build_rank, opts->tie_rank, (size_t)c->num_tris * sizeof(uint32_t)))) return rcu;
		d_rank = c->d_build_rank;
	}
	DeviceBvh r;
	r.nodes = c->d_nodes; r.prims = c->d_prims;
	r.nodes_capacity_bytes = c->d_nodes ? c->capacity[(void**)&c->d_nodes] : 0; r.prims_capacity_bytes = c->d_prims ? c->capacity[(void**)&c->d_prims] : 0;
	const int rc = build_wide_bvh_device(c->stream, c->arena, c->d_verts, c->d_tris, c->num_tris, c->d_spheres, c->num_spheres, d_rank,
	                                     (opts && opts->max_leaf_tris >= 1) ? opts->max_leaf_tris : 3, opts ? opts->ploc_radius : 0, &r);
	c->d_nodes = (uint4*)r.nodes; c->d_prims = (float4*)r.prims;   // the builder may have replaced them even when it failed later
	c->capacity[(void**)&c->d_nodes] = r.nodes_capacity_bytes; c->capacity[(void**)&c->d_prims] = r.prims_capacity_bytes;
	c->wt_valid = false;
	if (rc) { c->num_nodes = c->num_prims = 0; return rc; }
	c->bvh_depth = r.depth; c->num_nodes = r.num_nodes; c->num_prims = r.num_prims;
	if (info) {
		info->num_nodes = r.num_nodes; info->num_tri_records = r.num_prims;
		info->num_tris = c->num_tris; info->num_spheres = c->num_spheres;
		info->depth = r.depth; info->node_bytes = (int32_t)sizeof(Node8); info->tri_bytes = (int32_t)sizeof(Prim48);
		info->sah_cost = r.sah_cost; info->build_seconds = (double)r.ms_total * 1e-3;
	}
	c->build_ms[0] = r.ms_total; c->build_ms[1] = r.ms_sort; c->build_ms[2] = r.ms_ploc; c->build_ms[3] = r.ms_collapse; c->build_rounds = r.ploc_rounds;
	return CUPT_OK;
}

int cupt_read_bvh(cupt_context* c, void* nodes80, void* tris48, int64_t* counts)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	if (!c->d_nodes) return set_error(CUPT_ERR_INVALID, "no wide BVH on the device");
	CK(cudaSetDevice(c->device));
	if (counts) { counts[0] = c->num_nodes; counts[1] = c->num_prims; counts[2] = c->bvh_depth; counts[3] = c->build_rounds; }
	if (nodes80) CK(cudaMemcpyAsync(nodes80, c->d_nodes, (size_t)c->num_nodes * sizeof(Node8), cudaMemcpyDeviceToHost, c->stream));
	if (tris48) CK(cudaMemcpyAsync(tris48, c->d_prims, (size_t)c->num_prims * sizeof(Prim48), cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return CUPT_OK;
}

int cupt_debug_sort_pairs(cupt_context* c, uint64_t* keys, uint32_t* vals, int64_t n, int passes)
{
	if (!c || (n > 0 && (!keys || !vals))) return set_error(CUPT_ERR_INVALID, "bad arguments");
	CK(cudaSetDevice(c->device));
	return debug_sort_pairs(c->stream, keys, vals, n, passes);
}

int cupt_release_workspace(cupt_context* c)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	c->arena.destroy();
	return CUPT_OK;
}

int cupt_get_build_times(cupt_context* c, float* ms4)
{
	if (!c || !ms4) return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (int i = 0; i < 4; ++i) ms4[i] = c->build_ms[i];
	return CUPT_OK;
}

int cupt_set_bvh2(cupt_context* c, const cupt_bvh2_node* nodes, int size, int depth)
{
	if (!c || !nodes || size < 1) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (depth >= CUPT_MAX_BVH2_DEPTH)  // kernel.cu:627-631
		return set_error(CUPT_ERR_LIMIT, "BVH depth is too big (%d >= %d)", depth, CUPT_MAX_BVH2_DEPTH);
	CK(cudaSetDevice(c->device));
	int rc;
	if ((rc = upload(c, &c->d_bvh2, nodes, (size_t)size * sizeof(cupt_bvh2_node)))) return rc;
	// parent links for the strict mode's ancestor walk (the reference stores child links only, BVH.h:111-115)
	std::vector<uint32_t> parent((size_t)size, 0xffffffffu), leaf_parent;
	for (int i = 0; i < size; ++i)
		for (int k = 0; k < 2; ++k) {
			const uint32_t ch = k ? nodes[i].right : nodes[i].left;
			if (ch & CUPT_BVH_LEAF_FLAG) {
				const uint32_t tri = ch ^ CUPT_BVH_LEAF_FLAG;
				if (tri >= leaf_parent.size()) leaf_parent.resize((size_t)tri + 1, 0xffffffffu);
				leaf_parent[tri] = (uint32_t)i;
			} else if (ch < (uint32_t)size) parent[ch] = (uint32_t)i;
		}
	std::vector<cupt_bvh2_node> up(nodes, nodes + size);
	// chain[i] = 1 when every ancestor's box contains the box of its child on the way up from node i.  Then the
	// reference's slab test, whose comparisons are monotone in the box (IEEE rounding is monotone), accepts all of
	// them as soon as it accepts node i: the strict check needs ONE test instead of a walk to the root (k_verify_strict).
	std::vector<uint8_t> chain((size_t)size, 0);
	{
		std::vector<uint32_t> order;   // parents before children
		order.reserve((size_t)size);
		for (int i = 0; i < size; ++i) if (parent[(size_t)i] == 0xffffffffu) { chain[(size_t)i] = 1; order.push_back((uint32_t)i); }
		for (size_t q = 0; q < order.size(); ++q) {
			const cupt_bvh2_node& p = nodes[order[q]];
			for (int k = 0; k < 2; ++k) {
				const uint32_t ch = k ? p.right : p.left;
				if ((ch & CUPT_BVH_LEAF_FLAG) || ch >= (uint32_t)size || parent[ch] != order[q]) continue;
				const cupt_bvh2_node& n = nodes[ch];
				const bool nested = n.lo.x >= p.lo.x && n.lo.y >= p.lo.y && n.lo.z >= p.lo.z && n.hi.x <= p.hi.x && n.hi.y <= p.hi.y && n.hi.z <= p.hi.z;
				chain[ch] = (uint8_t)(nested && chain[order[q]]);
				order.push_back(ch);
			}
		}
	}
	for (int i = 0; i < size; ++i) { up[(size_t)i].left = parent[(size_t)i]; up[(size_t)i].right = chain[(size_t)i]; }
	if ((rc = upload(c, &c->d_bvh2_up, up.data(), up.size() * sizeof(cupt_bvh2_node)))) return rc;
	if ((rc = upload(c, &c->d_bvh2_leaf_parent, leaf_parent.data(), leaf_parent.size() * sizeof(uint32_t)))) return rc;
	c->bvh2_leaves = (int64_t)leaf_parent.size();
	c->bvh2_size = size; c->bvh2_depth = depth;
	CK(cudaStreamSynchronize(c->stream));
	return CUPT_OK;
}

int cupt_render(cupt_context* c, int first_sample, int num_samples)
{
	int rc = check_ready(c, true);
	if (rc) return rc;
	if (num_samples < 0 || first_sample < 0) return set_error(CUPT_ERR_INVALID, "bad sample range");
	CK(cudaSetDevice(c->device));
	if ((rc = ensure_accum(c))) return rc;
	const DevScene sc = make_scene(c);
	PassParams probe = make_pass(c, 0, 1, 0, true);
	if (c->tile_count > 0 && c->tile_first >= probe.tiles_x * probe.tiles_y) return set_error(CUPT_ERR_INVALID, "tile window starts beyond the image (%d tiles)", probe.tiles_x * probe.tiles_y);
	const bool bdpt = c->opts.integrator == CUPT_INTEGRATOR_BDPT;
	if (bdpt && c->num_spheres > 0) return set_error(CUPT_ERR_INVALID, "the BDPT integrator (kernel.cu:217-415) has no sphere support");
	int S = c->opts.samples_per_pass;
	// PT: 64 Mi paths x 72 B = 4.8 GB of path state: long passes amortise the tail of every extend launch;
	// BDPT keeps four ray records and four connection records per path (288 B): 16 Mi paths
	if (S <= 0) S = (int)std::max<int64_t>(1, (int64_t)((bdpt ? 16 : 64) << 20) / probe.p_pad);
	S = std::min(S, std::max(1, num_samples));
	const int64_t per_path = bdpt ? 4 : 1;
	// slot indices are 32-bit: a pass holds at most 2^31 - 1 ray records.  An image too large for even one sample per pass
	// (about 23 000 x 23 000 pixels with BDPT's four records per path) is refused rather than rendered with S = 0 passes.
	if ((int64_t)probe.p_pad * per_path > 0x7fffffff)
		return set_error(CUPT_ERR_LIMIT, "image of %d x %d pixels needs more than 2^31 - 1 ray records per sample (render it in tile windows: cupt_set_tile_window)", c->cam.pxlWidth, c->cam.pxlHeight);
	if ((int64_t)probe.p_pad * S * per_path > 0x7fffffff) S = (int)(0x7fffffff / ((int64_t)probe.p_pad * per_path));
	if ((rc = ensure_retrace_counts(c))) return rc;
	if ((rc = ensure_paths(c, (size_t)probe.p_pad * S * per_path))) return rc;
	const int grid = c->num_sms * 8;
	for (int done = 0; done < num_samples; done += S) {
		const int s_now = std::min(S, num_samples - done);
		const PassParams pp = make_pass(c, first_sample + done, s_now, 0, true);
		CK(cudaMemsetAsync(c->pb.counters, 0, 128 * sizeof(uint32_t), c->stream));
		StageEvent ev;
		if (bdpt) {
			// radianceAlongSingleStep, kernel.cu:217-415, as three extend rounds (bdpt.cuh)
			stage_begin(c, 0, ev);
			k_bdpt_gen<<<grid, 256, 0, c->stream>>>(sc, pp, c->pb);
			stage_end(c, ev);
			if ((rc = launch_extend(c, sc, 0, 0, false, &pp, 4))) return rc;
			stage_begin(c, 2, ev);
			k_bdpt_mid<<<grid, 256, 0, c->stream>>>(sc, pp, c->pb);
			stage_end(c, ev);
			if ((rc = launch_extend(c, sc, 1, 1, false, &pp, 4))) return rc;
			stage_begin(c, 2, ev);
			{
				const size_t nm = c->used[(void**)&c->d_mats] / (sizeof(MatF) + 3 * sizeof(double));   // (the f_s table lies behind the material table)
				k_bdpt_connect<<<grid, 256, 0, c->stream>>>(sc, pp, c->pb, reinterpret_cast<const double*>(reinterpret_cast<const uint8_t*>(c->d_mats) + nm * sizeof(MatF)));
			}
			stage_end(c, ev);
			if ((rc = launch_extend(c, sc, 2, 0, true, &pp, 4))) return rc;   // visibility rays: t limit + early stop
			stage_begin(c, 3, ev);
			k_bdpt_accumulate<<<(pp.p_pad + 255) / 256, 256, 0, c->stream>>>(pp, c->pb, c->d_accum);
			stage_end(c, ev);
			c->launches += 4;
			CK(cudaGetLastError());
			c->samples_accumulated += (uint64_t)s_now;
			continue;
		}
		stage_begin(c, 0, ev);
		// pinhole camera: the primary ray does not depend on the sample, so one trace per pixel serves the whole pass.
		// Wide traversal only: the lens offset is radius * cos(theta) = +-0, and where the film offset is 0 too (the
		// centre column / row) the ray direction comes out as +0 or -0 depending on the sample.  The wide traversal
		// returns the same hit for both; the reference's slab test divides by it (BVH.h:61-79) and does not, and the
		// reference-exact modes reproduce that.
		const bool share = c->opts.share_primary_hits && c->cam.radius == 0.f && s_now > 1 && c->opts.traversal == CUPT_TRAVERSAL_WIDE;
		k_raygen<<<grid, 256, 0, c->stream>>>(pp, c->pb, share);
		stage_end(c, ev);
		c->launches++;
		for (int depth = 0; depth < pp.max_depth; ++depth) {
			if (depth == 0 && share) {
				if ((rc = launch_extend(c, sc, 63, 1))) return rc;       // queue[1] / counters[63]: the first sample's slots
				stage_begin(c, 1, ev);
				k_broadcast_primary<<<grid, 256, 0, c->stream>>>(pp, c->pb);
				stage_end(c, ev);
				c->launches++;
			} else if ((rc = launch_extend(c, sc, depth, depth & 1, depth > 0 && pp.term_stop, &pp))) return rc;
			stage_begin(c, 2, ev);
			k_shade<<<grid, 256, 0, c->stream>>>(sc, pp, c->pb, depth, depth & 1);
			stage_end(c, ev);
			c->launches++;
		}
		stage_begin(c, 3, ev);
		k_accumulate<<<(pp.p_pad + 255) / 256, 256, 0, c->stream>>>(pp, c->pb, c->d_accum);
		stage_end(c, ev);
		c->launches++;
		CK(cudaGetLastError());
		c->samples_accumulated += (uint64_t)s_now;
	}
	return CUPT_OK;
}

// Image-tile partition (the alternative to sample ranges): cupt_render then traces only the pixels of tiles
// [first_tile, first_tile + num_tiles) — tiles of 8x4 pixels, numbered row-major, the order raygen lays warps out in —
// and leaves the rest of the accumulator untouched, so that the ranks' images are disjoint and their sum (or a gather)
// is the frame.  num_tiles <= 0 restores the whole image.
int cupt_set_tile_window(cupt_context* c, int first_tile, int num_tiles)
{
	if (!c || first_tile < 0) return set_error(CUPT_ERR_INVALID, "bad tile window");
	c->tile_first = num_tiles > 0 ? first_tile : 0;
	c->tile_count = num_tiles > 0 ? num_tiles : 0;
	return CUPT_OK;
}

int cupt_tile_count(const cupt_context* c, int* num_tiles)
{
	if (!c || !num_tiles || !c->have_cam) return set_error(CUPT_ERR_INVALID, "no camera set");
	*num_tiles = ((c->cam.pxlWidth + 7) / 8) * ((c->cam.pxlHeight + 3) / 4);
	return CUPT_OK;
}

int cupt_sync(cupt_context* c)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	drain_events(c);
	return CUPT_OK;
}

int cupt_reset_accum(cupt_context* c)
{
	if (!c || !c->have_cam) return set_error(CUPT_ERR_INVALID, "no camera set");
	CK(cudaSetDevice(c->device));
	int rc = ensure_accum(c);
	if (rc) return rc;
	CK(cudaMemsetAsync(c->d_accum, 0, c->accum_pixels * 3 * sizeof(float), c->stream));
	if (c->d_retrace_counts) CK(cudaMemsetAsync(c->d_retrace_counts, 0, c->retrace_pixels * sizeof(uint32_t), c->stream));
	c->samples_accumulated = 0;
	return CUPT_OK;
}

int cupt_read_retrace_counts(cupt_context* c, uint32_t* counts)
{
	if (!c || !counts || !c->have_cam) return set_error(CUPT_ERR_INVALID, "bad arguments");
	const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
	if (!c->d_retrace_counts || c->retrace_pixels != P)
		return set_error(CUPT_ERR_INVALID, "no re-trace counts: render in CUPT_TRAVERSAL_WIDE_STRICT with record_retraces = 1 first");
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(counts, c->d_retrace_counts, P * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return CUPT_OK;
}

int cupt_read_accum(cupt_context* c, float* rgb_host, int normalise)
{
	if (!c || !rgb_host || !c->d_accum) return set_error(CUPT_ERR_INVALID, "nothing rendered");
	CK(cudaSetDevice(c->device));
	const size_t n = c->accum_pixels * 3;
	CK(cudaMemcpyAsync(rgb_host, c->d_accum, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));   // (staging this read-back through pinned memory was measured: no gain)
	CK(cudaStreamSynchronize(c->stream));
	drain_events(c);
	if (normalise && c->samples_accumulated > 0) {
		const float inv = 1.0f / (float)c->samples_accumulated;
		for (size_t i = 0; i < n; ++i) rgb_host[i] *= inv;
	}
	return CUPT_OK;
}

int cupt_copy_accum_device(cupt_context* c, float* dst_device)
{
	if (!c || !dst_device || !c->d_accum) return set_error(CUPT_ERR_INVALID, "nothing rendered");
	CK(cudaSetDevice(c->device));
	CK(cudaMemcpyAsync(dst_device, c->d_accum, c->accum_pixels * 3 * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
	return CUPT_OK;
}

int cupt_add_accum_device(cupt_context* c, const float* src_device, int samples)
{
	if (!c || !src_device || samples < 0) return set_error(CUPT_ERR_INVALID, "bad arguments");
	int rc = check_ready(c, false);
	if (rc) return rc;
	CK(cudaSetDevice(c->device));
	if ((rc = ensure_accum(c))) return rc;
	const size_t n = c->accum_pixels * 3;
	k_add<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d_accum, src_device, n);
	CK(cudaGetLastError());
	c->launches++;
	c->samples_accumulated += (uint64_t)samples;
	return CUPT_OK;
}

static int primary_pass(cupt_context* c, int32_t* tri_id, float* t, float* rays6)
{
	int rc = check_ready(c, false);
	if (rc) return rc;
	CK(cudaSetDevice(c->device));
	const PassParams pp = make_pass(c, 0, 1, 1);
	if ((rc = ensure_paths(c, (size_t)pp.p_pad))) return rc;
	const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
	int32_t* d_ids = nullptr; float* d_t = nullptr; float* d_r = nullptr;
	if (tri_id) { CK(cudaMalloc((void**)&d_ids, P * sizeof(int32_t))); CK(cudaMalloc((void**)&d_t, P * sizeof(float))); }
	if (rays6) CK(cudaMalloc((void**)&d_r, P * 6 * sizeof(float)));
	CK(cudaMemsetAsync(c->pb.counters, 0, 128 * sizeof(uint32_t), c->stream));
	k_raygen<<<c->num_sms * 8, 256, 0, c->stream>>>(pp, c->pb);
	c->launches++;
	if ((rc = ensure_retrace_counts(c))) return rc;
	if (tri_id) { const DevScene sc = make_scene(c); if ((rc = launch_extend(c, sc, 0, 0, false, &pp))) return rc; }
	k_gather_primary<<<(pp.p_pad + 255) / 256, 256, 0, c->stream>>>(pp, c->pb, d_ids, d_t, d_r);
	c->launches++;
	CK(cudaGetLastError());
	if (tri_id) {
		CK(cudaMemcpyAsync(tri_id, d_ids, P * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
		CK(cudaMemcpyAsync(t, d_t, P * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	}
	if (rays6) CK(cudaMemcpyAsync(rays6, d_r, P * 6 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	drain_events(c);
	cudaFree(d_ids); cudaFree(d_t); cudaFree(d_r);
	return CUPT_OK;
}

int cupt_read_primary_hits(cupt_context* c, int32_t* tri_id, float* t)
{
	if (!tri_id || !t) return set_error(CUPT_ERR_INVALID, "null output");
	return primary_pass(c, tri_id, t, nullptr);
}

int cupt_read_image_rgb8(cupt_context* c, uint8_t* rgb, int mirror_x)
{
	if (!c || !rgb) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (!c->d_accum || c->samples_accumulated == 0) return set_error(CUPT_ERR_INVALID, "nothing has been rendered");
	CK(cudaSetDevice(c->device));
	const int W = c->cam.pxlWidth, H = c->cam.pxlHeight;
	const size_t bytes = (size_t)W * H * 3;
	int rc;
	if (c->capacity[(void**)&c->d_rgb8] < bytes) {
		if (c->d_rgb8) { CK(cudaFree(c->d_rgb8)); c->d_rgb8 = nullptr; }
		CK(cudaMalloc((void**)&c->d_rgb8, bytes));
		c->capacity[(void**)&c->d_rgb8] = bytes;
	}
	(void)rc;
	k_tonemap_rgb8<<<(W * H + 255) / 256, 256, 0, c->stream>>>(c->d_accum, 1.0f / (float)c->samples_accumulated, W, H, mirror_x, c->d_rgb8);
	c->launches++;
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(rgb, c->d_rgb8, bytes, cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	return CUPT_OK;
}

int cupt_read_heatmap(cupt_context* c, float* nodes_visited, float* prims_tested)
{
	if (!c || (!nodes_visited && !prims_tested)) return set_error(CUPT_ERR_INVALID, "bad arguments");
	int rc;
	const int saved = c->opts.traversal;
	c->opts.traversal = CUPT_TRAVERSAL_WIDE;           // the map is of the wide BVH, whatever mode renders
	rc = check_ready(c, false);
	c->opts.traversal = saved;
	if (rc) return rc;
	CK(cudaSetDevice(c->device));
	const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
	float* d = nullptr;
	CK(cudaMalloc((void**)&d, 2 * P * sizeof(float)));
	const DevScene sc = make_scene(c);
	const PassParams pp = make_pass(c, 0, 1, 1);
	const size_t smem = (size_t)sc.stack_cap * 128 * sizeof(uint2);
	CK(cudaFuncSetAttribute(k_heatmap, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	k_heatmap<<<(unsigned)((P + 127) / 128), 128, smem, c->stream>>>(sc, pp, d, d + P);
	c->launches++;
	CK(cudaGetLastError());
	if (nodes_visited) CK(cudaMemcpyAsync(nodes_visited, d, P * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	if (prims_tested) CK(cudaMemcpyAsync(prims_tested, d + P, P * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
	CK(cudaStreamSynchronize(c->stream));
	cudaFree(d);
	return CUPT_OK;
}

int cupt_camera_rays(cupt_context* c, float* rays_host)
{
	if (!c || !rays_host) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (!c->have_cam) return set_error(CUPT_ERR_INVALID, "no camera set");
	// raygen alone needs no acceleration structure
	const int saved = c->opts.traversal;
	int rc;
	{
		CK(cudaSetDevice(c->device));
		const PassParams pp = make_pass(c, 0, 1, 1);
		if ((rc = ensure_paths(c, (size_t)pp.p_pad))) return rc;
		const size_t P = (size_t)c->cam.pxlWidth * c->cam.pxlHeight;
		float* d_r = nullptr;
		CK(cudaMalloc((void**)&d_r, P * 6 * sizeof(float)));
		CK(cudaMemsetAsync(c->pb.counters, 0, 128 * sizeof(uint32_t), c->stream));
		k_raygen<<<c->num_sms * 8, 256, 0, c->stream>>>(pp, c->pb);
		k_gather_primary<<<(pp.p_pad + 255) / 256, 256, 0, c->stream>>>(pp, c->pb, nullptr, nullptr, d_r);
		c->launches += 2;
		CK(cudaGetLastError());
		CK(cudaMemcpyAsync(rays_host, d_r, P * 6 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
		CK(cudaStreamSynchronize(c->stream));
		cudaFree(d_r);
	}
	(void)saved;
	return CUPT_OK;
}

int cupt_trace_rays(cupt_context* c, const float* rays_host, int64_t n, int32_t* tri_id, float* t)
{
	if (!c || !rays_host || !tri_id || !t || n < 0) return set_error(CUPT_ERR_INVALID, "bad arguments");
	{ const int rc0 = check_ready(c, false, false); if (rc0) return rc0; }
	CK(cudaSetDevice(c->device));
	const int64_t chunk = 8 << 20;
	int rc;
	if ((rc = ensure_paths(c, (size_t)std::min<int64_t>(std::max<int64_t>(n, 1), chunk)))) return rc;
	float* d_r = nullptr; int32_t* d_ids = nullptr; float* d_t = nullptr;
	const size_t cap = (size_t)std::min<int64_t>(std::max<int64_t>(n, 1), chunk);
	CK(cudaMalloc((void**)&d_r, cap * 6 * sizeof(float)));
	CK(cudaMalloc((void**)&d_ids, cap * sizeof(int32_t)));
	CK(cudaMalloc((void**)&d_t, cap * sizeof(float)));
	const DevScene sc = make_scene(c);
	for (int64_t off = 0; off < n; off += chunk) {
		const uint32_t m = (uint32_t)std::min<int64_t>(chunk, n - off);
		CK(cudaMemcpyAsync(d_r, rays_host + 6 * off, (size_t)m * 6 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
		CK(cudaMemsetAsync(c->pb.counters, 0, 128 * sizeof(uint32_t), c->stream));
		CK(cudaMemcpyAsync(c->pb.counters, &m, sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
		k_load_rays<<<(m + 255) / 256, 256, 0, c->stream>>>(d_r, m, 0u, c->pb);
		c->launches++;
		if ((rc = launch_extend(c, sc, 0, 0))) return rc;
		k_store_hits<<<(m + 255) / 256, 256, 0, c->stream>>>(m, 0u, c->pb, d_ids, d_t);
		c->launches++;
		CK(cudaGetLastError());
		CK(cudaMemcpyAsync(tri_id + off, d_ids, (size_t)m * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
		CK(cudaMemcpyAsync(t + off, d_t, (size_t)m * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
		CK(cudaStreamSynchronize(c->stream));
	}
	drain_events(c);
	cudaFree(d_r); cudaFree(d_ids); cudaFree(d_t);
	return CUPT_OK;
}

int cupt_get_stats(cupt_context* c, cupt_stats* out)
{
	if (!c || !out) return set_error(CUPT_ERR_INVALID, "bad arguments");
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	drain_events(c);
	unsigned long long h[8];
	CK(cudaMemcpy(h, c->pb.stats, sizeof(h), cudaMemcpyDeviceToHost));
	memset(out, 0, sizeof(*out));
	out->rays = h[0]; out->nodes_visited = h[1]; out->tris_tested = h[2]; out->samples = h[3]; out->strict_retraces = h[4];
	out->kernel_launches = c->launches; out->extend_launches = c->extend_launches;
	out->ms_raygen = c->ms[0]; out->ms_extend = c->ms[1]; out->ms_shade = c->ms[2]; out->ms_accum = c->ms[3];
	return CUPT_OK;
}

int cupt_reset_stats(cupt_context* c)
{
	if (!c) return set_error(CUPT_ERR_INVALID, "null context");
	CK(cudaSetDevice(c->device));
	CK(cudaStreamSynchronize(c->stream));
	drain_events(c);
	CK(cudaMemsetAsync(c->pb.stats, 0, 8 * sizeof(unsigned long long), c->stream));
	c->launches = c->extend_launches = 0;
	for (int i = 0; i < 4; ++i) c->ms[i] = 0;
	return CUPT_OK;
}

}  // extern "C"

// ---- closest hits for caller rays WITHOUT any allocation (cupt_trace_rays_noalloc) ----------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  cupt_trace_rays allocates its ray / hit staging
// and, on first use, the path records and queues; this variant keeps all of that scratch in ONE file-scope device array of a
// fixed size, which every GPU that loads the library holds until the process exits:
//   rays6 [N x 24 B] | rayA [N x 16 B] | rayB [N x 16 B] | queue [N x 4 B] | ids [N x 4 B] | ts [N x 4 B] | counters [128 u32] | stats [8 u64]
// with N = CUPT_NOALLOC_MAX_RAYS = 65 536: 4 457 472 + 576 bytes.  In HOST code the array's name is only a stand-in for the
// symbol, not a device address, so the address is fetched once per device with cudaGetSymbolAddress (device current, return
// code checked) and that pointer is what the kernels get; device code (k_noalloc_begin) names the array directly.
// Every call shares the array, which keeps the last call's contents: the counters are cleared at the start of each call and
// nothing else is read before it is written.  Two calls that overlap on the GPU (two contexts / streams on one device) race.
#define CUPT_NOALLOC_OFF_RAYS6 ((size_t)0)
#define CUPT_NOALLOC_OFF_RAYA (CUPT_NOALLOC_OFF_RAYS6 + (size_t)CUPT_NOALLOC_MAX_RAYS * 24)
#define CUPT_NOALLOC_OFF_RAYB (CUPT_NOALLOC_OFF_RAYA + (size_t)CUPT_NOALLOC_MAX_RAYS * 16)
#define CUPT_NOALLOC_OFF_QUEUE (CUPT_NOALLOC_OFF_RAYB + (size_t)CUPT_NOALLOC_MAX_RAYS * 16)
#define CUPT_NOALLOC_OFF_IDS (CUPT_NOALLOC_OFF_QUEUE + (size_t)CUPT_NOALLOC_MAX_RAYS * 4)
#define CUPT_NOALLOC_OFF_TS (CUPT_NOALLOC_OFF_IDS + (size_t)CUPT_NOALLOC_MAX_RAYS * 4)
#define CUPT_NOALLOC_OFF_COUNTERS (CUPT_NOALLOC_OFF_TS + (size_t)CUPT_NOALLOC_MAX_RAYS * 4)
#define CUPT_NOALLOC_OFF_STATS (CUPT_NOALLOC_OFF_COUNTERS + 128 * sizeof(uint32_t))
#define CUPT_NOALLOC_BYTES (CUPT_NOALLOC_OFF_STATS + 8 * sizeof(unsigned long long))
static __device__ __align__(256) unsigned char g_noalloc_scratch[CUPT_NOALLOC_BYTES];

// queue sizes / fetch cursors of this call: counters[0] = n, everything else 0 (the array keeps the previous call's values)
static __global__ void __launch_bounds__(128) k_noalloc_begin(uint32_t n)
{
	uint32_t* counters = reinterpret_cast<uint32_t*>(g_noalloc_scratch + CUPT_NOALLOC_OFF_COUNTERS);
	counters[threadIdx.x] = threadIdx.x == 0u ? n : 0u;
	if (threadIdx.x < 8u) reinterpret_cast<unsigned long long*>(g_noalloc_scratch + CUPT_NOALLOC_OFF_STATS)[threadIdx.x] = 0ull;
}

namespace {
struct NoAllocDevice {
	unsigned char* base = nullptr;      // device address of g_noalloc_scratch on this device (cudaGetSymbolAddress)
	int blocks[2] = {0, 0};             // persistent grid of k_extend<mode, false, false>
	size_t smem[2] = {0, 0};            // the dynamic shared memory the grid was sized for
};
constexpr int kNoAllocMaxDevices = 64;
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
}  // namespace

extern "C" {

int cupt_trace_rays_noalloc(cupt_context* c, const float* rays, int64_t n, int32_t* tri_id, float* t, int synchronise)
{
	if (!c || n < 0 || (n > 0 && (!rays || !tri_id || !t))) return set_error(CUPT_ERR_INVALID, "bad arguments");
	if (n > CUPT_NOALLOC_MAX_RAYS)
		return set_error(CUPT_ERR_LIMIT, "cupt_trace_rays_noalloc: %lld rays exceed the fixed scratch of CUPT_NOALLOC_MAX_RAYS = %d rays per call "
		                 "(split the batch, or use cupt_trace_rays, which allocates)", (long long)n, CUPT_NOALLOC_MAX_RAYS);
	if (c->opts.traversal == CUPT_TRAVERSAL_WIDE_STRICT)
		return set_error(CUPT_ERR_INVALID, "cupt_trace_rays_noalloc: the strict mode needs a re-trace list that is allocated on demand; use CUPT_TRAVERSAL_WIDE or CUPT_TRAVERSAL_REF_BVH2");
	if (c->opts.tri_test != CUPT_TRI_REFEXACT)
		return set_error(CUPT_ERR_INVALID, "cupt_trace_rays_noalloc: CUPT_TRI_WATERTIGHT builds its vertex records on demand; use CUPT_TRI_REFEXACT");
	{ const int rc0 = check_ready(c, false, false); if (rc0) return rc0; }   // (allocates nothing with the reference-exact triangle test)
	if (c->device >= kNoAllocMaxDevices) return set_error(CUPT_ERR_LIMIT, "device index %d beyond %d", c->device, kNoAllocMaxDevices);
	CK(cudaSetDevice(c->device));
	NoAllocDevice& nd = g_noalloc_dev[c->device];
	if (!nd.base) {   // once per device, with that device current
		unsigned char* ptr = nullptr;
		CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
		if (!ptr) return set_error(CUPT_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the scratch array");
		nd.base = ptr;
	}
	if (n == 0) return CUPT_OK;
	const uint32_t m = (uint32_t)n;
	PathBuffers pb{};
	pb.rayA = reinterpret_cast<float4*>(nd.base + CUPT_NOALLOC_OFF_RAYA);
	pb.rayB = reinterpret_cast<float4*>(nd.base + CUPT_NOALLOC_OFF_RAYB);
	pb.queue[0] = pb.queue[1] = reinterpret_cast<uint32_t*>(nd.base + CUPT_NOALLOC_OFF_QUEUE);
	pb.counters = reinterpret_cast<uint32_t*>(nd.base + CUPT_NOALLOC_OFF_COUNTERS);
	pb.stats = reinterpret_cast<unsigned long long*>(nd.base + CUPT_NOALLOC_OFF_STATS);
	float* d_rays = reinterpret_cast<float*>(nd.base + CUPT_NOALLOC_OFF_RAYS6);
	int32_t* d_ids = reinterpret_cast<int32_t*>(nd.base + CUPT_NOALLOC_OFF_IDS);
	float* d_t = reinterpret_cast<float*>(nd.base + CUPT_NOALLOC_OFF_TS);
	const DevScene sc = make_scene(c);
	const int mode = c->opts.traversal == CUPT_TRAVERSAL_REF_BVH2 ? 1 : 0;
	const size_t smem = mode == 0 ? (size_t)sc.stack_cap * 128 * sizeof(uint2) + 128 * 16 : 0;
	void (*kern)(DevScene, PathBuffers, int, int, ExtendTuning) = mode == 0 ? k_extend<0, false, false> : k_extend<1, false, false>;
	if (nd.blocks[mode] == 0 || nd.smem[mode] != smem) {
		if (smem > 48 * 1024) CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		int per_sm = 0;
		CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
		if (per_sm < 1) return set_error(CUPT_ERR_LIMIT, "the extend kernel does not fit an SM with a traversal stack of %d entries", sc.stack_cap);
		nd.blocks[mode] = per_sm * c->num_sms; nd.smem[mode] = smem;
	}
	ExtendTuning tune;
	tune.fetch_thresh = c->opts.fetch_threshold > 0 ? std::min(c->opts.fetch_threshold, 31) : 24;
	tune.defer_pairs = c->opts.defer_pairs > 0 ? std::min(c->opts.defer_pairs, 64) : 20;
	tune.qbias_bits = CUPT_QBIAS_BITS;
	// the caller's pointers may be host or device memory (unified addressing decides); device pointers involve no host staging
	CK(cudaMemcpyAsync(d_rays, rays, (size_t)m * 6 * sizeof(float), cudaMemcpyDefault, c->stream));
	k_noalloc_begin<<<1, 128, 0, c->stream>>>(m);
	k_load_rays<<<(m + 255) / 256, 256, 0, c->stream>>>(d_rays, m, 0u, pb);
	kern<<<nd.blocks[mode], 128, smem, c->stream>>>(sc, pb, 0, 0, tune);
	k_store_hits<<<(m + 255) / 256, 256, 0, c->stream>>>(m, 0u, pb, d_ids, d_t);
	CK(cudaGetLastError());
	c->launches += 4; c->extend_launches++;
	CK(cudaMemcpyAsync(tri_id, d_ids, (size_t)m * sizeof(int32_t), cudaMemcpyDefault, c->stream));
	CK(cudaMemcpyAsync(t, d_t, (size_t)m * sizeof(float), cudaMemcpyDefault, c->stream));
	if (synchronise) CK(cudaStreamSynchronize(c->stream));
	return CUPT_OK;
}

}  // extern "C"

#include "group.cuh"
