// cupt_group — several GPUs behind one handle (include/cupt.h "multi-GPU"), SURVEY.md §8(b) "Threading": one context per
// GPU, one host thread per context, an NCCL communicator owned by the group.  The reference renders on device 0 only
// (kernel.cu:569, render loop kernel.cu:709-736); what a group replaces is that loop for N devices:
//   * the scene and the BVH are staged through host memory ONCE, on the member with rank 0, and reach the other members by
//     ncclBroadcast over NVLink (instead of N host stagings of the same 28 MB);
//   * cupt_group_render gives member r the sample range sharding.sample_range(r, N, samples) of every pixel (Philox is
//     keyed by the global sample index, so the ranges are disjoint by construction) and sums the f32 accumulators onto
//     rank 0 with ONE ncclReduce — the path's only exchange step; the other members' accumulators are cleared afterwards,
//     so rank 0 holds the running total of a progressive render;
//   * the image is read from rank 0 (f32 sums, or tonemapped to 3 bytes per pixel on the device).
// A group's members may be threads of one process (cupt_group_create_local: ncclCommInitAll) or one member per process
// (cupt_group_create_rank: ncclCommInitRank with an id the caller distributes, e.g. under torchrun / MPI).
// NCCL is dlopen()ed when the first group is made, so a process that never makes one neither needs nor loads it, and a
// process that already holds an NCCL (PyTorch's) shares that copy.
//
// This file is part of context.cu's translation unit (it needs the context's internals).
#pragma once
#include <nccl.h>
#include <dlfcn.h>
#include <condition_variable>
#include <functional>
#include <mutex>

namespace {

struct NcclApi {
	void* lib = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int nccl_load()
{
	std::lock_guard<std::mutex> lk(g_nccl_mu);
	if (g_nccl.lib) return CUPT_OK;
	void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
	if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
	if (!h) return set_error(CUPT_ERR_INVALID, "cupt_group needs NCCL: %s", dlerror());
#define CUPT_NCCL_SYM(field, name)                                                                              \
	if (!(*(void**)(&g_nccl.field) = dlsym(h, name))) return set_error(CUPT_ERR_INVALID, "NCCL symbol %s not found", name)
	CUPT_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
	CUPT_NCCL_SYM(CommInitRank, "ncclCommInitRank");
	CUPT_NCCL_SYM(CommInitAll, "ncclCommInitAll");
	CUPT_NCCL_SYM(CommDestroy, "ncclCommDestroy");
	CUPT_NCCL_SYM(Reduce, "ncclReduce");
	CUPT_NCCL_SYM(Broadcast, "ncclBroadcast");
	CUPT_NCCL_SYM(GroupStart, "ncclGroupStart");
	CUPT_NCCL_SYM(GroupEnd, "ncclGroupEnd");
	CUPT_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef CUPT_NCCL_SYM
	g_nccl.lib = h;
	return CUPT_OK;
}

#define NK(call)                                                                                             \
	do {                                                                                                     \
		ncclResult_t r_ = (call);                                                                            \
		if (r_ != ncclSuccess)                                                                               \
			return set_error(CUPT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
	} while (0)

// what a non-root member has to know before it can receive the arrays: their sizes and the context's scalar fields
struct GroupHeader {
	uint64_t bytes[12];
	int64_t num_tris, num_nodes, num_prims, bvh2_leaves;
	int32_t num_lights, num_spheres, bvh_depth, bvh2_size, bvh2_depth, pad;
	float total_light_area, emis_lo[3], emis_hi[3];
};

struct GroupMember {
	cupt_context* ctx = nullptr;
	ncclComm_t comm = nullptr;
	int rank = 0;
	GroupHeader* d_hdr = nullptr;     // device scratch for the header broadcast
	GroupHeader* h_hdr = nullptr;     // pinned
	std::thread th;
	int rc = 0;
	std::string err;
};

}  // namespace

struct cupt_group {
	int nranks = 0;                                  // members of the whole group (all processes)
	std::vector<GroupMember> local;                  // members this process drives
	std::mutex mu;
	std::condition_variable cv_job, cv_done;
	std::function<int(GroupMember&)> job;
	uint64_t job_id = 0;
	int pending = 0;
	bool stop = false;
};

namespace {

void group_worker(cupt_group* g, int index)
{
	GroupMember& m = g->local[(size_t)index];
	cudaSetDevice(m.ctx->device);
	uint64_t seen = 0;
	for (;;) {
		std::function<int(GroupMember&)> fn;
		{
			std::unique_lock<std::mutex> lk(g->mu);
			g->cv_job.wait(lk, [&] { return g->stop || g->job_id != seen; });
			if (g->stop) return;
			seen = g->job_id;
			fn = g->job;
		}
		m.rc = fn(m);
		m.err = m.rc ? cupt_last_error() : "";
		{
			std::lock_guard<std::mutex> lk(g->mu);
			if (--g->pending == 0) g->cv_done.notify_all();
		}
	}
}

// run fn on every local member's thread, wait for all, return the first failure (message re-raised on this thread)
int group_run(cupt_group* g, std::function<int(GroupMember&)> fn)
{
	{
		std::lock_guard<std::mutex> lk(g->mu);
		g->job = std::move(fn);
		g->pending = (int)g->local.size();
		++g->job_id;
	}
	g->cv_job.notify_all();
	{
		std::unique_lock<std::mutex> lk(g->mu);
		g->cv_done.wait(lk, [&] { return g->pending == 0; });
	}
	for (GroupMember& m : g->local)
		if (m.rc) return set_error(m.rc, "group member %d (device %d): %s", m.rank, m.ctx->device, m.err.c_str());
	return CUPT_OK;
}

int group_finish_create(cupt_group* g)
{
	for (size_t i = 0; i < g->local.size(); ++i) g->local[i].th = std::thread(group_worker, g, (int)i);
	return group_run(g, [](GroupMember& m) -> int {
		CK(cudaMalloc((void**)&m.d_hdr, sizeof(GroupHeader)));
		CK(cudaHostAlloc((void**)&m.h_hdr, sizeof(GroupHeader), cudaHostAllocDefault));
		return CUPT_OK;
	});
}

}  // namespace

namespace {

// the device arrays a member receives by broadcast, in the order of GroupHeader::bytes
enum { kScene0 = 0, kScene1 = 7, kBvh0 = 7, kBvh1 = 9, kBvh20 = 9, kBvh21 = 12 };
void** group_array(cupt_context* c, int i)
{
	switch (i) {
	case 0: return (void**)&c->d_verts;
	case 1: return (void**)&c->d_tris;
	case 2: return (void**)&c->d_trishade;
	case 3: return (void**)&c->d_mats;
	case 4: return (void**)&c->d_lights;
	case 5: return (void**)&c->d_light_cdf;
	case 6: return (void**)&c->d_spheres;
	case 7: return (void**)&c->d_nodes;
	case 8: return (void**)&c->d_prims;
	case 9: return (void**)&c->d_bvh2;
	case 10: return (void**)&c->d_bvh2_up;
	default: return (void**)&c->d_bvh2_leaf_parent;
	}
}

// Arrays [first, last) and the scalar fields that go with them travel from rank 0 to every other member.  root_rc: the
// outcome of rank 0's own upload (only rank 0's value matters); it is part of the header so that a failure on rank 0 fails
// the call on every member instead of leaving the others waiting in a broadcast.
int group_replicate(GroupMember& m, int first, int last, int root_rc)
{
	cupt_context* c = m.ctx;
	GroupHeader& h = *m.h_hdr;
	if (m.rank == 0) {
		memset(&h, 0, sizeof(h));
		for (int i = 0; i < 12; ++i) { void** p = group_array(c, i); h.bytes[i] = *p ? (uint64_t)c->used[p] : 0; }
		h.bytes[7] = (uint64_t)c->num_nodes * sizeof(Node8); h.bytes[8] = (uint64_t)c->num_prims * sizeof(Prim48);   // (the device builder fills these two itself)
		h.num_tris = c->num_tris; h.num_nodes = c->num_nodes; h.num_prims = c->num_prims; h.bvh2_leaves = c->bvh2_leaves;
		h.num_lights = c->num_lights; h.num_spheres = c->num_spheres; h.bvh_depth = c->bvh_depth; h.bvh2_size = c->bvh2_size; h.bvh2_depth = c->bvh2_depth;
		h.pad = root_rc;
		h.total_light_area = c->total_light_area;
		for (int a = 0; a < 3; ++a) { h.emis_lo[a] = c->emis_lo[a]; h.emis_hi[a] = c->emis_hi[a]; }
		CK(cudaMemcpyAsync(m.d_hdr, &h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
	}
	NK(g_nccl.Broadcast(m.d_hdr, m.d_hdr, sizeof(GroupHeader), ncclUint8, 0, m.comm, c->stream));
	if (m.rank != 0) {
		CK(cudaMemcpyAsync(&h, m.d_hdr, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
		CK(cudaStreamSynchronize(c->stream));
	}
	if (h.pad != 0) return m.rank == 0 ? h.pad : set_error(h.pad, "the upload on rank 0 failed");
	if (m.rank != 0)
		for (int i = first; i < last; ++i) {
			void** p = group_array(c, i);
			int rc = ensure_device(c, (char**)p, (size_t)h.bytes[i]);
			if (rc) return rc;   // (the other members then wait in the broadcast below: an allocation failure here is fatal for the group)
		}
	NK(g_nccl.GroupStart());
	for (int i = first; i < last; ++i)
		if (h.bytes[i]) {
			void* p = *group_array(c, i);
			ncclResult_t r = g_nccl.Broadcast(p, p, (size_t)h.bytes[i], ncclUint8, 0, m.comm, c->stream);
			if (r != ncclSuccess) { g_nccl.GroupEnd(); return set_error(CUPT_ERR_CUDA, "ncclBroadcast failed: %s", g_nccl.GetErrorString(r)); }
		}
	NK(g_nccl.GroupEnd());
	if (m.rank != 0) {
		if (first == kScene0) {
			c->num_tris = h.num_tris; c->num_lights = h.num_lights; c->num_spheres = h.num_spheres; c->total_light_area = h.total_light_area;
			for (int a = 0; a < 3; ++a) { c->emis_lo[a] = h.emis_lo[a]; c->emis_hi[a] = h.emis_hi[a]; }
		} else if (first == kBvh0) {
			c->num_nodes = h.num_nodes; c->num_prims = h.num_prims; c->bvh_depth = h.bvh_depth;
		} else {
			c->bvh2_size = h.bvh2_size; c->bvh2_depth = h.bvh2_depth; c->bvh2_leaves = h.bvh2_leaves;
		}
		c->wt_valid = false;
	}
	return CUPT_OK;
}

bool group_has_root(const cupt_group* g) { return !g->local.empty() && g->local[0].rank == 0; }

}  // namespace

extern "C" {

int cupt_group_unique_id(void* id128)
{
	if (!id128) return set_error(CUPT_ERR_INVALID, "null output");
	int rc = nccl_load();
	if (rc) return rc;
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
	NK(g_nccl.GetUniqueId((ncclUniqueId*)id128));
	return CUPT_OK;
}

void cupt_group_destroy(cupt_group* g)
{
	if (!g) return;
	{
		std::lock_guard<std::mutex> lk(g->mu);
		g->stop = true;
	}
	g->cv_job.notify_all();
	for (GroupMember& m : g->local) {
		if (m.th.joinable()) m.th.join();
		if (m.ctx) {
			cudaSetDevice(m.ctx->device);
			cudaStreamSynchronize(m.ctx->stream);
			if (m.comm) g_nccl.CommDestroy(m.comm);
			cudaFree(m.d_hdr);
			if (m.h_hdr) cudaFreeHost(m.h_hdr);
			cupt_destroy(m.ctx);
		}
	}
	delete g;
}

int cupt_group_create_local(const int* devices, int n, cupt_group** out)
{
	if (!devices || n < 1 || n > 64 || !out) return set_error(CUPT_ERR_INVALID, "bad arguments");
	int rc = nccl_load();
	if (rc) return rc;
	cupt_group* g = new (std::nothrow) cupt_group();
	if (!g) return set_error(CUPT_ERR_NOMEM, "out of memory");
	g->nranks = n;
	g->local.resize((size_t)n);
	for (int i = 0; i < n; ++i) {
		g->local[(size_t)i].rank = i;
		if ((rc = cupt_create(devices[i], &g->local[(size_t)i].ctx))) { cupt_group_destroy(g); return rc; }
	}
	std::vector<ncclComm_t> comms((size_t)n);
	ncclResult_t r = g_nccl.CommInitAll(comms.data(), n, devices);
	if (r != ncclSuccess) { cupt_group_destroy(g); return set_error(CUPT_ERR_CUDA, "ncclCommInitAll over %d devices failed: %s", n, g_nccl.GetErrorString(r)); }
	for (int i = 0; i < n; ++i) g->local[(size_t)i].comm = comms[(size_t)i];
	if ((rc = group_finish_create(g))) { cupt_group_destroy(g); return rc; }
	*out = g;
	return CUPT_OK;
}

int cupt_group_create_rank(int device, int rank, int nranks, const void* id128, cupt_group** out)
{
	if (!id128 || !out || nranks < 1 || rank < 0 || rank >= nranks) return set_error(CUPT_ERR_INVALID, "bad arguments");
	int rc = nccl_load();
	if (rc) return rc;
	cupt_group* g = new (std::nothrow) cupt_group();
	if (!g) return set_error(CUPT_ERR_NOMEM, "out of memory");
	g->nranks = nranks;
	g->local.resize(1);
	g->local[0].rank = rank;
	if ((rc = cupt_create(device, &g->local[0].ctx))) { cupt_group_destroy(g); return rc; }
	ncclUniqueId id;
	memcpy(&id, id128, sizeof(id));
	cudaSetDevice(device);
	ncclResult_t r = g_nccl.CommInitRank(&g->local[0].comm, nranks, id, rank);
	if (r != ncclSuccess) { cupt_group_destroy(g); return set_error(CUPT_ERR_CUDA, "ncclCommInitRank (rank %d of %d) failed: %s", rank, nranks, g_nccl.GetErrorString(r)); }
	if ((rc = group_finish_create(g))) { cupt_group_destroy(g); return rc; }
	*out = g;
	return CUPT_OK;
}

int cupt_group_info(const cupt_group* g, int* out3)
{
	if (!g || !out3) return set_error(CUPT_ERR_INVALID, "bad arguments");
	out3[0] = g->nranks; out3[1] = (int)g->local.size(); out3[2] = g->local.empty() ? -1 : g->local[0].rank;
	return CUPT_OK;
}

cupt_context* cupt_group_context(cupt_group* g, int local_index)
{
	if (!g || local_index < 0 || local_index >= (int)g->local.size()) { set_error(CUPT_ERR_INVALID, "no such group member"); return nullptr; }
	return g->local[(size_t)local_index].ctx;
}

int cupt_group_set_scene(cupt_group* g, const cupt_scene* s)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	if (group_has_root(g) && !s) return set_error(CUPT_ERR_INVALID, "the process that holds rank 0 must pass the scene");
	return group_run(g, [s](GroupMember& m) -> int {
		const int rc0 = m.rank == 0 ? cupt_set_scene(m.ctx, s) : 0;
		return group_replicate(m, kScene0, kScene1, rc0);
	});
}

int cupt_group_set_bvh(cupt_group* g, const cupt_bvh* b)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	if (group_has_root(g) && !b) return set_error(CUPT_ERR_INVALID, "the process that holds rank 0 must pass the BVH");
	return group_run(g, [b](GroupMember& m) -> int {
		const int rc0 = m.rank == 0 ? cupt_set_bvh(m.ctx, b) : 0;
		return group_replicate(m, kBvh0, kBvh1, rc0);
	});
}

// the wide BVH built on rank 0's GPU (cupt_build_bvh_device) and broadcast
int cupt_group_build_bvh_device(cupt_group* g, const cupt_bvh_opts* opts, cupt_bvh_info* info)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	return group_run(g, [opts, info](GroupMember& m) -> int {
		const int rc0 = m.rank == 0 ? cupt_build_bvh_device(m.ctx, opts, info) : 0;
		return group_replicate(m, kBvh0, kBvh1, rc0);
	});
}

int cupt_group_set_bvh2(cupt_group* g, const cupt_bvh2_node* nodes, int size, int depth)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	if (group_has_root(g) && !nodes) return set_error(CUPT_ERR_INVALID, "the process that holds rank 0 must pass the nodes");
	return group_run(g, [=](GroupMember& m) -> int {
		const int rc0 = m.rank == 0 ? cupt_set_bvh2(m.ctx, nodes, size, depth) : 0;
		return group_replicate(m, kBvh20, kBvh21, rc0);
	});
}

int cupt_group_set_camera(cupt_group* g, const cupt_camera* cam)
{
	if (!g || !cam) return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (GroupMember& m : g->local) { const int rc = cupt_set_camera(m.ctx, cam); if (rc) return rc; }
	return CUPT_OK;
}

int cupt_group_set_render_opts(cupt_group* g, const cupt_render_opts* o)
{
	if (!g || !o) return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (GroupMember& m : g->local) { const int rc = cupt_set_render_opts(m.ctx, o); if (rc) return rc; }
	return CUPT_OK;
}

int cupt_group_reset_accum(cupt_group* g)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	return group_run(g, [](GroupMember& m) -> int { return cupt_reset_accum(m.ctx); });
}

int cupt_group_render(cupt_group* g, int first_sample, int num_samples)
{
	if (!g || first_sample < 0 || num_samples < 0) return set_error(CUPT_ERR_INVALID, "bad arguments");
	const int n = g->nranks;
	// every member renders its share (the same split as cudapathtracer_b200/sharding.py: contiguous, sizes differ by <= 1) ...
	int rc = group_run(g, [=](GroupMember& m) -> int {
		const int base = num_samples / n, extra = num_samples % n;
		const int first = m.rank * base + std::min(m.rank, extra), count = base + (m.rank < extra ? 1 : 0);
		return cupt_render(m.ctx, first_sample + first, count);
	});
	if (rc || n == 1) return rc;
	// ... and the f32 sums meet on rank 0: the one exchange step of the path
	return group_run(g, [=](GroupMember& m) -> int {
		cupt_context* c = m.ctx;
		const int base = num_samples / n, extra = num_samples % n;
		const int count = base + (m.rank < extra ? 1 : 0);
		const size_t cnt = c->accum_pixels * 3;
		NK(g_nccl.Reduce(c->d_accum, c->d_accum, cnt, ncclFloat, ncclSum, 0, m.comm, c->stream));
		if (m.rank == 0) c->samples_accumulated += (uint64_t)(num_samples - count);
		else {
			CK(cudaMemsetAsync(c->d_accum, 0, cnt * sizeof(float), c->stream));
			c->samples_accumulated = 0;
		}
		return CUPT_OK;
	});
}

int cupt_group_sync(cupt_group* g)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	return group_run(g, [](GroupMember& m) -> int { return cupt_sync(m.ctx); });
}

// the image lives on rank 0: a process that does not hold it gets 1 (nothing written)
int cupt_group_read_accum(cupt_group* g, float* rgb_host, int normalise)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	if (!group_has_root(g)) return 1;
	return cupt_read_accum(g->local[0].ctx, rgb_host, normalise);
}

int cupt_group_read_image_rgb8(cupt_group* g, uint8_t* rgb, int mirror_x)
{
	if (!g) return set_error(CUPT_ERR_INVALID, "null group");
	if (!group_has_root(g)) return 1;
	return cupt_read_image_rgb8(g->local[0].ctx, rgb, mirror_x);
}

// sums over the members this process drives
int cupt_group_get_stats(cupt_group* g, cupt_stats* out)
{
	if (!g || !out) return set_error(CUPT_ERR_INVALID, "bad arguments");
	memset(out, 0, sizeof(*out));
	for (GroupMember& m : g->local) {
		cupt_stats s;
		const int rc = cupt_get_stats(m.ctx, &s);
		if (rc) return rc;
		out->rays += s.rays; out->samples += s.samples; out->nodes_visited += s.nodes_visited; out->tris_tested += s.tris_tested;
		out->kernel_launches += s.kernel_launches; out->extend_launches += s.extend_launches; out->strict_retraces += s.strict_retraces;
		out->ms_raygen = std::max(out->ms_raygen, s.ms_raygen); out->ms_extend = std::max(out->ms_extend, s.ms_extend);
		out->ms_shade = std::max(out->ms_shade, s.ms_shade); out->ms_accum = std::max(out->ms_accum, s.ms_accum);
	}
	return CUPT_OK;
}

}  // extern "C"
