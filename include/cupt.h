/* cupt.h — C-ABI of the B200-native path-tracing hot path (drop-in for CulDeVu/CUDAPathTracer's render loop).
 *
 * The reference has no FFI: its "API" is the set of PODs, globals and free functions that main()
 * (reference kernel.cu:565-790) uses inside one translation unit.  Every entry point below names the
 * reference interface it replaces (file:line relative to the reference tree).  All functions return
 * 0 on success and a negative cupt_status on failure (the reference printf()s and continues,
 * kernel.cu:37-42, or exit(0)s, BVH.h:339-340; this library never exits and never throws across
 * the ABI); cupt_last_error() gives the message for the calling thread.
 *
 * Signatures carry plain pointers and sizes only.  POD layouts are byte-identical to the reference's
 * (sizes pinned by tests/test_abi.py against the reference compiled in oracle/_ref).
 */
#ifndef CUPT_H
#define CUPT_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUPT_VERSION 100

typedef enum cupt_status {
	CUPT_OK = 0,
	CUPT_ERR_INVALID = -1,   /* bad argument / bad call order */
	CUPT_ERR_CUDA = -2,      /* CUDA runtime error (message carries cudaGetErrorString) */
	CUPT_ERR_NO_DEVICE = -3, /* no CUDA device: there is NO CPU fallback */
	CUPT_ERR_IO = -4,        /* file could not be opened / written */
	CUPT_ERR_PARSE = -5,     /* OBJ / MTL parse error */
	CUPT_ERR_LIMIT = -6,     /* BVH depth / index limit (kernel.cu:627-631, BVH.h:337-341) */
	CUPT_ERR_NOMEM = -7
} cupt_status;

/* ---- PODs, byte-identical to the reference ---------------------------------------------------- */
typedef struct cupt_vec3 { float x, y, z; } cupt_vec3;                       /* vec3.h:4-8, 12 B   */
typedef struct cupt_triangle {                                               /* modelLoader.h:14-19, 28 B */
	int32_t v0, v1, v2;
	cupt_vec3 norm;
	int32_t mat;
} cupt_triangle;
typedef struct cupt_material {                                               /* modelLoader.h:21-25 + color.h:4-6, 48 B */
	double albedo[3];
	double emission[3];   /* the reference spells it `emmision` */
} cupt_material;
typedef struct cupt_camera {                                                 /* camera.h:26-35, 32 B */
	cupt_vec3 pos;
	float distFromFilm, focalLength, radius;
	int32_t pxlWidth, pxlHeight;
} cupt_camera;
typedef struct cupt_bvh2_node {                                              /* BVH.h:111-115, 32 B */
	cupt_vec3 lo, hi;
	uint32_t left, right;   /* inner index, or triangle | CUPT_BVH_LEAF_FLAG (limits.h:7) */
} cupt_bvh2_node;
typedef struct cupt_sphere {                                                 /* sphere.h:7-13, 64 B */
	cupt_vec3 pos;
	float rad;
	double diffuse[3];
	double emm[3];
} cupt_sphere;

#define CUPT_BVH_LEAF_FLAG 0x80000000u   /* limits.h:7 */
#define CUPT_MAX_BVH_INDEX 0x7fffffffu   /* limits.h:6 */
#define CUPT_MAX_FLOAT 100000.0f         /* limits.h:3 — the reference's "infinity": miss = {id -1, t 1e5} */
#define CUPT_MAX_BVH2_DEPTH 64           /* kernel.cu:35 */

typedef struct cupt_scene cupt_scene;       /* host scene: the reference's globals verts/tris/mats/lights/totalLightArea (modelLoader.h:43-47) */
typedef struct cupt_bvh cupt_bvh;           /* host image of the wide quantised BVH (replaces BVH.h:133-474) */
typedef struct cupt_context cupt_context;   /* one per GPU: device scene, BVH, wavefront state, accumulator */

const char* cupt_last_error(void);
int cupt_version(void);
/* sizes of the PODs above as compiled into the library, in declaration order (7 values) */
int cupt_struct_sizes(int* out7);

/* ---- scene ingest (host) — replaces loadOBJ + the globals, modelLoader.h:43-47,125-210 -------- */
int cupt_scene_create(cupt_scene** out);
void cupt_scene_destroy(cupt_scene* s);
int cupt_scene_clear(cupt_scene* s);
/* loadOBJ(filename, origin, scale, flipNormals) — modelLoader.h:125.  `mtl_dir` is the MTL base path the
 * reference hard-codes as "models/" (modelLoader.h:132); NULL means "models/".  Appends, like the reference. */
int cupt_scene_load_obj(cupt_scene* s, const char* path, const char* mtl_dir,
                        const float origin[3], float scale, int flip_normals);
/* SURVEY 8f rank 2 — ingest fidelity beyond the reference's loader: */
enum { CUPT_OBJ_FACE_MATERIALS = 1,   /* material_ids[k] per triangle instead of material_ids[0] per shape (modelLoader.h:189);
                                         the vendored loader splits shapes at usemtl, so this only guards against a parser that does not */
       CUPT_OBJ_DEDUP_MATERIALS = 2 };/* one table entry per distinct material instead of the table appended twice per file (modelLoader.h:138-159) */
int cupt_scene_load_obj_ex(cupt_scene* s, const char* path, const char* mtl_dir, const float origin[3], float scale,
                           int flip_normals, int flags);   /* flags = 0: exactly cupt_scene_load_obj */
/* binary scene cache: the arrays verbatim with a checksum; load REPLACES the scene's contents (CUPT_ERR_PARSE on a
 * foreign, truncated or corrupted file) */
int cupt_scene_save(const cupt_scene* s, const char* path);
int cupt_scene_load(cupt_scene* s, const char* path);
/* the same append for an in-memory shape: materials first (appended twice like modelLoader.h:138-159 when
 * duplicate_table != 0), then one call per shape with ONE material per shape (modelLoader.h:189). */
int cupt_scene_add_materials(cupt_scene* s, const cupt_material* mats, int n, int duplicate_table, int* first_index);
int cupt_scene_add_shape(cupt_scene* s, const float* positions, int nverts, const uint32_t* indices, int ntris,
                         int material, const float origin[3], float scale, int flip_normals);
/* per-face materials: like add_shape with material taken per triangle */
int cupt_scene_add_shape_facemats(cupt_scene* s, const float* positions, int nverts, const uint32_t* indices,
                                  int ntris, const int32_t* face_materials, const float origin[3], float scale,
                                  int flip_normals);
/* spheres: sphere.h:7-13 has the POD only; the intersector is new (parity unpinned) */
/* BASELINE config 5 ("mixed diffuse/specular"): mark a material as a perfect mirror.  NOT in the reference (all its
 * materials are diffuse, kernel.cu:101-104); honoured by CUPT_INTEGRATOR_PT only; parity unpinned. */
int cupt_scene_set_specular(cupt_scene* s, int material, int specular);
int cupt_scene_add_spheres(cupt_scene* s, const cupt_sphere* spheres, int n);
/* raw access to the arrays in reference layout */
int cupt_scene_set_arrays(cupt_scene* s, const cupt_vec3* verts, int nverts, const cupt_triangle* tris, int ntris,
                          const cupt_material* mats, int nmats, const int32_t* lights, int nlights,
                          float total_light_area);
int cupt_scene_counts(const cupt_scene* s, int* out5 /* verts, tris, mats, lights, spheres */);
int cupt_scene_get_arrays(const cupt_scene* s, cupt_vec3* verts, cupt_triangle* tris, cupt_material* mats,
                          int32_t* lights);
float cupt_scene_total_light_area(const cupt_scene* s);

/* ---- reference-layout BVH2 — buildBVH(), BVH.h:443-474 (compatibility + tie-break ranks) ------- */
/* Returns a malloc'ed BVH_array{root,size,depth} image; reproduces the 3x3x3 grid heuristic bit for bit. */
int cupt_bvh2_build(const cupt_scene* s, cupt_bvh2_node** nodes, int* size, int* depth);
void cupt_bvh2_free(cupt_bvh2_node* nodes);
/* rank[tri] = position of tri in the left-first leaf order of trace() (kernel.cu:139-151): the reference's
 * tie-break (`t < closestT` strict, kernel.cu:127) expressed so any other tree can reproduce it. */
int cupt_bvh2_leaf_order(const cupt_bvh2_node* nodes, int size, int ntris, uint32_t* rank);

/* ---- wide BVH (host build) — replaces BVH.h:133-328 -------------------------------------------- */
typedef struct cupt_bvh_opts {
	int sah_bins;          /* bins per axis of the binned-SAH BVH2 builder (default 32) */
	int max_leaf_tris;     /* triangles per BVH8 child leaf, 1..3 (default 3) */
	int threads;           /* host threads, 0 = all */
	const uint32_t* tie_rank;  /* optional rank per triangle (cupt_bvh2_leaf_order); NULL = triangle index */
	int ploc_radius;       /* cupt_build_bvh_device only: neighbour search radius of the clustering: 1, 2, 4 (default, also for any other value), 8, 16 or 32 */
	int reinsert_passes;   /* cupt_bvh_build only: sweeps of insertion-based optimisation over the binary tree before the collapse (every
	                          subtree is re-inserted where the tree's total surface area drops most).  0 = default (2 sweeps up to 2^20
	                          primitives, none above: the search is single-threaded), < 0 = none, k > 0 = k sweeps (at most 16) */
	int greedy_collapse;   /* cupt_bvh_build only: 1 = open the largest child until there are eight (the first collapse rule, kept for
	                          comparison: tools/build_compare.py); default 0 = cost-optimal cut */
} cupt_bvh_opts;
typedef struct cupt_bvh_info {
	int64_t num_nodes;       /* 80-byte BVH8 nodes */
	int64_t num_tri_records; /* 48-byte triangle records */
	int64_t num_tris, num_spheres;
	int32_t depth;
	int32_t node_bytes, tri_bytes;
	double sah_cost;
	double build_seconds;
} cupt_bvh_info;
int cupt_bvh_build(const cupt_scene* s, const cupt_bvh_opts* opts, cupt_bvh** out);
void cupt_bvh_destroy(cupt_bvh* b);
int cupt_bvh_get_info(const cupt_bvh* b, cupt_bvh_info* info);
/* raw images (for tests: structural invariants are checked on the host) */
int cupt_bvh_get_arrays(const cupt_bvh* b, void* nodes80, void* tris48);

/* ---- device context — replaces the cudaMalloc/cudaMemcpy block, kernel.cu:633-700 -------------- */
int cupt_create(int device, cupt_context** out);
void cupt_destroy(cupt_context* c);
int cupt_set_stream(cupt_context* c, void* cuda_stream);          /* default: a stream owned by the context */
int cupt_set_scene(cupt_context* c, const cupt_scene* s);         /* kernel.cu:664-684 */
int cupt_set_bvh(cupt_context* c, const cupt_bvh* b);             /* kernel.cu:686-691 (wide layout) */
int cupt_set_bvh2(cupt_context* c, const cupt_bvh2_node* nodes, int size, int depth); /* kernel.cu:686-691 verbatim */
/* GPU-side construction of the wide BVH from the scene cupt_set_scene uploaded: replaces buildBVH (BVH.h:133-474)
 * AND the upload (kernel.cu:686-691) in one call; Morton sort + parallel locally-ordered clustering + 8-wide collapse
 * on the device (csrc/bvh_device.cu).  opts: max_leaf_tris and tie_rank are honoured (sah_bins, threads: host builder
 * only); info may be NULL.  The tree differs from cupt_bvh_build's, the hits do not. */
int cupt_build_bvh_device(cupt_context* c, const cupt_bvh_opts* opts, cupt_bvh_info* info);
/* copy the device's wide BVH back (tests: structural invariants); counts = {nodes, records, depth, PLOC rounds}; any pointer may be NULL */
int cupt_read_bvh(cupt_context* c, void* nodes80, void* tris48, int64_t* counts);
/* test hook: the stable LSD radix sort the build uses for the Morton order (csrc/radix_sort.cuh), on host pairs, in
 * place; sorts on key bits [63 - 8 * passes, 63), passes = 1..8 */
int cupt_debug_sort_pairs(cupt_context* c, uint64_t* keys, uint32_t* vals, int64_t n, int passes);
/* the build workspace (about 270 B per primitive) stays allocated between builds; this frees it */
int cupt_release_workspace(cupt_context* c);
/* CUDA-event times of the last cupt_build_bvh_device in ms: {total, boxes+Morton+sort, clustering, collapse+emit} */
int cupt_get_build_times(cupt_context* c, float* ms4);
int cupt_set_camera(cupt_context* c, const cupt_camera* cam);     /* kernel.cu:642-651 */

enum { CUPT_INTEGRATOR_PT = 2 /* radianceAlongSingleStep2, kernel.cu:417-515 */,
       CUPT_INTEGRATOR_BDPT = 1 /* radianceAlongSingleStep, kernel.cu:217-415: what drawPixel calls at HEAD (kernel.cu:549); max_depth is ignored */ };
/* CUPT_TRAVERSAL_WIDE: closest hit = argmin over all primitives (ties by the reference's visiting order).  Inside cupt_render
 * the LAST segment of a path — whose hit is only asked for its emission, kernel.cu:453-457 — may end at the first hit known to
 * lie in front of every emitter; the image is the same, cupt_read_primary_hits / cupt_trace_rays always return the closest hit.
 * (cupt_render_opts.no_terminal_stop = 1 switches that early stop off.) */
enum { CUPT_TRAVERSAL_WIDE = 0, CUPT_TRAVERSAL_REF_BVH2 = 1 /* trace(), kernel.cu:112-161 op for op */,
       /* wide traversal, then every hit is checked against the reference tree's ancestor boxes with the reference's
        * slab arithmetic and re-traced the reference's way if the reference would not have reached it: results equal
        * trace() bit for bit on every ray.  Needs cupt_set_bvh AND cupt_set_bvh2, a wide BVH built with the
        * cupt_bvh2_leaf_order ranks, and a scene without spheres. */
       CUPT_TRAVERSAL_WIDE_STRICT = 2 };
enum { CUPT_TRI_REFEXACT = 0 /* triIntersect, modelLoader.h:49-82, bit-identical */,
       CUPT_TRI_WATERTIGHT = 1 /* Woop-Benthin-Wald test on exact vertices (wide traversal only): no ray slips between two
                                  triangles that share an edge; NOT the reference's arithmetic, outside the parity statement */ };

typedef struct cupt_render_opts {
	int32_t max_depth;        /* NUM_BOUNCES (kernel.cu:33) made run-time: path segments per sample */
	int32_t integrator;       /* CUPT_INTEGRATOR_* */
	int32_t traversal;        /* CUPT_TRAVERSAL_* */
	int32_t tri_test;         /* CUPT_TRI_* */
	uint64_t seed;            /* kernel.cu:532 uses 1234 */
	int32_t samples_per_pass; /* samples of every pixel in flight per wavefront pass (0 = auto) */
	int32_t collect_stats;    /* 1: count node visits / triangle tests (slower; for the roofline's bytes/ray) */
	int32_t time_stages;      /* 1: CUDA events around every stage launch */
	int32_t fetch_threshold;  /* extend: refill idle lanes of a warp from the ray queue when <= this many are busy (0 = default 24) */
	int32_t defer_pairs;      /* extend: hold primitive tests one traversal step while a warp has fewer (ray, primitive) pairs than this (0 = default 20, 1 = never) */
	int32_t share_primary_hits; /* 1: with a pinhole camera (radius == 0, the reference's setting, kernel.cu:646) every sample of a
	                             * pixel has the same primary ray: trace it once per pixel and pass and reuse the hit.  Same image
	                             * bit for bit, fewer rays traced (and counted).  Default 0; CUPT_INTEGRATOR_PT and CUPT_TRAVERSAL_WIDE
	                             * only (the reference's slab test distinguishes the +0 / -0 direction components that
	                             * different samples get on the centre column and row). */
	int32_t no_terminal_stop; /* 1: last segments are traced to their closest hit like every other ray (default 0: early stop, see
	                           * CUPT_TRAVERSAL_WIDE above; never used by the strict / reference modes) */
	int32_t record_retraces;  /* CUPT_TRAVERSAL_WIDE_STRICT: 1 = count, per pixel, the rays of its paths whose wide hit the reference
	                           * would not have reached (cupt_read_retrace_counts); the counts accumulate until cupt_reset_accum */
} cupt_render_opts;
int cupt_default_render_opts(cupt_render_opts* o);
int cupt_set_render_opts(cupt_context* c, const cupt_render_opts* o);

/* drawPixel<<<>>> loop, kernel.cu:709-736: adds samples [first_sample, first_sample+num_samples) of every
 * pixel to the float sum accumulator.  Asynchronous on the context's stream. */
int cupt_render(cupt_context* c, int first_sample, int num_samples);
int cupt_sync(cupt_context* c);                                   /* cudaDeviceSynchronize, kernel.cu:716 */
/* Multi-GPU by image tiles instead of sample ranges (the reference renders the whole image on device 0, kernel.cu:569,
 * 709-736): cupt_render traces only the pixels of tiles [first_tile, first_tile + num_tiles) — 8x4-pixel tiles numbered
 * row-major — and leaves the other pixels of the accumulator untouched.  num_tiles <= 0: the whole image again. */
int cupt_set_tile_window(cupt_context* c, int first_tile, int num_tiles);
int cupt_tile_count(const cupt_context* c, int* num_tiles);       /* tiles of the current camera's image */
int cupt_reset_accum(cupt_context* c);                            /* setupImgBuffer, kernel.cu:520-526 */
/* image readback, kernel.cu:760: row-major P*3 floats; normalise != 0 divides by the samples accumulated */
int cupt_read_accum(cupt_context* c, float* rgb_host, int normalise);
/* device-side view of the accumulator (P*3 f32 sums) for the multi-GPU reduce: copies into caller memory */
int cupt_copy_accum_device(cupt_context* c, float* dst_device);
int cupt_add_accum_device(cupt_context* c, const float* src_device, int samples);
/* parity hooks: primary hit triangle id and raw t per pixel (row-major), and closest hits for caller rays */
int cupt_read_primary_hits(cupt_context* c, int32_t* tri_id, float* t);
int cupt_trace_rays(cupt_context* c, const float* rays_host /* n*6: o, dir */, int64_t n, int32_t* tri_id, float* t);
int cupt_camera_rays(cupt_context* c, float* rays_host /* P*6 */);
/* cupt_trace_rays without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as cupt_trace_rays bit for bit.  All scratch (ray / hit staging, path records, queue, counters) lives in one
 * file-scope device array of CUPT_NOALLOC_MAX_RAYS * 68 + 576 bytes (about 4.3 MiB) that every GPU which loads the library
 * holds until the process exits; the call performs no cudaMalloc / cudaFree / allocating library call.
 *   rays, tri_id, t   host OR device pointers (n*6 floats in, n ids and n distances out); with device pointers nothing is
 *                     staged through the host
 *   n                 0 .. CUPT_NOALLOC_MAX_RAYS; more is refused with CUPT_ERR_LIMIT (split the batch)
 *   synchronise       != 0: wait for the context's stream before returning (needed before reading host outputs)
 * CUPT_TRAVERSAL_WIDE or CUPT_TRAVERSAL_REF_BVH2 with CUPT_TRI_REFEXACT (the strict mode and the watertight test allocate on
 * demand and are refused).  The array is shared by every call on a device and keeps the last call's contents (what must start
 * at zero is cleared by the call): calls that OVERLAP on one GPU — two contexts or two streams — race; serialise them. */
#define CUPT_NOALLOC_MAX_RAYS 65536
int cupt_trace_rays_noalloc(cupt_context* c, const float* rays, int64_t n, int32_t* tri_id, float* t, int synchronise);
/* parity hook (record_retraces): per pixel, row-major, how many rays of the paths rendered since cupt_reset_accum had a
 * wide-BVH hit the reference's trace() would not have reached (BVH.h:65,79: its slab test is not conservative) and
 * were therefore re-traced.  Pixels with 0 carry paths on which the plain wide mode and the reference agree ray by ray. */
int cupt_read_retrace_counts(cupt_context* c, uint32_t* counts);

typedef struct cupt_stats {
	uint64_t rays;            /* extend invocations (one ray = one trace(), counted not nominal) */
	uint64_t samples;         /* pixel-samples started */
	uint64_t nodes_visited;   /* wide-node fetches (collect_stats) */
	uint64_t tris_tested;     /* triangle records fetched (collect_stats) */
	uint64_t kernel_launches; /* launches of this library's kernels */
	double ms_raygen, ms_extend, ms_shade, ms_accum;  /* time_stages */
	uint64_t extend_launches;
	uint64_t strict_retraces; /* CUPT_TRAVERSAL_WIDE_STRICT: rays re-traced with the reference's algorithm */
} cupt_stats;
int cupt_get_stats(cupt_context* c, cupt_stats* out);
int cupt_reset_stats(cupt_context* c);

/* ---- several GPUs — replaces the single-device loop (kernel.cu:569 `cudaGetDeviceProperties(&p, 0)`, kernel.cu:709-736) -----
 * A group is one context per GPU, one host thread per context and an NCCL communicator owned by the group (SURVEY 8b
 * "Threading").  The scene and the BVH are staged through host memory once, on rank 0, and reach the other members by
 * ncclBroadcast over NVLink; cupt_group_render gives member r the r-th of N contiguous sample ranges of every pixel (disjoint
 * Philox sample indices, replicated BVH) and sums the f32 accumulators onto rank 0 with one ncclReduce — the path's only
 * exchange step; the image is read from rank 0.  NCCL (libnccl.so.2) is dlopen()ed by the first group call: a process that
 * makes no group does not need it.
 *   cupt_group_create_local   all members are threads of this process (devices[i] = CUDA device of rank i)
 *   cupt_group_create_rank    this process drives ONE member of an nranks-wide group (one process per GPU: torchrun, MPI);
 *                             id128 = the 128 bytes cupt_group_unique_id wrote on one process, distributed by the caller
 * Group calls are collective: every process of the group makes the same calls in the same order.  A call that fails on one
 * process only (e.g. out of memory) can leave the others waiting in NCCL, as with any collective. */
typedef struct cupt_group cupt_group;
int cupt_group_unique_id(void* id128);
int cupt_group_create_local(const int* devices, int n, cupt_group** out);
int cupt_group_create_rank(int device, int rank, int nranks, const void* id128, cupt_group** out);
void cupt_group_destroy(cupt_group* g);
int cupt_group_info(const cupt_group* g, int* out3 /* members in the group, members driven by this process, rank of the first of those */);
cupt_context* cupt_group_context(cupt_group* g, int local_index);   /* a member's context, for per-context calls (stats, hooks) between group calls */
/* cupt_set_scene / cupt_set_bvh / cupt_build_bvh_device / cupt_set_bvh2 on rank 0, then broadcast; processes that do not hold
 * rank 0 may pass NULL for the scene / BVH / nodes */
int cupt_group_set_scene(cupt_group* g, const cupt_scene* s);
int cupt_group_set_bvh(cupt_group* g, const cupt_bvh* b);
int cupt_group_build_bvh_device(cupt_group* g, const cupt_bvh_opts* opts, cupt_bvh_info* info);
int cupt_group_set_bvh2(cupt_group* g, const cupt_bvh2_node* nodes, int size, int depth);
int cupt_group_set_camera(cupt_group* g, const cupt_camera* cam);            /* the same camera on every member */
int cupt_group_set_render_opts(cupt_group* g, const cupt_render_opts* o);
int cupt_group_reset_accum(cupt_group* g);
/* samples [first_sample, first_sample + num_samples) of every pixel, split over the members; asynchronous.  Afterwards rank 0's
 * accumulator holds the sum of everything rendered since cupt_group_reset_accum and the other members' are zero. */
int cupt_group_render(cupt_group* g, int first_sample, int num_samples);
int cupt_group_sync(cupt_group* g);
/* the image, from rank 0 (cupt_read_accum / cupt_read_image_rgb8).  Returns 1 and writes nothing on a process that does not hold rank 0. */
int cupt_group_read_accum(cupt_group* g, float* rgb_host, int normalise);
int cupt_group_read_image_rgb8(cupt_group* g, uint8_t* rgb, int mirror_x);
int cupt_group_get_stats(cupt_group* g, cupt_stats* out);                    /* counters summed (stage times: max) over this process's members */

/* ---- output — kernel.cu:763-778 ------------------------------------------------------------------ */
/* P3 PPM: x mirrored, Reinhard c/(c+1) (color.h:59-62), gamma 1/2.2 (color.h:68-71), (int)(c*255) truncation */
int cupt_write_ppm(const char* path, const float* rgb_mean, int width, int height);
/* the same output stage on the device: mean = sum / samples, normalized (color.h:59-62), gammaCorrect with (float)(1/2.2)
 * (color.h:68-71, kernel.cu:774), (int)(c*255); 3 bytes per pixel come back instead of 12.  mirror_x = 1 flips x like
 * the reference's writer (kernel.cu:768).  rgb: width*height*3 bytes, row-major. */
int cupt_read_image_rgb8(cupt_context* c, uint8_t* rgb, int mirror_x);
int cupt_write_ppm8(const char* path, const uint8_t* rgb, int width, int height);   /* P3 text like kernel.cu:763-778, no mirroring */
/* the float image itself as a Portable Float Map (binary "PF", little-endian, rows bottom to top): what cupt_read_accum returned,
 * bit for bit, for tools that want the radiance and not the reference's 8-bit tonemapped text (kernel.cu:763-778 has no such output) */
int cupt_write_pfm(const char* path, const float* rgb, int width, int height);
/* traversal-cost heat map of the primary rays: wide-BVH nodes visited / primitives tested per pixel (the reference's
 * per-ray step counter, kernel.cu:121,132,152, which drew keepers/"sponza bvh.ppm").  Either pointer may be NULL. */
int cupt_read_heatmap(cupt_context* c, float* nodes_visited, float* prims_tested);
int cupt_tonemap(const float* rgb_mean, int64_t npixels, int32_t* out_rgb255);

#ifdef __cplusplus
}
#endif
#endif /* CUPT_H */
