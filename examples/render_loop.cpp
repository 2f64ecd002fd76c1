// The reference's main() (kernel.cu:565-790) written against libcupt.so: load OBJ files, build the BVH, upload,
// render N samples, read the image back, write image.ppm, print timing and counted Mrays/s.  Pure C++ host code over
// the C-ABI (include/cupt.h); see INTEGRATION.md for the line-by-line correspondence.
//
//   render_loop <scene.obj> <mtl_dir/> [width height spp max_depth [out.ppm [strict|wide [gpus]]]]
//
// gpus >= 1: the same loop over a cupt_group of that many GPUs (devices 0 .. gpus-1; one host thread per GPU, scene and BVH
// broadcast over NVLink, sample ranges split over the members, one ncclReduce): the image is the single-GPU one.
//
// Build: g++ -std=c++17 -O2 examples/render_loop.cpp -Iinclude -Lcudapathtracer_b200 -lcupt -Wl,-rpath,$PWD/cudapathtracer_b200 -o examples/render_loop
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "cupt.h"

#define CHECK(call)                                                                        \
	do {                                                                                   \
		if ((call) != CUPT_OK) {                                                           \
			fprintf(stderr, "%s failed: %s\n", #call, cupt_last_error());                  \
			return 1;                                                                      \
		}                                                                                  \
	} while (0)

int main(int argc, char** argv)
{
	if (argc < 3) { fprintf(stderr, "usage: %s scene.obj mtl_dir/ [width height spp max_depth [out.ppm [strict|wide [gpus]]]]\n", argv[0]); return 2; }
	const int width = argc > 3 ? atoi(argv[3]) : 512, height = argc > 4 ? atoi(argv[4]) : 512;   // IMAGE_WIDTH/HEIGHT, kernel.cu:28-29
	const int spp = argc > 5 ? atoi(argv[5]) : 100, max_depth = argc > 6 ? atoi(argv[6]) : 3;    // NUM_SAMPLES, NUM_BOUNCES, kernel.cu:32-33
	const char* out = argc > 7 ? argv[7] : "image.ppm";
	const bool strict = argc > 8 && !strcmp(argv[8], "strict");
	const int gpus = argc > 9 ? atoi(argv[9]) : 0;

	cupt_scene* scene;
	CHECK(cupt_scene_create(&scene));
	const float origin[3] = {0, 0, 0};
	CHECK(cupt_scene_load_obj(scene, argv[1], argv[2], origin, 1.0f, 0));                        // loadOBJ, kernel.cu:591
	int n[5];
	CHECK(cupt_scene_counts(scene, n));
	printf("scene: %d verts, %d tris, %d mats, %d lights, light area %f\n", n[0], n[1], n[2], n[3], cupt_scene_total_light_area(scene));

	cupt_bvh2_node* bvh2; int size2, depth2;
	CHECK(cupt_bvh2_build(scene, &bvh2, &size2, &depth2));                                      // buildBVH, kernel.cu:601
	if (depth2 >= CUPT_MAX_BVH2_DEPTH) { printf("BVH depth is too big\n"); return 0; }         // kernel.cu:627-631
	std::vector<uint32_t> rank((size_t)n[1]);
	CHECK(cupt_bvh2_leaf_order(bvh2, size2, n[1], rank.data()));
	cupt_bvh_opts bo = {};
	bo.tie_rank = rank.data();
	cupt_bvh* wide;
	CHECK(cupt_bvh_build(scene, &bo, &wide));
	cupt_bvh_info info;
	CHECK(cupt_bvh_get_info(wide, &info));
	printf("reference-layout BVH: %d nodes, depth %d; wide BVH: %lld nodes, depth %d, built in %.3f s\n", size2, depth2,
	       (long long)info.num_nodes, info.depth, info.build_seconds);

	cupt_camera cam = {{0, 1, 3}, 1.0f, 3.0f, 0.0f, width, height};                            // kernel.cu:642-648
	if (gpus >= 1) {
		std::vector<int> devices((size_t)gpus);
		for (int i = 0; i < gpus; ++i) devices[(size_t)i] = i;
		cupt_group* grp;
		CHECK(cupt_group_create_local(devices.data(), gpus, &grp));
		CHECK(cupt_group_set_scene(grp, scene));
		CHECK(cupt_group_set_bvh(grp, wide));
		if (strict) CHECK(cupt_group_set_bvh2(grp, bvh2, size2, depth2));
		CHECK(cupt_group_set_camera(grp, &cam));
		cupt_render_opts gro;
		CHECK(cupt_default_render_opts(&gro));
		gro.max_depth = max_depth;
		gro.traversal = strict ? CUPT_TRAVERSAL_WIDE_STRICT : CUPT_TRAVERSAL_WIDE;
		CHECK(cupt_group_set_render_opts(grp, &gro));
		CHECK(cupt_group_render(grp, 0, gpus));                                                 // warm-up (one sample per member)
		CHECK(cupt_group_reset_accum(grp));
		for (int i = 0; i < gpus; ++i) CHECK(cupt_reset_stats(cupt_group_context(grp, i)));
		const auto g0 = std::chrono::steady_clock::now();
		CHECK(cupt_group_render(grp, 0, spp));
		CHECK(cupt_group_sync(grp));
		const double gms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - g0).count();
		cupt_stats gst;
		CHECK(cupt_group_get_stats(grp, &gst));
		printf("%d GPUs\ntotal ms: %f\nms per sample: %f\nMrays/s (counted): %f\n", gpus, gms, gms / spp, (double)gst.rays / (gms * 1000.0));
		std::vector<float> gimg((size_t)width * height * 3);
		CHECK(cupt_group_read_accum(grp, gimg.data(), 1));
		CHECK(cupt_write_ppm(out, gimg.data(), width, height));
		printf("wrote %s\n", out);
		cupt_group_destroy(grp); cupt_bvh_destroy(wide); cupt_bvh2_free(bvh2); cupt_scene_destroy(scene);
		return 0;
	}
	cupt_context* ctx;
	CHECK(cupt_create(0, &ctx));
	CHECK(cupt_set_scene(ctx, scene));
	CHECK(cupt_set_bvh(ctx, wide));
	if (strict) CHECK(cupt_set_bvh2(ctx, bvh2, size2, depth2));
	CHECK(cupt_set_camera(ctx, &cam));
	cupt_render_opts ro;
	CHECK(cupt_default_render_opts(&ro));
	ro.max_depth = max_depth;
	ro.traversal = strict ? CUPT_TRAVERSAL_WIDE_STRICT : CUPT_TRAVERSAL_WIDE;
	CHECK(cupt_set_render_opts(ctx, &ro));

	CHECK(cupt_render(ctx, 0, 1));                                                              // warm-up sample (not in the image)
	CHECK(cupt_reset_accum(ctx));
	CHECK(cupt_reset_stats(ctx));
	const auto t0 = std::chrono::steady_clock::now();
	CHECK(cupt_render(ctx, 0, spp));                                                            // the drawPixel loop, kernel.cu:709-736
	CHECK(cupt_sync(ctx));
	const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	cupt_stats st;
	CHECK(cupt_get_stats(ctx, &st));
	printf("total ms: %f\nms per sample: %f\nMrays/s (counted): %f\n", ms, ms / spp, (double)st.rays / (ms * 1000.0));   // kernel.cu:753-757

	std::vector<float> img((size_t)width * height * 3);
	CHECK(cupt_read_accum(ctx, img.data(), 1));                                                 // kernel.cu:760-761
	CHECK(cupt_write_ppm(out, img.data(), width, height));                                      // kernel.cu:763-778
	printf("wrote %s\n", out);
	cupt_destroy(ctx); cupt_bvh_destroy(wide); cupt_bvh2_free(bvh2); cupt_scene_destroy(scene);
	return 0;
}
