// A driver in the shape of the reference's main() (kernel.cu:565-790) — loadOBJ, buildBVH, the camera POD, the depth guard,
// the sample loop, the PPM — compiled against include/cupt_compat.h with the REFERENCE's own vec3.h / color.h / camera.h:
// the test that the source-level shim is a file that compiles and runs, not prose.  What used to be the cudaMalloc /
// cudaMemcpy block and the drawPixel loop is the handful of cupt_* calls in the middle.
//
//   cd <dir with models/> && reference_main [model.obj [width height samples [out.ppm]]]
//
// Build (tests/test_gpu_parity.py, __graft_entry__.build() when the reference tree is present):
//   g++ -std=c++17 -O2 examples/reference_main.cpp -Iinclude -iquote <reference dir> -Lcudapathtracer_b200 -lcupt -Wl,-rpath,... -o examples/reference_main
#include "cupt_compat.h"
#include <cstdlib>

#define MAX_BVH_DEPTH 64        // kernel.cu:35
#define NUM_BOUNCES 3           // kernel.cu:33

int main(int argc, char** argv)
{
	const char* model = argc > 1 ? argv[1] : "models/CornellBox-Original.obj";   // kernel.cu:591
	const int IMAGE_WIDTH = argc > 3 ? atoi(argv[2]) : 512, IMAGE_HEIGHT = argc > 3 ? atoi(argv[3]) : 512;   // kernel.cu:28-29
	const int NUM_SAMPLES = argc > 4 ? atoi(argv[4]) : 100;                      // kernel.cu:32
	const char* out = argc > 5 ? argv[5] : "image.ppm";

	loadOBJ(model, vec3(), 1);
	printf("%d verts, %d tris, %d mats, %d lights, light area %f\n", (int)verts.size(), (int)tris.size(), (int)mats.size(), (int)lights.size(), totalLightArea);

	BVH_array bvh = buildBVH();
	printf("bvh size %d depth %d\n", bvh.size, bvh.depth);
	if (bvh.depth >= MAX_BVH_DEPTH) {                                             // kernel.cu:627-631
		printf("BVH depth is too big\n");
		return 0;
	}

	camera cam;                                                                   // kernel.cu:642-648
	cam.pos = vec3(0, 1, 3);
	cam.distFromFilm = 1;
	cam.focalLength = 3;
	cam.radius = 0.0;
	cam.pxlWidth = IMAGE_WIDTH;
	cam.pxlHeight = IMAGE_HEIGHT;

	// was: cudaMalloc / cudaMemcpy of RNG states, camera, image, verts, tris, mats, lights, BVH (kernel.cu:637-697)
	cupt_context* ctx;
	if (cupt_create(0, &ctx) != 0) { printf("%s\n", cupt_last_error()); return 1; }
	cupt_set_scene(ctx, cupt_compat_scene());
	std::vector<uint32_t> rank(tris.size());
	cupt_bvh2_leaf_order(bvh.root, bvh.size, (int)tris.size(), rank.data());    // the reference's tie-break order
	cupt_bvh_opts bo = {};
	bo.tie_rank = rank.data();
	cupt_bvh* wide;
	if (cupt_bvh_build(cupt_compat_scene(), &bo, &wide) != 0) { printf("%s\n", cupt_last_error()); return 1; }
	cupt_set_bvh(ctx, wide);
	cupt_set_camera(ctx, (const cupt_camera*)&cam);
	cupt_render_opts ro;
	cupt_default_render_opts(&ro);
	ro.max_depth = NUM_BOUNCES;
	cupt_set_render_opts(ctx, &ro);

	// was: for (sampleNum = 1; sampleNum < NUM_SAMPLES; ++sampleNum) { cudaDeviceSynchronize(); drawPixel<<<...>>>(...); }  (kernel.cu:709-736)
	if (cupt_render(ctx, 0, NUM_SAMPLES - 1) != 0 || cupt_sync(ctx) != 0) { printf("%s\n", cupt_last_error()); return 1; }

	cupt_stats st;
	cupt_get_stats(ctx, &st);
	printf("rays traced: %llu\n", (unsigned long long)st.rays);

	std::vector<float> img((size_t)IMAGE_WIDTH * IMAGE_HEIGHT * 3);
	cupt_read_accum(ctx, img.data(), 1);                                          // kernel.cu:760-761
	// the writer of kernel.cu:763-778, with the reference's own color functions: x mirrored, Reinhard, gamma 1 / 2.2, (int)(c * 255)
	FILE* f = fopen(out, "w");
	if (!f) return 1;
	fprintf(f, "P3 %d %d 255\n", IMAGE_WIDTH, IMAGE_HEIGHT);
	for (int y = 0; y < IMAGE_HEIGHT; ++y)
		for (int x = IMAGE_WIDTH - 1; x >= 0; --x) {
			const float* p = &img[((size_t)y * IMAGE_WIDTH + x) * 3];
			color c = gammaCorrect(normalized(color(p[0], p[1], p[2])), 1 / 2.2);
			fprintf(f, "%d %d %d ", (int)(c.r * 255), (int)(c.g * 255), (int)(c.b * 255));
		}
	fclose(f);
	printf("wrote %s\n", out);
	cupt_destroy(ctx); cupt_bvh_destroy(wide); cupt_bvh2_free(bvh.root);
	return 0;
}
