// Output stage on the device: the tonemap of the reference's PPM writer (kernel.cu:763-778,
// color.h:59-71) as a kernel, so that 3 bytes per pixel cross PCIe instead of 12, and the traversal-cost heat map the
// reference draws from its per-ray step counter (kernel.cu:121,132,152; `keepers/sponza bvh.ppm`).
#pragma once
#include "kernels.cuh"

namespace cupt {

// mean = sum / samples; Reinhard c/(c+1) (color.h:59-62); gamma with the exponent (float)(1/2.2) in double pow
// (color.h:68-71, kernel.cu:774); (int)(c*255) truncation; x mirrored like kernel.cu:768 when `mirror`
__global__ void __launch_bounds__(256) k_tonemap_rgb8(const float* __restrict__ accum, const float inv_samples, const int W, const int H,
                                                      const int mirror, uint8_t* __restrict__ out)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= W * H) return;
	const int y = p / W, x = p - y * W;
	const float* a = accum + 3 * (size_t)p;
	const double g = (double)(float)(1 / 2.2);
	uint8_t* o = out + 3 * ((size_t)y * W + (mirror ? W - 1 - x : x));
#pragma unroll
	for (int k = 0; k < 3; ++k) {
		double c = (double)__fmul_rn(a[k], inv_samples);
		c = c / (c + 1);
		c = pow(c, g);
		const int v = (int)(c * 255);
		o[k] = (uint8_t)min(255, max(0, v));
	}
}

// per pixel: wide-BVH nodes visited and primitives tested by the primary ray (one thread per pixel, the single-ray
// traversal of traverse.cuh; a debug output, not a hot path)
__global__ void __launch_bounds__(128) k_heatmap(const DevScene sc, const PassParams pp, float* __restrict__ nodes_out, float* __restrict__ prims_out)
{
	extern __shared__ uint2 hm_stack[];
	const int W = pp.cam.pxlWidth, H = pp.cam.pxlHeight;
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= W * H) return;
	const int y = p / W, x = p - y * W;
	f3 o, d;
	camera_ray(pp.cam, x, y, 0.5f, 0.5f, o, d);
	Hit hit;
	uint32_t n_nodes = 0, n_tris = 0;
	traverse_wide<true>(sc.nodes, sc.prims, o, d, hit, hm_stack + threadIdx.x, blockDim.x, sc.stack_cap, n_nodes, n_tris);
	if (nodes_out) nodes_out[p] = (float)n_nodes;
	if (prims_out) prims_out[p] = (float)n_tris;
}

}  // namespace cupt
