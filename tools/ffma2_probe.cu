// Developer probe (GPU box): issue throughput of scalar FFMA against packed FFMA2 (fma.rn.f32x2) on sm_100a, and of the
// node step's per-child pattern (6 PRMT + 6 FFMA + 2 FMNMX3 ... vs 6 PRMT + 3 FFMA2 ...).  Build and run:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/ffma2_probe tools/ffma2_probe.cu && /tmp/ffma2_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

template <int MODE>   // 0: 16 independent scalar FFMA chains; 1: 8 independent FFMA2 chains (the same 16 values); 2: FFMA + PRMT mix; 3: FFMA2 + PRMT mix
__global__ void __launch_bounds__(256) k(float* out, int iters, float s, float o, uint32_t w)
{
	float a[16];
#pragma unroll
	for (int i = 0; i < 16; ++i) a[i] = (float)(threadIdx.x + i);
	f32x2 p[8];
#pragma unroll
	for (int i = 0; i < 8; ++i) p[i] = pack2(a[2 * i], a[2 * i + 1]);
	const f32x2 s2 = pack2(s, s), o2 = pack2(o, o);
	uint32_t acc = w;
	for (int it = 0; it < iters; ++it) {
		if (MODE == 0 || MODE == 2) {
#pragma unroll
			for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, o);
		} else {
#pragma unroll
			for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], s2, o2);
		}
		if (MODE >= 2) {
#pragma unroll
			for (int i = 0; i < 16; ++i) acc = __byte_perm(acc, w + i, 0x7604 + (i & 3) * 16);
		}
	}
	float r = 0;
	if (MODE == 0 || MODE == 2) { for (int i = 0; i < 16; ++i) r += a[i]; }
	else { for (int i = 0; i < 8; ++i) { float lo, hi; unpack2(p[i], lo, hi); r += lo + hi; } }
	out[blockIdx.x * blockDim.x + threadIdx.x] = r + __uint_as_float(acc);
}

template <int MODE>
float run(float* d, int iters)
{
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	k<MODE><<<148 * 8, 256>>>(d, 16, 1.0001f, 0.5f, 0x01020304u);
	cudaEventRecord(a);
	k<MODE><<<148 * 8, 256>>>(d, iters, 1.0001f, 0.5f, 0x01020304u);
	cudaEventRecord(b); cudaEventSynchronize(b);
	float ms; cudaEventElapsedTime(&ms, a, b);
	return ms;
}

int main()
{
	float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
	const int iters = 20000;
	const double fmas = 148.0 * 8 * 256 * 16.0 * iters;   // scalar-equivalent FMAs in every mode
	const float t0 = run<0>(d, iters), t1 = run<1>(d, iters), t2 = run<2>(d, iters), t3 = run<3>(d, iters);
	printf("scalar FFMA x16        : %7.2f ms  %6.2f TFMA/s\n", t0, fmas / t0 * 1e-9);
	printf("packed FFMA2 x8        : %7.2f ms  %6.2f TFMA/s (scalar-equivalent)\n", t1, fmas / t1 * 1e-9);
	printf("16 FFMA + 16 PRMT      : %7.2f ms\n", t2);
	printf(" 8 FFMA2 + 16 PRMT     : %7.2f ms\n", t3);
	return 0;
}
