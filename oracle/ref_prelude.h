// TEST INFRASTRUCTURE — prelude force-included ahead of the (sed-patched) reference kernel.cu when
// oracle/build_ref.sh builds oracle/_ref/libcupt_ref.so.  Nothing here is product code.
//
// It (a) makes cuRAND's device generators callable on the host (SURVEY.md §8c patch 7),
//    (b) replaces the reference's `curandState` by a switchable generator so that the one RNG entry
//        point of the reference, nrand() (kernel.cu:56-59), can draw either from XORWOW exactly as
//        shipped (kernel.cu:532: seed 1234, subsequence = pixel index) or from a counter-based
//        Philox4x32-10 stream keyed by (seed; pixel, sample) that the CUDA path can reproduce under
//        any execution order,
//    (c) declares the run-time bounce count and the per-thread ray counter used by the patches.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#define QUALIFIERS static __forceinline__ __host__ __device__
#include <curand_kernel.h>

struct OracleRng {
	int kind;  // 0 = XORWOW (as shipped), 1 = Philox4x32-10 keyed stream
	curandStateXORWOW_t xorwow;
	curandStatePhilox4_32_10_t philox;
};

static inline __host__ __device__ void curand_init(unsigned long long seed, unsigned long long subsequence,
                               unsigned long long offset, OracleRng* s)
{
	if (s->kind == 0)
		curand_init(seed, subsequence, offset, &s->xorwow);
	else
		curand_init(seed, subsequence, offset, &s->philox);
}
static inline __host__ __device__ float curand_uniform(OracleRng* s)
{
	if (s->kind == 0)
		return curand_uniform(&s->xorwow);
	return curand_uniform(&s->philox);
}
// every later mention of curandState in kernel.cu now names the switchable generator
#define curandState OracleRng

// kernel.cu:33 made run-time (patch 8) and the added ray counter (patch 9).  The device pass of nvcc
// still parses the __host__ __device__ bodies; it is never executed, so it sees constants instead.
static int g_oracle_num_bounces = 3;
static thread_local unsigned long long g_oracle_ray_count = 0;
#ifdef __CUDA_ARCH__
#define ORACLE_NUM_BOUNCES 3
#define ORACLE_COUNT_RAY() ((void)0)
#else
#define ORACLE_NUM_BOUNCES g_oracle_num_bounces
#define ORACLE_COUNT_RAY() (++g_oracle_ray_count)
#endif
