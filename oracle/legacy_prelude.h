// TEST / BASELINE INFRASTRUCTURE — prelude of the "legacy GPU" build: the reference's kernel.cu rebuilt for
// sm_100a with only its compile patches (oracle/build_ref.sh, second half).  Unlike ref_prelude.h nothing of the
// reference's device code is replaced: cuRAND XORWOW state as shipped, NUM_BOUNCES 3, the live BDPT integrator.
// The one addition is an optional ray counter at the top of trace() (kernel.cu:114), off during timed runs.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
__device__ unsigned long long g_legacy_rays;
__constant__ int g_legacy_count_on;
#ifdef __CUDA_ARCH__
#define ORACLE_COUNT_RAY() do { if (g_legacy_count_on) atomicAdd(&g_legacy_rays, 1ull); } while (0)
#else
#define ORACLE_COUNT_RAY() ((void)0)
#endif
