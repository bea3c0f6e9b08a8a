// ulvt/utils/common.cuh -- drop-in for the reference header of the same name (utils/common.cuh:3-32):
// the ASSERT / CUDA_CHECK macros, the two tile constants and check_gpu_capabilities(), which here
// forwards to the C ABI (b2ntt_check_gpu_capabilities, include/b2ntt.h).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <iostream>

#include "../../b2ntt.h"

#if defined(STRIP_ASSERTIONS)
#define ASSERT(cond) (cond)
#else
#define ASSERT(cond)                                                                                   \
	do {                                                                                               \
		if (!(cond)) {                                                                                 \
			std::cerr << "Assertion failed at " << __FILE__ << ":" << __LINE__ << " inside " << __FUNCTION__ \
			          << "\nCondition: " << #cond << std::endl;                                        \
			abort();                                                                                   \
		}                                                                                              \
	} while (0)
#endif

#if defined(STRIP_CUDA_CHECK)
#define CUDA_CHECK(call) (call)
#else
#define CUDA_CHECK(call)                                                                               \
	do {                                                                                               \
		cudaError_t b2_err_ = (call);                                                                  \
		if (b2_err_ != cudaSuccess)                                                                    \
			std::cerr << "Cuda error at " << __FILE__ << ":" << __LINE__ << " inside " << __FUNCTION__ \
			          << "\nError: " << cudaGetErrorString(b2_err_) << std::endl;                      \
	} while (0)
#endif

#define MAX_SHARED_MEM (1 << 15)
#define MAX_STAGES_PER_KERNEL 11

// utils/common.cu:6-43 -- device 0, true when the device can run the library
inline bool check_gpu_capabilities() { return b2ntt_check_gpu_capabilities(0) > 0; }
