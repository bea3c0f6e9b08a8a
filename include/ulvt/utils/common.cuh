// ulvt/utils/common.cuh -- what user code takes from the reference's header (utils/common.cuh:3-32): ASSERT and CUDA_CHECK
// (with their STRIP_* switches), the two tile constants, and check_gpu_capabilities(), which forwards to the C ABI
// (b2ntt_check_gpu_capabilities, include/b2ntt.h).  The macros are thin wrappers over two inline reporters.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <iostream>

#include "../../b2ntt.h"

namespace b2shim {
[[noreturn]] inline void assertion_failed(const char* what, const char* file, int line, const char* func) {
	std::fprintf(stderr, "Assertion failed at %s:%d inside %s\nCondition: %s\n", file, line, func, what);
	std::abort();
}
// CUDA errors are reported, not propagated (the reference's CUDA_CHECK prints and carries on)
inline void report_cuda(cudaError_t status, const char* file, int line, const char* func) {
	if (status != cudaSuccess) std::fprintf(stderr, "Cuda error at %s:%d inside %s\nError: %s\n", file, line, func, cudaGetErrorString(status));
}
}  // namespace b2shim

#if defined(STRIP_ASSERTIONS)
#define ASSERT(cond) ((void)(cond))
#else
#define ASSERT(cond) ((cond) ? (void)0 : b2shim::assertion_failed(#cond, __FILE__, __LINE__, __func__))
#endif

#if defined(STRIP_CUDA_CHECK)
#define CUDA_CHECK(call) ((void)(call))
#else
#define CUDA_CHECK(call) b2shim::report_cuda((call), __FILE__, __LINE__, __func__)
#endif

constexpr int MAX_SHARED_MEM = 32 * 1024;     // bytes of static shared memory the reference kernels assumed
constexpr int MAX_STAGES_PER_KERNEL = 11;     // not consulted by the pass scheduler

// utils/common.cu:6-43 -- device 0, true when the device can run the library.  A compiled symbol of ulvt_gpu, as in the
// reference (defined in csrc/b2_compat.cu over b2ntt_check_gpu_capabilities).
bool check_gpu_capabilities();
