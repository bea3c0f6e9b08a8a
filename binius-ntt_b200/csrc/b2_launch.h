// b2_launch.h -- host-callable launchers of the CUDA kernels (one per translation unit), used by
// the C ABI implementation in b2ntt.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

struct PassParams;
struct UpperParams;
struct BottomParams;
struct BslParams;

// element-wise path for transforms with fewer than two 128-byte lines per row
struct TinyParams {
	const uint32_t* src;
	uint32_t* dst;
	const uint32_t* consts;  // device copy of the [32][kConstPitch] constants
	int log_h, log_rate, limbs_log;
	int rows;            // columns x cosets (forward) or columns (inverse)
	int src_row_shift;   // log_rate on the forward fan-out
	uint32_t coset_mask;
	uint32_t coset0;
	int inverse;
};

int launch_pass(const PassParams& P, cudaStream_t stream);
int launch_upper(const UpperParams& U, cudaStream_t stream);  // b2_upper.cu
int launch_upper_pipe(const UpperParams& U, cudaStream_t stream);  // b2_upper_pipe.cu: k = 4, 6 as a producer / consumer pipeline
int launch_bottom(const BottomParams& B, cudaStream_t stream);  // b2_bottom.cu
int launch_bsl(const BslParams& L, cudaStream_t stream);        // b2_bsl.cu
int launch_tiny_fanout(const TinyParams& T, cudaStream_t stream);
int launch_tiny_layers(const TinyParams& T, cudaStream_t stream);

// b2_misc.cu
int launch_bitslice(uint32_t* d, uint64_t n_blocks, int width_bits, int dir, cudaStream_t stream);
int launch_mul(int field_bits, int layout, const uint32_t* a, const uint32_t* b, uint32_t* out, uint64_t n, cudaStream_t stream);
int launch_mul_repeat(int field_bits, uint32_t* io, uint64_t n_threads, int iters, cudaStream_t stream);

// Every ABI call runs on its object's device and leaves the caller's current device as it found it
// (the reference uses "the current device" implicitly, utils/common.cu:18-20; a plan here is bound to one).
struct DeviceGuard {
	int prev = -1, want;
	cudaError_t err = cudaSuccess;
	explicit DeviceGuard(int device) : want(device) {
		err = cudaGetDevice(&prev);
		if (err == cudaSuccess && prev != want) err = cudaSetDevice(want);
	}
	~DeviceGuard() {
		if (prev >= 0 && prev != want) cudaSetDevice(prev);
	}
	DeviceGuard(const DeviceGuard&) = delete;
	DeviceGuard& operator=(const DeviceGuard&) = delete;
};
// device that owns a device pointer, -1 for host / unknown memory (cudaPointerGetAttributes)
inline int pointer_device(const void* p) {
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return -1;
	}
	return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? a.device : -1;
}

// records the CUDA error text for b2ntt_last_cuda_error() and returns B2NTT_ECUDA (b2ntt.cu)
int report_cuda_error(cudaError_t e, const char* where);
// records `text` the same way and returns B2NTT_ELIMIT: a size beyond the fixed scratch of a no-allocation entry point (b2ntt.cu)
int report_limit_error(const char* text);

}  // namespace b2
