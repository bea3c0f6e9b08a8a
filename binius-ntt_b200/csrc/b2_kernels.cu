// b2_kernels.cu -- __global__ wrappers of the additive-NTT pass phases (b2_pass.cuh) and the
// element-wise kernels for transforms smaller than two lines.  sm_100a only.
//
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160).
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_pass.cuh"

namespace b2 {

// One CTA = 256 threads = up to 512 lines (64 KiB) of the data set, grouped as 2^(9-k) tiles of
// 2^k lines.  Every thread keeps two bundles (64 registers) and runs one butterfly per layer;
// between layers the bundles are exchanged through swizzled shared memory.
template <bool INV>
__global__ void __launch_bounds__(kCtaThreads, 1) pass_kernel(const __grid_constant__ PassParams P) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	PassSmem& sm = *reinterpret_cast<PassSmem*>(smem_raw);
	const int tid = threadIdx.x;
	pass_prologue(P, sm, tid, kCtaThreads);
	__syncthreads();
	Bundles st;
#pragma unroll 1
	for (int step = 0; step < P.k; step++) {
		const int i = pass_layer_bit(P, step);
		pass_layer<INV>(P, sm, blockIdx.x, tid, i, step == 0, step == P.k - 1, st);
		__syncthreads();
	}
}

int launch_pass(const PassParams& P, cudaStream_t stream) {
	static bool attr_set = false;
	if (!attr_set) {
		cudaFuncSetAttribute(pass_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PassSmem));
		cudaFuncSetAttribute(pass_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PassSmem));
		attr_set = true;
	}
	const uint64_t tiles_per_cta = kCtaThreads >> (P.k - 1);
	const uint64_t ctas = (P.n_tiles + tiles_per_cta - 1) / tiles_per_cta;
	if (ctas == 0 || ctas > 0x7fffffffull) return -1;
	if (P.inverse)
		pass_kernel<true><<<(unsigned)ctas, kCtaThreads, sizeof(PassSmem), stream>>>(P);
	else
		pass_kernel<false><<<(unsigned)ctas, kCtaThreads, sizeof(PassSmem), stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ------------------------------------------------------------------------------------------------
// tiny transforms (fewer than two lines per row): one layer per launch, compact arithmetic.
// rows x 2^(log_h-1) butterflies; limb l of an element meets the same twiddle.
// ------------------------------------------------------------------------------------------------
__global__ void tiny_fanout_kernel(TinyParams T) {
	// out[row][i] = in[row >> log_rate][i]
	const uint64_t n = (uint64_t)T.rows << (T.log_h + T.limbs_log);
	for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n; g += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t row = g >> (T.log_h + T.limbs_log), off = g & (((uint64_t)1 << (T.log_h + T.limbs_log)) - 1);
		T.dst[g] = T.src[((row >> T.src_row_shift) << (T.log_h + T.limbs_log)) + off];
	}
}

__global__ void tiny_layer_kernel(TinyParams T, int stage) {
	const uint64_t per_row = (uint64_t)1 << (T.log_h - 1);
	const uint64_t n = (uint64_t)T.rows * per_row;
	const int limbs = 1 << T.limbs_log;
	for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n; g += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t row = (uint32_t)(g / per_row);
		const uint32_t t = (uint32_t)(g % per_row);
		const uint32_t j = t >> stage, k = t & ((1u << stage) - 1);
		const uint32_t u = (j << (stage + 1)) | k, v = u | (1u << stage);
		const uint32_t ind = ((T.coset0 + (row & T.coset_mask)) << (T.log_h - 1 - stage)) | j;
		const uint32_t w = xor_select(T.consts + stage * kConstPitch, ind);
		uint32_t* base = T.dst + ((uint64_t)row << (T.log_h + T.limbs_log));
		for (int l = 0; l < limbs; l++) {
			uint32_t a = base[(u << T.limbs_log) + l], b = base[(v << T.limbs_log) + l];
			if (T.inverse) {
				b ^= a;
				a ^= gf32_mul(w, b);
			} else {
				a ^= gf32_mul(w, b);
				b ^= a;
			}
			base[(u << T.limbs_log) + l] = a, base[(v << T.limbs_log) + l] = b;
		}
	}
}

static unsigned tiny_blocks(const TinyParams& T) {
	const uint64_t n = (uint64_t)T.rows << (T.log_h + T.limbs_log);
	unsigned blocks = (unsigned)((n + 255) / 256);
	return blocks > 148 * 8 ? 148 * 8 : blocks;
}

int launch_tiny_fanout(const TinyParams& T, cudaStream_t stream) {
	if (T.src == T.dst && !T.src_row_shift) return 0;
	tiny_fanout_kernel<<<tiny_blocks(T), 256, 0, stream>>>(T);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_tiny_layers(const TinyParams& T, cudaStream_t stream) {
	for (int n = 0; n < T.log_h; n++) {
		const int stage = T.inverse ? n : T.log_h - 1 - n;
		tiny_layer_kernel<<<tiny_blocks(T), 256, 0, stream>>>(T, stage);
	}
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace b2
