// b2_kernels.cu -- __global__ wrappers of the additive-NTT pass phases (b2_pass.cuh) and the
// element-wise kernels for transforms smaller than two lines.  sm_100a only.
//
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160).
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_pass.cuh"

#ifndef B2_WARPS_PER_SM
#define B2_WARPS_PER_SM 8  // resident warps per SM the pass kernels are compiled for (register budget 65536 / (32 * this))
#endif

namespace b2 {

// A CTA of T threads works through groups of 2T lines (grid-stride, persistent): coalesced
// cp.async copy-in, k butterfly layers with the bundles exchanged through swizzled shared memory,
// coalesced copy-out.  Several small CTAs per SM (4 x 64, 2 x 128 or 1 x 256 threads at ~250
// registers per thread) run out of phase, so one CTA's copies overlap another's logic.
template <bool INV, int T>
__global__ void __launch_bounds__(T, B2_WARPS_PER_SM * 32 / T) pass_kernel(const __grid_constant__ PassParams P) {
	extern __shared__ __align__(16) uint32_t smem_raw[];
	PassSmem sm = pass_smem_view(smem_raw, P.group_lines);
	const int tid = threadIdx.x;
	pass_prologue(P, sm, tid);
#pragma unroll 1
	for (uint64_t grp = blockIdx.x; grp < P.n_groups; grp += gridDim.x) {
		group_prologue(P, sm, grp, tid);
		group_copy<false>(P, sm, grp, tid);
		__syncthreads();
#pragma unroll 1
		for (int step = 0; step < P.k; step++) {
			const int i = pass_layer_bit(P, step);
#pragma unroll 1
			for (int vt = tid; vt < P.group_lines / 2; vt += T) pass_layer<INV>(P, sm, grp, vt, i, step == 0, step == P.k - 1);
			__syncthreads();
		}
		group_copy<true>(P, sm, grp, tid);
		__syncthreads();
	}
}

template <bool INV, int T>
static int launch_pass_t(const PassParams& P, cudaStream_t stream) {
	int sms = 0;
	const size_t smem = pass_smem_words(P.group_lines, 1) * sizeof(uint32_t);
	{
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
		cudaFuncSetAttribute(pass_kernel<INV, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	}
	uint64_t ctas = (uint64_t)sms * (B2_WARPS_PER_SM * 32 / T) * 4;  // a few waves of groups per CTA slot for balance
	if (ctas > P.n_groups) ctas = P.n_groups;
	if (ctas == 0) return -1;
	pass_kernel<INV, T><<<(unsigned)ctas, T, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_pass(const PassParams& P, cudaStream_t stream) {
	switch (P.cta_threads) {
		case 64: return P.inverse ? launch_pass_t<true, 64>(P, stream) : launch_pass_t<false, 64>(P, stream);
		case 128: return P.inverse ? launch_pass_t<true, 128>(P, stream) : launch_pass_t<false, 128>(P, stream);
		case 256: return P.inverse ? launch_pass_t<true, 256>(P, stream) : launch_pass_t<false, 256>(P, stream);
	}
	return -1;
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
