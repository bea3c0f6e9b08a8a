// b2_bsl.cu -- __global__ wrapper of the block-sliced lowest pass (b2_bsl.cuh): a tile of 32 consecutive 4 KB blocks per CTA,
// the stages below one block with twiddle tables shared by the whole tile.  sm_100a only.
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for those stages.
#include <cuda_runtime.h>

#include "b2_bsl.cuh"
#include "b2_launch.h"

namespace b2 {

template <bool INV>
__global__ void __launch_bounds__(kBslThreads, 1) bsl_kernel(const __grid_constant__ BslParams P) {
	extern __shared__ __align__(16) uint32_t bslsm_raw[];
	BslSmem sm = bsl_smem_view(bslsm_raw);
	const int tid = threadIdx.x, S = bsl_stages(P);
	for (int i = tid; i < 1024; i += kBslThreads) sm.consts[i] = P.consts[i];
	__syncthreads();
	bsl_prologue(P, sm, tid, kBslThreads);
	const uint64_t per = (P.n_tiles + gridDim.x - 1) / gridDim.x;
	const uint64_t t0 = per * blockIdx.x, t1 = t0 + per < P.n_tiles ? t0 + per : P.n_tiles;
#pragma unroll 1
	for (uint64_t tile = t0; tile < t1; tile++) {
		const BslTile t = bsl_tile_of(P, tile);
		bsl_copy<false>(P, sm, t, tid);
		bsl_tile_twiddles(P, sm, t, tid);
		copy_wait_all();
		__syncthreads();
		bsl_transpose(sm, tid);
		__syncthreads();
#pragma unroll 1
		for (int step = 0; step < S; step++) {
			const int s = INV ? step : S - 1 - step;
			if (bsl_has_table(P, s)) {
				bsl_build_tables(P, sm, s, tid);
				__syncthreads();
			}
			bsl_stage<INV>(P, sm, s, tid);
			__syncthreads();
		}
		bsl_transpose(sm, tid);
		__syncthreads();
		bsl_copy<true>(P, sm, t, tid);
		__syncthreads();
	}
}

template <bool INV>
static int launch_bsl_t(const BslParams& P, cudaStream_t stream) {
	static int sms_of[64] = {};
	int dev = 0;
	const size_t smem = bsl_smem_words() * sizeof(uint32_t);
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64) return -1;
	if (!sms_of[dev]) {  // once per device
		int sms = 0;
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
		if (cudaFuncSetAttribute(bsl_kernel<INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
		sms_of[dev] = sms;
	}
	uint64_t ctas = (uint64_t)sms_of[dev] * 4;  // a few chunks per SM for balance
	if (ctas > P.n_tiles) ctas = P.n_tiles;
	if (ctas == 0) return -1;
	bsl_kernel<INV><<<(unsigned)ctas, kBslThreads, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_bsl(const BslParams& P, cudaStream_t stream) {
	if (P.tiles_per_row_log < 0 || (P.limbs_log != 0 && P.limbs_log != 2)) return -1;
	return P.inverse ? launch_bsl_t<true>(P, stream) : launch_bsl_t<false>(P, stream);
}

}  // namespace b2
