// b2_wupper.cu -- upper passes with WARP-OWNED groups: the same layers, twiddle tables and radix-4 units as
// b2_upper.cu, but every warp of the CTA walks its own groups of 128 lines (2^k super-lines of 2^(7-k) lines)
// from copy-in to copy-out with warp-level synchronisation only.  The coefficient table of a (coset row, hi)
// block is shared by the eight warps of the CTA and rebuilt behind a CTA barrier only when the block changes, so
// the warps drift apart and one warp's copy latency is covered by the logic of the other warp of its SM
// sub-partition (b2_upper.cu keeps four warps in lock-step on one 512-line group: ALU pipe 63-73 % busy, r1p).
// sm_100a only.  Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the layers with stride >= 32 lines.
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_upper.cuh"

namespace b2 {

constexpr int kWUpperWarps = 8;

B2_HD size_t wupper_smem_words() { return (size_t)kLeafTableWords + 64 + 1024 + (size_t)kWUpperWarps * kWGroupLines * kUSlotWords; }

template <bool INV>
__global__ void __launch_bounds__(kWUpperWarps * 32, 1) wupper_kernel(const __grid_constant__ UpperParams P) {
	extern __shared__ __align__(16) uint32_t wsmem_raw[];
	UpperSmem sm = upper_smem_view(wsmem_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	sm.slots += (size_t)warp * kWGroupLines * kUSlotWords;  // this warp's 128 slots
	for (int i = threadIdx.x; i < 1024; i += kWUpperWarps * 32) sm.consts[i] = P.consts[i];
	const int k = P.k, ls = P.ls, nph = k >> 1, bshift = P.q0 - ls;  // 2^bshift groups share (coset row, hi)
	const int u = lane >> ls;  // radix-4 unit of the lane inside a layer pair
	const uint64_t per = (P.n_groups + gridDim.x - 1) / gridDim.x;
	const uint64_t G0 = per * blockIdx.x, G1 = G0 + per < P.n_groups ? G0 + per : P.n_groups;
	uint64_t g = G0;
#pragma unroll 1
	while (g < G1) {
		const uint64_t blk_end = ((g >> bshift) + 1) << bshift;
		const uint64_t seg1 = blk_end < G1 ? blk_end : G1;
		{
			const UGroup c = upper_group(P, g);
			__syncthreads();  // every warp is done with the previous block's table (first round: consts are in place)
			upper_twiddles(P, sm, c, (int)threadIdx.x, sm.consts);
			__syncthreads();
			upper_leaves(P, sm, (int)threadIdx.x, kWUpperWarps * 32);
			__syncthreads();
		}
#pragma unroll 1
		for (uint64_t gg = g + warp; gg < seg1; gg += kWUpperWarps) {
			const UGroup c = upper_group(P, gg);
			wupper_copy<false>(P, sm.slots, c, lane);
			copy_wait_all();
			__syncwarp();
#pragma unroll 1
			for (int e = 0; e < nph; e++) {
				const int j = INV ? nph - 1 - e : e, b1 = k - 1 - 2 * j, b0 = b1 - 1;
				const int xbase = ((u >> b0) << (b1 + 1)) | (u & ((1 << b0) - 1));
				upper_pair<INV>(P, sm, lane, xbase, b0, 0, b1, e == 0, e == nph - 1);
				__syncwarp();
			}
			wupper_copy<true>(P, sm.slots, c, lane);
			__syncwarp();
		}
		g = seg1;
	}
}

template <bool INV>
static int launch_wupper_t(const UpperParams& P, cudaStream_t stream) {
	static int sms_of[64] = {};
	int dev = 0;
	const size_t smem = wupper_smem_words() * sizeof(uint32_t);
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64) return -1;
	if (!sms_of[dev]) {  // once per device: attribute queries and the opt-in shared-memory size
		int sms = 0;
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
		if (cudaFuncSetAttribute(wupper_kernel<INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
		sms_of[dev] = sms;
	}
	uint64_t ctas = (uint64_t)sms_of[dev];
	const uint64_t units = (P.n_groups + kWUpperWarps - 1) / kWUpperWarps;
	if (ctas > units) ctas = units;
	if (ctas == 0) return -1;
	wupper_kernel<INV><<<(unsigned)ctas, kWUpperWarps * 32, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_wupper(const UpperParams& P, cudaStream_t stream) {
	if (P.k != 2 && P.k != 4 && P.k != 6) return -1;
	if (P.ls != 7 - P.k || P.q0 < P.ls) return -1;
	if ((P.lpitch != kLeafRows5 && P.lpitch != kLeafRows4) || (P.k == 6 && P.lpitch != kLeafRows4)) return -1;
	return P.inverse ? launch_wupper_t<true>(P, stream) : launch_wupper_t<false>(P, stream);
}

}  // namespace b2
