// b2_bottom.cu -- __global__ wrapper of the lowest pass (b2_bottom.cuh): a 512-line tile per CTA, phase A
// (strides >= 32 lines, warp-uniform twiddles) and phase B (strides < 32 lines and the layers inside a
// line, warp-local).  sm_100a only.  Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160).
#include <cuda_runtime.h>

#include "b2_bottom.cuh"
#include "b2_launch.h"

namespace b2 {

template <bool INV, bool SWAP>
__global__ void __launch_bounds__(kBottomTeams * 128, 1) bottom_kernel(const __grid_constant__ BottomParams B) {
	extern __shared__ __align__(16) uint32_t bsmem_raw[];
	const int team = threadIdx.x >> 7, tid = threadIdx.x & 127, warp = tid >> 5, lane = tid & 31, T = 128;
	BottomSmem bs = bottom_smem_view(bsmem_raw, team);
	const UpperParams& P = B.up;
	const int nline = 5 - B.batch_bits;
	// all four phase-A layers of class 5 (the usual case at large log_n): no coefficient table, the butterflies take the
	// warp-uniform twiddle word through the static GF(16) rows, and the barrier in front of the table build goes away
	const bool nt = P.cls[0] == 5 && P.cls[1] == 5 && P.cls[2] == 5 && P.cls[3] == 5;
	for (int i = threadIdx.x; i < 1024; i += kBottomTeams * 128) bs.us.consts[i] = P.consts[i];
	__syncthreads();
	bottom_prologue(B, bs, (int)threadIdx.x, kBottomTeams * 128, bs.us.consts);
	const uint64_t per = (P.n_groups + gridDim.x - 1) / gridDim.x;
	const uint64_t g0 = per * blockIdx.x, g1 = g0 + per < P.n_groups ? g0 + per : P.n_groups;
#pragma unroll 1
	for (uint64_t gg = g0; gg < g1; gg += kBottomTeams) {
		const uint64_t grp = gg + team;
		const bool active = grp < g1;  // uniform over the team; every barrier below is met by all warps of the CTA
		const UGroup c = upper_group(P, active ? grp : gg);  // hi = tile index of the row
		if (active) {
			// every warp moves its own quad: forward the quad of phase A's first layer pair (x = warp + 4i), inverse the
			// super-lines 4 warp .. 4 warp + 3 of phase B; cp.async stays in flight while the tables are built
			upper_copy<false>(P, bs.us, c, 0, INV ? warp << 2 : warp, INV ? 1 : 4, 4, 0, 1, lane);
			upper_twiddles(P, bs.us, c, tid, bs.us.consts);
			bottom_tile_tables(B, bs, c, tid - 16, bs.us.consts);
		}
		if (!nt) {
			__syncthreads();
			if (active) upper_leaves(P, bs.us, tid, T);
		}
		copy_wait_all();
		__syncthreads();
		if (!INV) {
			if (active) {
				if (nt)
					upper_pair<false, true>(P, bs.us, lane, warp, 2, 0, 3, true, false);
				else
					upper_pair<false, false>(P, bs.us, lane, warp, 2, 0, 3, true, false);
			}
			__syncthreads();
			if (active) {
				if (nt)
					upper_pair<false, true>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
				else
					upper_pair<false, false>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
			}
			__syncwarp();
		}
		if (active) {
#pragma unroll 1
			for (int iter = 0; iter < 2; iter++) {
#pragma unroll 1
				for (int n = 0; n < nline; n++) {
					bottom_phase_b<INV, SWAP>(B, bs, warp, lane, iter, n);
					__syncwarp();
				}
			}
		}
		if (INV) {
			if (active) {
				if (nt)
					upper_pair<true, true>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
				else
					upper_pair<true, false>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
			}
			__syncthreads();
			if (active) {
				if (nt)
					upper_pair<true, true>(P, bs.us, lane, warp, 2, 0, 3, false, true);
				else
					upper_pair<true, false>(P, bs.us, lane, warp, 2, 0, 3, false, true);
			}
		}
		__syncwarp();
		if (active) upper_copy<true>(P, bs.us, c, 0, INV ? warp : warp << 2, INV ? 4 : 1, 4, 0, 1, lane);
		__syncthreads();  // the slots and the tables are free again
	}
}

template <bool INV, bool SWAP>
static int launch_bottom_t(const BottomParams& B, cudaStream_t stream) {
	int sms = 0, dev = 0;
	const size_t smem = bottom_smem_words() * sizeof(uint32_t);
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	cudaFuncSetAttribute(bottom_kernel<INV, SWAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	uint64_t ctas = (uint64_t)sms * 4;  // one CTA of two teams per SM, a few chunks each for balance
	const uint64_t pairs = (B.up.n_groups + kBottomTeams - 1) / kBottomTeams;
	if (ctas > pairs) ctas = pairs;
	if (ctas == 0) return -1;
	bottom_kernel<INV, SWAP><<<(unsigned)ctas, kBottomTeams * 128, smem, stream>>>(B);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_bottom(const BottomParams& B, cudaStream_t stream) {
	// the exchanged in-line form needs the compact image on the in-line side of the pass (its element permutation is undone by the
	// chunk-wise loads / stores of the compact words)
	const bool swap = B.swap_inline && (B.up.inverse ? !B.up.src_sliced : !B.up.dst_sliced);
	if (B.up.inverse) return swap ? launch_bottom_t<true, true>(B, stream) : launch_bottom_t<true, false>(B, stream);
	return swap ? launch_bottom_t<false, true>(B, stream) : launch_bottom_t<false, false>(B, stream);
}

}  // namespace b2
