// b2_bottom_kernel.cuh -- __global__ wrapper of the lowest pass (b2_bottom.cuh): a 512-line tile per CTA, phase A
// (strides >= 32 lines, warp-uniform twiddles) and phase B (strides < 32 lines and the layers inside a
// line, warp-local).  sm_100a only.  Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160).
#pragma once
#include <cuda_runtime.h>

#include "b2_bottom.cuh"
#include "b2_launch.h"

namespace b2 {

// SWAP: the in-line layers in the exchanged form; ALL5: every stage of the pass has class-5 twiddles (any large transform), so the
// instantiation carries one circuit per phase and none of the coefficient-table or small-class code (decided by launch_bottom)
template <bool INV, bool SWAP, bool ALL5>
__global__ void __launch_bounds__(kBottomTeams * 128, 1) bottom_kernel(const __grid_constant__ BottomParams B) {
	extern __shared__ __align__(16) uint32_t bsmem_raw[];
	const int team = threadIdx.x >> 7, tid = threadIdx.x & 127, warp = tid >> 5, lane = tid & 31, T = 128;
	BottomSmem bs = bottom_smem_view(bsmem_raw, team);
	const UpperParams& P = B.up;
	const int nline = 5 - B.batch_bits;
	// all four phase-A layers of class 5 (the usual case at large log_n): no coefficient table, the butterflies take the
	// warp-uniform twiddle word through the static GF(16) rows, and the barrier in front of the table build goes away
	const bool nt = ALL5 || (P.cls[0] == 5 && P.cls[1] == 5 && P.cls[2] == 5 && P.cls[3] == 5);
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
		if constexpr (!ALL5) {
			if (!nt) {
				__syncthreads();
				if (active) upper_leaves(P, bs.us, tid, T);
			}
		}
		copy_wait_all();
		__syncthreads();
		if (!INV) {
			if (active) {
				if (ALL5 || nt)
					upper_pair<false, true>(P, bs.us, lane, warp, 2, 0, 3, true, false);
				else if constexpr (!ALL5)
					upper_pair<false, false>(P, bs.us, lane, warp, 2, 0, 3, true, false);
			}
			__syncthreads();
			if (active) {
				if (ALL5 || nt)
					upper_pair<false, true>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
				else if constexpr (!ALL5)
					upper_pair<false, false>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
			}
			__syncwarp();
		}
		if (active) {
#pragma unroll 1
			for (int iter = 0; iter < 2; iter++) {
#pragma unroll 1
				for (int n = 0; n < nline; n++) {
					bottom_phase_b<INV, SWAP, ALL5>(B, bs, warp, lane, iter, n);
					__syncwarp();
				}
			}
		}
		if (INV) {
			if (active) {
				if (ALL5 || nt)
					upper_pair<true, true>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
				else if constexpr (!ALL5)
					upper_pair<true, false>(P, bs.us, lane, warp << 2, 0, 0, 1, false, false);
			}
			__syncthreads();
			if (active) {
				if (ALL5 || nt)
					upper_pair<true, true>(P, bs.us, lane, warp, 2, 0, 3, false, true);
				else if constexpr (!ALL5)
					upper_pair<true, false>(P, bs.us, lane, warp, 2, 0, 3, false, true);
			}
		}
		__syncwarp();
		if (active) upper_copy<true>(P, bs.us, c, 0, INV ? warp : warp << 2, INV ? 4 : 1, 4, 0, 1, lane);
		__syncthreads();  // the slots and the tables are free again
	}
}

template <bool INV, bool SWAP, bool ALL5>
int launch_bottom_t(const BottomParams& B, cudaStream_t stream) {
	int sms = 0, dev = 0;
	const size_t smem = bottom_smem_words() * sizeof(uint32_t);
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	cudaFuncSetAttribute(bottom_kernel<INV, SWAP, ALL5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	uint64_t ctas = (uint64_t)sms * 4;  // one CTA of two teams per SM, a few chunks each for balance
	const uint64_t pairs = (B.up.n_groups + kBottomTeams - 1) / kBottomTeams;
	if (ctas > pairs) ctas = pairs;
	if (ctas == 0) return -1;
	bottom_kernel<INV, SWAP, ALL5><<<(unsigned)ctas, kBottomTeams * 128, smem, stream>>>(B);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// one translation unit per instantiation (b2_bottom_i*.cu): the six kernels compile in parallel instead of 12 minutes in a row
#define B2_BOTTOM_INSTANCE(NAME, INV, SWAP, ALL5) \
	int NAME(const BottomParams& B, cudaStream_t stream) { return launch_bottom_t<INV, SWAP, ALL5>(B, stream); }
int launch_bottom_f00(const BottomParams& B, cudaStream_t stream);  // forward, plain
int launch_bottom_f10(const BottomParams& B, cudaStream_t stream);  // forward, exchanged in-line form
int launch_bottom_f11(const BottomParams& B, cudaStream_t stream);  // forward, exchanged form, all stages of class 5
int launch_bottom_i00(const BottomParams& B, cudaStream_t stream);  // inverse ...
int launch_bottom_i10(const BottomParams& B, cudaStream_t stream);
int launch_bottom_i11(const BottomParams& B, cudaStream_t stream);

}  // namespace b2
