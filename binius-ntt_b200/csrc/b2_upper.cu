// b2_upper.cu -- __global__ wrapper of the upper passes (b2_upper.cuh): warp-uniform twiddles, constant
// multiplication from shared-memory leaf masks, radix-4 blocks held in registers.  sm_100a only.
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the layers with stride >= 32 lines.
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_upper.cuh"

namespace b2 {

// Each CTA (4 warps) walks a contiguous chunk of groups, so the leaf-mask table is rebuilt only when
// (coset row, hi) changes.  k = 4: one group (16 super-lines) per CTA at a time, two radix-4 phases
// with one exchange through shared memory.  k = 2: one group (4 super-lines) per WARP, no exchange,
// per-warp tables, warp-level synchronisation only.
template <bool INV>
__global__ void __launch_bounds__(kUpperWarps * 32, 2) upper_kernel(const __grid_constant__ UpperParams P) {
	extern __shared__ __align__(16) uint32_t usmem_raw[];
	UpperSmem sm = upper_smem_view(usmem_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const bool k4 = P.k == 4;  // uniform over the grid
	// a "worker" is the CTA (k = 4) or the warp (k = 2); its threads, its slots, its tables
	const int wt = k4 ? (int)threadIdx.x : lane, wn = k4 ? kUpperWarps * 32 : 32;
	const int slot_off = k4 ? 0 : warp * 4, nsl = k4 ? 16 : 4;
	if (!k4) sm.leaves += warp * 3 * 81, sm.tw += warp * 4;
	const uint64_t nw = k4 ? gridDim.x : (uint64_t)gridDim.x * kUpperWarps;
	const uint64_t wid = k4 ? blockIdx.x : (uint64_t)blockIdx.x * kUpperWarps + warp;
	const uint64_t per = (P.n_groups + nw - 1) / nw;
	const uint64_t g0 = per * wid, g1 = g0 + per < P.n_groups ? g0 + per : P.n_groups;
	auto sync = [&]() {
		if (k4)
			__syncthreads();
		else
			__syncwarp();
	};
	for (int i = threadIdx.x; i < 1024; i += kUpperWarps * 32) sm.consts[i] = P.consts[i];
	__syncthreads();
	uint64_t key = ~uint64_t(0);
#pragma unroll 1
	for (uint64_t grp = g0; grp < g1; grp++) {
		const UGroup c = upper_group(P, grp);
		// every warp moves its own quad of super-lines (k = 4: the quad of its first layer pair in, the quad of its
		// last layer pair out; k = 2: the whole group), so copies need no CTA barrier and overlap other warps' logic
		const bool high_first = k4 && !INV;
		upper_copy<false>(P, sm, c, slot_off, k4 ? (high_first ? warp : warp << 2) : 0, high_first ? 4 : 1, 4, lane, 32);
		const uint64_t nk = ((uint64_t)c.row << 40) ^ c.hi;
		if (nk != key) {  // uniform over the worker
			key = nk;
			upper_twiddles(P, sm, c, wt, sm.consts);
			sync();
			upper_leaves(P, sm, wt, wn);
			sync();
		}
		copy_wait_all();
		__syncwarp();
		const int nph = k4 ? 2 : 1;
#pragma unroll 1
		for (int ph = 0; ph < nph; ph++) {
			const bool high = k4 && (INV ? ph == 1 : ph == 0);
			upper_pair<INV>(P, sm, lane, high ? warp : (k4 ? warp << 2 : 0), high ? 2 : 0, slot_off, high ? 3 : 1, ph == 0, ph == nph - 1);
			if (ph + 1 < nph)
				__syncthreads();
			else
				__syncwarp();
		}
		const bool high_last = k4 && INV;
		upper_copy<true>(P, sm, c, slot_off, k4 ? (high_last ? warp : warp << 2) : 0, high_last ? 4 : 1, 4, lane, 32);
		sync();  // the slots are free again
	}
}

template <bool INV>
static int launch_upper_t(const UpperParams& P, cudaStream_t stream) {
	int sms = 0, dev = 0;
	const size_t smem = upper_smem_words(P.k) * sizeof(uint32_t);
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	cudaFuncSetAttribute(upper_kernel<INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	const uint64_t units = P.k == 4 ? P.n_groups : (P.n_groups + kUpperWarps - 1) / kUpperWarps;
	uint64_t ctas = (uint64_t)sms * 2 * 4;  // a few chunks per CTA slot for balance
	if (ctas > units) ctas = units;
	if (ctas == 0) return -1;
	upper_kernel<INV><<<(unsigned)ctas, kUpperWarps * 32, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_upper(const UpperParams& P, cudaStream_t stream) {
	if (P.k != 2 && P.k != 4) return -1;
	return P.inverse ? launch_upper_t<true>(P, stream) : launch_upper_t<false>(P, stream);
}

}  // namespace b2
