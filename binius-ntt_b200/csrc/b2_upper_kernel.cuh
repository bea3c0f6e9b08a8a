// b2_upper_kernel.cuh -- __global__ wrapper of the upper passes (b2_upper.cuh): warp-uniform twiddles, constant
// multiplication from shared-memory leaf masks, radix-4 blocks held in registers.  sm_100a only.
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the layers with stride >= 32 lines.
#pragma once
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_upper.cuh"

namespace b2 {

// Each CTA (4 warps) walks a contiguous chunk of groups, so the leaf-mask table is rebuilt only when
// (coset row, hi) changes.  k = 4: one group (16 super-lines) per CTA at a time, two radix-4 phases
// with one exchange through shared memory.  k = 2: one group (4 super-lines) per WARP, no exchange,
// per-warp tables, warp-level synchronisation only.
// CLS: 0 = any mix of twiddle classes; 4 / 5 = every layer of the pass has that class (k >= 4): one circuit, no dispatch
template <bool INV, int CLS>
__global__ void __launch_bounds__(kUpperWarps * 32, 2) upper_kernel(const __grid_constant__ UpperParams P) {
	extern __shared__ __align__(16) uint32_t usmem_raw[];
	UpperSmem sm = upper_smem_view(usmem_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < 1024; i += kUpperWarps * 32) sm.consts[i] = P.consts[i];
	upper_rows16_fill(sm, (int)threadIdx.x, kUpperWarps * 32);
	__syncthreads();
	if (CLS != 0 || P.k >= 4) {  // the lean instantiations are launched for k >= 4 only
		// The CTA works on one group of 2^k super-lines (512 lines) at a time, k / 2 layer pairs with a CTA barrier between
		// them.  A layer pair splits the group into 2^(k-2) units of 4 super-lines; a warp runs 2^(k-4) units at once,
		// lane = (unit of the warp) << ls | (line of the super-line).  Pair j (bits k-1-2j, k-2-2j) of unit u touches
		// x = u_hi | a << (k-2-2j) | u_lo, a = 0..3: the low k-2-2j bits of u below the pair, the rest above.
		const int k = P.k, ls = P.ls, upw = 1 << (k - 4), nph = k >> 1;
		const int u = warp * upw + (lane >> ls);
		const uint64_t per = (P.n_groups + gridDim.x - 1) / gridDim.x;
		const uint64_t g0 = per * blockIdx.x, g1 = g0 + per < P.n_groups ? g0 + per : P.n_groups;
		uint64_t key = ~uint64_t(0);
#pragma unroll 1
		for (uint64_t grp = g0; grp < g1; grp++) {
			const UGroup c = upper_group(P, grp);
			// every warp moves the super-lines of its own units: those of its first layer pair in, those of its last layer
			// pair out (first pair of the forward transform: x = a << (k-2) | u; last: x = u << 2 | a), so copies need
			// no CTA barrier and overlap other warps' logic
			if (!INV)
				upper_copy<false>(P, sm, c, 0, warp * upw, 1, upw, 1 << (k - 2), 4, lane);
			else
				upper_copy<false>(P, sm, c, 0, warp * upw * 4, 1, upw * 4, 0, 1, lane);
			const uint64_t nk = ((uint64_t)c.row << 40) ^ c.hi;
			if (nk != key) {  // uniform over the CTA
				key = nk;
				upper_twiddles(P, sm, c, (int)threadIdx.x, sm.consts);
				__syncthreads();
				upper_leaves(P, sm, (int)threadIdx.x, kUpperWarps * 32);
				__syncthreads();
			}
			copy_wait_all();
			__syncwarp();
#pragma unroll 1
			for (int e = 0; e < nph; e++) {
				const int j = INV ? nph - 1 - e : e, b1 = k - 1 - 2 * j, b0 = b1 - 1;
				const int xbase = ((u >> b0) << (b1 + 1)) | (u & ((1 << b0) - 1));
				upper_pair<INV, 0, CLS>(P, sm, lane, xbase, b0, 0, b1, e == 0, e == nph - 1);
				if (e + 1 < nph)
					__syncthreads();
				else
					__syncwarp();
			}
			if (!INV)
				upper_copy<true>(P, sm, c, 0, warp * upw * 4, 1, upw * 4, 0, 1, lane);
			else
				upper_copy<true>(P, sm, c, 0, warp * upw, 1, upw, 1 << (k - 2), 4, lane);
			__syncthreads();  // the slots are free again
		}
	} else if constexpr (CLS == 0) {
		// k = 2: one group (4 super-lines of 32 lines) per WARP, per-warp tables (3 twiddles) and slots, warp-level sync only
		sm.leaves += warp * 3 * P.lpitch, sm.tw += warp * 4;
		const int slot_off = warp * 4;
		const uint64_t nw = (uint64_t)gridDim.x * kUpperWarps, wid = (uint64_t)blockIdx.x * kUpperWarps + warp;
		const uint64_t per = (P.n_groups + nw - 1) / nw;
		const uint64_t g0 = per * wid, g1 = g0 + per < P.n_groups ? g0 + per : P.n_groups;
		uint64_t key = ~uint64_t(0);
#pragma unroll 1
		for (uint64_t grp = g0; grp < g1; grp++) {
			const UGroup c = upper_group(P, grp);
			upper_copy<false>(P, sm, c, slot_off, 0, 1, 4, 0, 1, lane);
			const uint64_t nk = ((uint64_t)c.row << 40) ^ c.hi;
			if (nk != key) {  // uniform over the warp
				key = nk;
				upper_twiddles(P, sm, c, lane, sm.consts);
				__syncwarp();
				upper_leaves(P, sm, lane, 32);
			}
			copy_wait_all();
			__syncwarp();
			upper_pair<INV>(P, sm, lane, 0, 0, slot_off, 1, true, true);
			__syncwarp();
			upper_copy<true>(P, sm, c, slot_off, 0, 1, 4, 0, 1, lane);
			__syncwarp();
		}
	}
}

template <bool INV, int CLS>
int launch_upper_t(const UpperParams& P, cudaStream_t stream) {
	int sms = 0, dev = 0;
	const size_t smem = upper_smem_words(P.k) * sizeof(uint32_t);
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	cudaFuncSetAttribute(upper_kernel<INV, CLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	const uint64_t units = P.k >= 4 ? P.n_groups : (P.n_groups + kUpperWarps - 1) / kUpperWarps;
	uint64_t ctas = (uint64_t)sms * 2 * 4;  // a few chunks per CTA slot for balance
	if (ctas > units) ctas = units;
	if (ctas == 0) return -1;
	upper_kernel<INV, CLS><<<(unsigned)ctas, kUpperWarps * 32, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// one translation unit per instantiation (b2_upper_*.cu), compiled in parallel
#define B2_UPPER_INSTANCE(NAME, INV, CLS) \
	int NAME(const UpperParams& P, cudaStream_t stream) { return launch_upper_t<INV, CLS>(P, stream); }
int launch_upper_f0(const UpperParams& P, cudaStream_t stream);
int launch_upper_f4(const UpperParams& P, cudaStream_t stream);
int launch_upper_f5(const UpperParams& P, cudaStream_t stream);
int launch_upper_i0(const UpperParams& P, cudaStream_t stream);
int launch_upper_i4(const UpperParams& P, cudaStream_t stream);
int launch_upper_i5(const UpperParams& P, cudaStream_t stream);

}  // namespace b2
