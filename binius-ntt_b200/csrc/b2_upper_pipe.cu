// b2_upper_pipe.cu -- the upper passes (k = 4, 6) as a producer / consumer pipeline: one CTA per SM, a producer warp that
// streams groups into a ring of three slot buffers with asynchronous copies (cp.async, completion counted on an mbarrier),
// and two teams of four warps that run the butterfly layers of b2_upper.cuh on alternate groups.  sm_100a only.
//
// Why: in upper_kernel (b2_upper.cu) a CTA loads a group, waits, computes, stores; with two CTAs per SM the memory phases
// overlap the other CTA's logic only by accident.  Measured on the headline transform (profiles/README.md, round 2): with the
// global copies removed the three upper passes take 2.83 / 3.41 / 3.36 ms instead of 3.80 / 4.32 / 3.92 -- 2.4 ms of exposed
// memory time.  Here the group after next is already in flight while a team computes.
//
// Shared memory: 3 x 72 KB of slots leave no room for coefficient tables, so every layer takes its warp-uniform twiddle WORD
// through the static GF(16) rows (butterfly_lane_dispatch, b2_constmat.cuh).
//
// Ring protocol (groups i = 0, 1, .. of the CTA's chunk; buffer i % 3, team i & 1):
//   full[b]  (count 32): every producer lane issues its 128 16-byte copies of the group and lets the barrier count their
//            completion (cp.async.mbarrier.arrive.noinc).  [cp.async.bulk per 128-byte line was tried first: the padded slots
//            allow no larger bulk copy, and 512 bulk copies per group issued one lane at a time made the producer the limit:
//            7.5 ms per pass];
//   empty[b] (count 4): each warp of the team arrives after copying its part of the result out of the buffer.
// The teams drift half a group apart by themselves: a team that finds its next group not yet loaded waits once, and from then
// on its loads are issued (when the OTHER team frees a buffer) half a period before it needs them.
//
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the layers with stride >= 32 lines, like b2_upper.cu.
#include <cuda_runtime.h>

#include "b2_launch.h"
#include "b2_upper.cuh"

namespace b2 {

constexpr int kPipeTeams = 2, kPipeBuffers = 3;
constexpr int kPipeThreads = kPipeTeams * 128 + 32;  // + the producer warp

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// the barrier receives one arrival from this thread when all its earlier cp.async copies have landed
__device__ __forceinline__ void mbar_arrive_on_copies(uint64_t* bar) {
	asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile(
	    "{\n"
	    ".reg .pred p;\n"
	    "WAIT_LOOP:\n"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
	    "@p bra WAIT_DONE;\n"
	    "bra WAIT_LOOP;\n"
	    "WAIT_DONE:\n"
	    "}\n" ::"r"(smem_addr(bar)),
	    "r"(parity)
	    : "memory");
}
__device__ __forceinline__ void team_barrier(int team) { asm volatile("bar.sync %0, 128;" ::"r"(team + 1) : "memory"); }

struct PipeSmem {
	uint32_t* consts;  // [32][32]
	Coef4* rows16;     // [4][16]
	uint32_t* tw;      // [teams][64]
	uint64_t* full;    // [3]
	uint64_t* empty;   // [3]
	uint32_t* slots;   // [3][512][kUSlotWords]
};
constexpr size_t kPipeSlotWords = (size_t)512 * kUSlotWords;
constexpr size_t kPipeSmemWords = 1024 + 256 + kPipeTeams * 64 + 16 + kPipeBuffers * kPipeSlotWords;
__device__ __forceinline__ PipeSmem pipe_smem_view(uint32_t* base) {
	PipeSmem sm;
	sm.consts = base;
	sm.rows16 = reinterpret_cast<Coef4*>(base + 1024);
	sm.tw = base + 1024 + 256;
	sm.full = reinterpret_cast<uint64_t*>(sm.tw + kPipeTeams * 64);
	sm.empty = sm.full + 4;
	sm.slots = sm.tw + kPipeTeams * 64 + 16;
	return sm;
}

template <bool INV>
__global__ void __launch_bounds__(kPipeThreads, 1) upper_pipe_kernel(const __grid_constant__ UpperParams P) {
	extern __shared__ __align__(128) uint32_t psmem_raw[];
	const PipeSmem ps = pipe_smem_view(psmem_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < 1024; i += kPipeThreads) ps.consts[i] = P.consts[i];
	for (int i = threadIdx.x; i < 64; i += kPipeThreads) ps.rows16[i] = lane_row_entry(i);
	if (threadIdx.x == 0) {
		for (int b = 0; b < kPipeBuffers; b++) mbar_init(ps.full + b, 32), mbar_init(ps.empty + b, 4);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	const uint64_t per = (P.n_groups + gridDim.x - 1) / gridDim.x;
	const uint64_t g0 = per * blockIdx.x, g1 = g0 + per < P.n_groups ? g0 + per : P.n_groups;
	const int k = P.k, ls = P.ls;
	if (warp == kPipeTeams * 4) {
		// ---- producer: group i into buffer i % 3 once the team that used it three groups ago has drained it
		int buf = 0;
		uint32_t round = 0;  // how many times the ring has wrapped
#pragma unroll 1
		for (uint64_t grp = g0; grp < g1; grp++) {
			if (round) mbar_wait(ps.empty + buf, (round - 1) & 1);
			const UGroup c = upper_group(P, grp);
			UpperSmem sb;
			sb.slots = ps.slots + buf * kPipeSlotWords;
			upper_copy<false>(P, sb, c, 0, 0, 1, 1 << k, 0, 1, lane);
			mbar_arrive_on_copies(ps.full + buf);
			if (++buf == kPipeBuffers) buf = 0, round++;
		}
		return;
	}
	// ---- consumers: team t runs the groups g0 + t, g0 + t + 2, ..
	const int team = warp >> 2, tw = warp & 3, ttid = threadIdx.x & 127;
	UpperSmem sm;
	sm.leaves = nullptr, sm.consts = ps.consts, sm.rows16 = ps.rows16, sm.tw = ps.tw + team * 64;
	const int upw = 1 << (k - 4), nph = k >> 1;
	const int u = tw * upw + (lane >> ls);
	uint64_t key = ~uint64_t(0);
	uint64_t i = team;
#pragma unroll 1
	for (uint64_t grp = g0 + team; grp < g1; grp += kPipeTeams, i += kPipeTeams) {
		const int buf = (int)(i % kPipeBuffers);
		const uint32_t round = (uint32_t)(i / kPipeBuffers);
		sm.slots = ps.slots + buf * kPipeSlotWords;
		const UGroup c = upper_group(P, grp);
		const uint64_t nk = ((uint64_t)c.row << 40) ^ c.hi;
		if (nk != key) {  // uniform over the team
			key = nk;
			team_barrier(team);  // the previous group's butterflies have read the old words
			upper_twiddles(P, sm, c, ttid, ps.consts);
			team_barrier(team);
		}
		mbar_wait(ps.full + buf, round & 1);
#pragma unroll 1
		for (int e = 0; e < nph; e++) {
			const int j = INV ? nph - 1 - e : e, b1 = k - 1 - 2 * j, b0 = b1 - 1;
			const int xbase = ((u >> b0) << (b1 + 1)) | (u & ((1 << b0) - 1));
			upper_pair<INV, 2>(P, sm, lane, xbase, b0, 0, b1, e == 0, e == nph - 1);
			if (e + 1 < nph)
				team_barrier(team);
			else
				__syncwarp();
		}
		if (!INV)
			upper_copy<true>(P, sm, c, 0, tw * upw * 4, 1, upw * 4, 0, 1, lane);
		else
			upper_copy<true>(P, sm, c, 0, tw * upw, 1, upw, 1 << (k - 2), 4, lane);
		// the warp's reads of the buffer are complete (the data sits in registers on its way out); order its generic-proxy
		// accesses before the producer's next asynchronous writes, then hand the buffer back
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		__syncwarp();
		if (lane == 0) mbar_arrive(ps.empty + buf);
	}
}

template <bool INV>
static int launch_upper_pipe_t(const UpperParams& P, cudaStream_t stream) {
	static int sms_of[64] = {};
	int dev = 0;
	const size_t smem = kPipeSmemWords * sizeof(uint32_t);
	cudaGetDevice(&dev);
	if (dev < 0 || dev >= 64) return -1;
	if (!sms_of[dev]) {  // once per device
		int sms = 0;
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
		if (cudaFuncSetAttribute(upper_pipe_kernel<INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
		sms_of[dev] = sms;
	}
	uint64_t ctas = (uint64_t)sms_of[dev];
	if (ctas > P.n_groups) ctas = P.n_groups;
	if (ctas == 0) return -1;
	upper_pipe_kernel<INV><<<(unsigned)ctas, kPipeThreads, smem, stream>>>(P);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_upper_pipe(const UpperParams& P, cudaStream_t stream) {
	if (P.k != 4 && P.k != 6) return -1;
	if (P.ls != (P.k == 6 ? 3 : 5)) return -1;
	return P.inverse ? launch_upper_pipe_t<true>(P, stream) : launch_upper_pipe_t<false>(P, stream);
}

}  // namespace b2
