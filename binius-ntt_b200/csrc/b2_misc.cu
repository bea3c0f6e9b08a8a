// b2_misc.cu -- layout conversion and element-wise tower-multiply kernels.  sm_100a only.
//
// Replaces transpose_kernel<W> / untranspose_kernel<W> (utils/bitslicing.cuh:89-104), the smoke
// kernels of finite_fields/tests/test_kernels.cu:9-60 and bitsliced_repeat
// (finite_fields/tests/profiling/kernels/bitsliced_repeat.cu:6-33).
#include <cuda_runtime.h>
#include <stdlib.h>

#include "b2_bundle.cuh"
#include "b2_launch.h"
#include "b2_pass.cuh"
#include "b2_mul128.cuh"

namespace b2 {

// ------------------------------------------------------------------------------------------------
// compact <-> little-endian bit-sliced, blocks of 32 elements of L = width/32 limbs.
// compact: word e*L + k = limb k of element e; bit-sliced: word 32*k + b = bit b of limb k of the
// 32 elements, element e at bit position e (utils/bitslicing.cuh:32-47).
// A warp converts 32 blocks at a time through shared memory so that global accesses are whole
// 128-byte lines; each thread transposes one (block, limb) 32x32 bit matrix in registers.
// ------------------------------------------------------------------------------------------------
template <int L>
__global__ void __launch_bounds__(128) bitslice_kernel(uint32_t* __restrict__ d, uint64_t n_blocks, int dir) {
	constexpr int W = 32 * L;               // words per block
	__shared__ uint32_t sm[4][32 * 33];     // per warp: 32 lines of 32 words, padded
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uint32_t* s = sm[warp];
	// a "unit" = one (block, limb) matrix; units of a block are contiguous in the sliced image only
	const uint64_t n_groups = (n_blocks * L + 31) / 32;  // groups of 32 units (= 32/L blocks)
	for (uint64_t grp = (uint64_t)blockIdx.x * 4 + warp; grp < n_groups; grp += (uint64_t)gridDim.x * 4) {
		// 32/L blocks x W words = 1024 words, contiguous in memory
		const uint64_t base = grp * 1024;
		const uint64_t total = n_blocks * W;
		uint32_t r[32];
		// coalesced load of 32 lines into shared memory (line i, word lane)
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint64_t g = base + (uint64_t)i * 32 + lane;
			s[i * 33 + lane] = g < total ? d[g] : 0u;
		}
		__syncwarp();
		// unit = lane: block lb = lane / L, limb k = lane % L
		const int lb = lane / L, k = lane % L;
		if (dir == 0) {
			// gather limb k of elements 0..31 of block lb: compact word (e*L + k) of the block
#pragma unroll
			for (int e = 0; e < 32; e++) {
				const int wi = lb * W + e * L + k;
				r[e] = s[(wi >> 5) * 33 + (wi & 31)];
			}
		} else {
#pragma unroll
			for (int b = 0; b < 32; b++) {
				const int wi = lb * W + 32 * k + b;
				r[b] = s[(wi >> 5) * 33 + (wi & 31)];
			}
		}
		transpose32(r);
		__syncwarp();
		if (dir == 0) {
#pragma unroll
			for (int b = 0; b < 32; b++) {
				const int wi = lb * W + 32 * k + b;
				s[(wi >> 5) * 33 + (wi & 31)] = r[b];
			}
		} else {
#pragma unroll
			for (int e = 0; e < 32; e++) {
				const int wi = lb * W + e * L + k;
				s[(wi >> 5) * 33 + (wi & 31)] = r[e];
			}
		}
		__syncwarp();
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint64_t g = base + (uint64_t)i * 32 + lane;
			if (g < total) d[g] = s[i * 33 + lane];
		}
		__syncwarp();
	}
}

int launch_bitslice(uint32_t* d, uint64_t n_blocks, int width_bits, int dir, cudaStream_t stream) {
	if (n_blocks == 0) return 0;
	const int L = width_bits / 32;
	const uint64_t n_groups = (n_blocks * L + 31) / 32;
	unsigned blocks = (unsigned)((n_groups + 3) / 4);
	if (blocks > 148 * 16) blocks = 148 * 16;
	switch (L) {
		case 1: bitslice_kernel<1><<<blocks, 128, 0, stream>>>(d, n_blocks, dir); break;
		case 2: bitslice_kernel<2><<<blocks, 128, 0, stream>>>(d, n_blocks, dir); break;
		case 4: bitslice_kernel<4><<<blocks, 128, 0, stream>>>(d, n_blocks, dir); break;
		default: return -1;
	}
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ------------------------------------------------------------------------------------------------
// element-wise products
// ------------------------------------------------------------------------------------------------
template <int H>
// the output may alias either input (include/b2ntt.h), so none of the multiply kernels declares __restrict__
__global__ void mul_compact_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n) {
	for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n; g += (uint64_t)gridDim.x * blockDim.x)
		o[g] = (uint32_t)Tower<H>::mul(a[g], b[g]);
}

// compact GF(2^32) (multiply<5>, finite_fields/binary_tower.cuh): a warp stages 32 lines per operand through shared
// memory (coalesced), each lane bit-slices its own line of 32 elements in registers, multiplies bit-sliced and
// transposes the products back -- ~1.5k logic instructions per 32 products instead of one tower recursion per element.
__global__ void __launch_bounds__(128) mul_compact32_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n) {
	__shared__ uint32_t sm[4][32 * 33];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uint32_t* s = sm[warp];
	const uint64_t n_groups = (n + 1023) / 1024;
	for (uint64_t grp = (uint64_t)blockIdx.x * 4 + warp; grp < n_groups; grp += (uint64_t)gridDim.x * 4) {
		const uint64_t base = grp * 1024;
		uint32_t x[32], y[32], z[32];
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint64_t g = base + (uint64_t)i * 32 + lane;
			s[i * 33 + lane] = g < n ? a[g] : 0u;
		}
		__syncwarp();
#pragma unroll
		for (int e = 0; e < 32; e++) x[e] = s[lane * 33 + e];
		__syncwarp();
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint64_t g = base + (uint64_t)i * 32 + lane;
			s[i * 33 + lane] = g < n ? b[g] : 0u;
		}
		__syncwarp();
#pragma unroll
		for (int e = 0; e < 32; e++) y[e] = s[lane * 33 + e];
		__syncwarp();
		transpose32(x), transpose32(y);
		Sliced<5>::mul(x, y, z);
		transpose32(z);
#pragma unroll
		for (int e = 0; e < 32; e++) s[lane * 33 + e] = z[e];
		__syncwarp();
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint64_t g = base + (uint64_t)i * 32 + lane;
			if (g < n) o[g] = s[i * 33 + lane];
		}
		__syncwarp();
	}
}

// bit-sliced GF(2^32): one thread per block of 32 products, everything in registers
__global__ void __launch_bounds__(128) mul_sliced32_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n_blocks) {
	for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < n_blocks; g += (uint64_t)gridDim.x * blockDim.x) {
		uint32_t x[32], y[32], z[32];
		const uint4* pa = reinterpret_cast<const uint4*>(a + g * 32);
		const uint4* pb = reinterpret_cast<const uint4*>(b + g * 32);
#pragma unroll
		for (int i = 0; i < 8; i++) {
			const uint4 u = pa[i], v = pb[i];
			x[4 * i] = u.x, x[4 * i + 1] = u.y, x[4 * i + 2] = u.z, x[4 * i + 3] = u.w;
			y[4 * i] = v.x, y[4 * i + 1] = v.y, y[4 * i + 2] = v.z, y[4 * i + 3] = v.w;
		}
		Sliced<5>::mul(x, y, z);
		uint4* po = reinterpret_cast<uint4*>(o + g * 32);
#pragma unroll
		for (int i = 0; i < 8; i++) po[i] = make_uint4(z[4 * i], z[4 * i + 1], z[4 * i + 2], z[4 * i + 3]);
	}
}

// bit-sliced GF(2^128), 32 blocks (16 KB per operand) per CTA iteration: operands arrive by coalesced cp.async into the
// padded rows of b2_mul128.cuh, cta_mul128 multiplies, the result rows leave by coalesced 16-byte stores.
#ifndef B2_MUL128_CTAS
#define B2_MUL128_CTAS 2  // 96 registers per thread (a few spilled words), two CTAs = 18 warps per SM: 9.1 ms vs 11.7 ms at one CTA
#endif
// COMPACT = true: the operands are compact GF(2^128) elements (multiply<7>, finite_fields/binary_tower.cuh): a block row is
// 32 elements x 4 limbs; warps 0..7 bit-slice the staged rows in place (lane = 4 * block + limb: conflict-free gathers),
// the product runs bit-sliced as above and warps 0..3 transpose the result rows back before the store.  n_chunks = number
// of valid 16-byte chunks (= elements when compact), so a ragged last block is neither read nor written past its end.
template <bool COMPACT>
__global__ void __launch_bounds__(kMulThreads, B2_MUL128_CTAS) mul_sliced128_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n_blocks, uint64_t n_chunks) {
	extern __shared__ __align__(16) uint32_t msm[];
	const MulSmem m = mul_smem_view(msm);
	uint32_t *sa = m.sa, *sb = m.sb;
	const int j = threadIdx.x >> 5, lb = threadIdx.x & 31;
	for (uint64_t g0 = (uint64_t)blockIdx.x * 32; g0 < n_blocks; g0 += (uint64_t)gridDim.x * 32) {
		const int nact = n_blocks - g0 < 32 ? (int)(n_blocks - g0) : 32;
		for (int ch = threadIdx.x; ch < nact * 32; ch += kMulThreads) {  // 16-byte chunks: 32 per block row
			const int row = ch >> 5, c4 = (ch & 31) * 4;
			if ((g0 + row) * 32 + (ch & 31) >= n_chunks) continue;
			const unsigned da = (unsigned)__cvta_generic_to_shared(sa + row * kMulPitch + c4), db = (unsigned)__cvta_generic_to_shared(sb + row * kMulPitch + c4);
			asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(a + (g0 + row) * 128 + c4) : "memory");
			asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(db), "l"(b + (g0 + row) * 128 + c4) : "memory");
		}
		asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
		__syncthreads();
		if (COMPACT) {
			if (j < 8) {
				uint32_t r[32];
				uint32_t* row = (j < 4 ? sa : sb) + (8 * (j & 3) + (lb >> 2)) * kMulPitch;
				const int k = lb & 3;
#pragma unroll
				for (int e = 0; e < 32; e++) r[e] = row[4 * e + k];
				transpose32(r);
				__syncwarp();
#pragma unroll
				for (int i = 0; i < 32; i++) row[32 * k + i] = r[i];
			}
			__syncthreads();
		}
		cta_mul128(sa, sb, m.sp, m.sz, nact);
		if (COMPACT) {
			if (j < 4) {
				uint32_t r[32];
				uint32_t* row = sa + (8 * j + (lb >> 2)) * kMulPitch;
				const int k = lb & 3;
#pragma unroll
				for (int i = 0; i < 32; i++) r[i] = row[32 * k + i];
				transpose32(r);
				__syncwarp();
#pragma unroll
				for (int e = 0; e < 32; e++) row[4 * e + k] = r[e];
			}
			__syncthreads();
		}
		for (int ch = threadIdx.x; ch < nact * 32; ch += kMulThreads) {
			const int row = ch >> 5, c4 = (ch & 31) * 4;
			if ((g0 + row) * 32 + (ch & 31) >= n_chunks) continue;
			*reinterpret_cast<uint4*>(o + (g0 + row) * 128 + c4) = *reinterpret_cast<const uint4*>(sa + row * kMulPitch + c4);
		}
		__syncthreads();
	}
}

// ---- bit-sliced GF(2^128), warp-owned form (config 5) --------------------------------------------------------------------------
// A warp owns 32 consecutive blocks, lane = block, and runs all nine GF(2^32) products of its blocks itself: no CTA barrier, no
// operand staging.  The operand of product j is a subset sum of the block's four limbs; the nine subsets are visited along a
// Gray path ({0} {01} {1} {13} {3} {23} {2} {02} {0123}), so x and y change by ONE limb (read straight from global memory /
// L1, 8 x 16 bytes per thread) between products instead of being rebuilt.  Each product is XORed into the lane's private
// result row in shared memory (4 limbs x 32 words, pitch 132: conflict-free 16-byte accesses), which the warp then stores
// coalesced.  With a5 = the GF(2^32)-level alpha (A below) the result limbs are
//   L0 = P0+P1+P3+P4                      L1 = P0+P1+P2+P3+P4+P5 + A(P1+P4)
//   L2 = P0+P1+P5+P6+P7 + A P4            L3 = P0+P1+P2+P5+P6+P7+P8 + A(P1+P3+P5+P7) + A A P4
// (P0 = a0 b0, P1 = a1 b1, P2 = (a0+a1)(b0+b1), P3..P5 the same on limbs 2, 3, P6..P8 on the sums a0+a2, a1+a3: two Karatsuba
// levels over GF(2^32), multiply_unrolled<7>, finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh:4-5).
constexpr int kWMulWarps = 8, kWMulPitch = 132, kWStagePitch = 36;
constexpr int kWMulWarpWords = 32 * kWMulPitch + 2 * 32 * kWStagePitch;  // result rows + one staged limb of a and of b
constexpr size_t kWMulSmemBytes = (size_t)kWMulWarps * kWMulWarpWords * sizeof(uint32_t);

__device__ __forceinline__ void xor_limb(const uint32_t* g, uint32_t* w) {  // w ^= the 32 words at g
	const uint4* q = reinterpret_cast<const uint4*>(g);
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const uint4 u = q[i];
		w[4 * i] ^= u.x, w[4 * i + 1] ^= u.y, w[4 * i + 2] ^= u.z, w[4 * i + 3] ^= u.w;
	}
}
template <bool FIRST>
__device__ __forceinline__ void acc_limb(uint32_t* row, const uint32_t* v) {  // row (^)= v
	uint4* q = reinterpret_cast<uint4*>(row);
#pragma unroll
	for (int i = 0; i < 8; i++) {
		uint4 u = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
		if (!FIRST) {
			const uint4 o = q[i];
			u.x ^= o.x, u.y ^= o.y, u.z ^= o.z, u.w ^= o.w;
		}
		q[i] = u;
	}
}
// limb `limb` of the 32 blocks starting at block `first`, of both operands, into the warp's staging rows: coalesced 16-byte
// cp.async, four 128-byte rows per warp instruction; rows past nact are not touched
__device__ __forceinline__ void stage_limb(uint32_t* sta, uint32_t* stb, const uint32_t* a, const uint32_t* b, uint64_t first, int nact, int limb, int lane) {
	const int piece = lane & 7, r4 = lane >> 3;
#pragma unroll
	for (int it = 0; it < 8; it++) {
		const int r = it * 4 + r4;
		if (r < nact) {
			const uint64_t off = (first + r) * 128 + 32 * limb + piece * 4;
			copy16_g2s(sta + r * kWStagePitch + piece * 4, a + off);
			copy16_g2s(stb + r * kWStagePitch + piece * 4, b + off);
		}
	}
	asm volatile("cp.async.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(kWMulWarps * 32, 1) mul_sliced128_warp_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n_blocks) {
	extern __shared__ __align__(16) uint32_t wsm[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	uint32_t* rows = wsm + (size_t)warp * kWMulWarpWords;
	uint32_t *sta = rows + 32 * kWMulPitch, *stb = sta + 32 * kWStagePitch;
	uint32_t* row = rows + lane * kWMulPitch;  // this lane's result block
	const uint64_t n_sets = (n_blocks + 31) / 32, stride = (uint64_t)gridDim.x * kWMulWarps;
	uint64_t set = (uint64_t)blockIdx.x * kWMulWarps + warp;
	if (set < n_sets) stage_limb(sta, stb, a, b, set * 32, n_blocks - set * 32 < 32 ? (int)(n_blocks - set * 32) : 32, 0, lane);
	for (; set < n_sets; set += stride) {
		const uint64_t first = set * 32;
		const int nact = n_blocks - first < 32 ? (int)(n_blocks - first) : 32;
		const bool act = lane < nact;
		uint32_t x[32], y[32], t[32], q[32];
#pragma unroll
		for (int i = 0; i < 32; i++) x[i] = 0, y[i] = 0;
		// The limbs arrive in the order 0 1 0 3 1 2 3 0 1 3 (one toggle of the Gray path each, the last subset needs two); a
		// product follows every arrival but the ninth.  ONE product call site walked nine times: nine inlined copies of the
		// circuit spill and thrash the instruction cache.  The next limb (or limb 0 of the warp's next 32 blocks) is copied in
		// while the product runs.
		// step:            P0      P2      P1      P7      P4      P5      P3      P6      P8
		// t        -> limbs 0123   13      02      2       0       12      01      23      3
		// t + A t  -> limbs -      -       13      3       1       3       -       -       -
		// A t      -> limbs -      -       -       -       2       -       3       -       -
		// A A t    -> limbs -      -       -       -       3       -       -       -       -
		const uint64_t order = 0x3103213010ull;  // nibble i = limb of arrival i
		const uint64_t tm = 0x8C36145AFull, rm = 0x828A00ull, qm = 0x8040000ull, am = 0x80000ull;  // nibble per step, bit l = limb l
#pragma unroll 1
		for (int i = 0; i < 10; i++) {
			asm volatile("cp.async.wait_group 0;" ::: "memory");
			__syncwarp();
			if (act) xor_limb(sta + lane * kWStagePitch, x), xor_limb(stb + lane * kWStagePitch, y);
			__syncwarp();  // the staging rows are free again
			if (i < 9)
				stage_limb(sta, stb, a, b, first, nact, (int)((order >> (4 * (i + 1))) & 15), lane);
			else if (set + stride < n_sets)
				stage_limb(sta, stb, a, b, (set + stride) * 32, n_blocks - (set + stride) * 32 < 32 ? (int)(n_blocks - (set + stride) * 32) : 32, 0, lane);
			if (i == 8 || !act) continue;
			const int st = i < 8 ? i : 8;
			Sliced<5>::mul(x, y, t);
			const int mt = (int)((tm >> (4 * st)) & 15), mr = (int)((rm >> (4 * st)) & 15), mq = (int)((qm >> (4 * st)) & 15), ma = (int)((am >> (4 * st)) & 15);
#pragma unroll 1
			for (int l = 0; l < 4; l++)
				if ((mt >> l) & 1) {
					if (st == 0)
						acc_limb<true>(row + 32 * l, t);
					else
						acc_limb<false>(row + 32 * l, t);
				}
			if (mr | mq | ma) {
				Sliced<5>::mul_alpha(t, q);
#pragma unroll 1
				for (int l = 0; l < 4; l++)
					if ((mq >> l) & 1) acc_limb<false>(row + 32 * l, q);
				if (mr) {
#pragma unroll
					for (int k = 0; k < 32; k++) t[k] ^= q[k];
#pragma unroll 1
					for (int l = 0; l < 4; l++)
						if ((mr >> l) & 1) acc_limb<false>(row + 32 * l, t);
				}
				if (ma) {
					Sliced<5>::mul_alpha(q, t);
#pragma unroll 1
					for (int l = 0; l < 4; l++)
						if ((ma >> l) & 1) acc_limb<false>(row + 32 * l, t);
				}
			}
		}
		__syncwarp();  // every lane has finished its row; all operand reads of this set are done (o may alias a or b)
#pragma unroll 4
		for (int r = 0; r < nact; r++)  // one block (512 bytes) per warp instruction
			*reinterpret_cast<uint4*>(o + (first + r) * 128 + lane * 4) = *reinterpret_cast<const uint4*>(rows + r * kWMulPitch + lane * 4);
		__syncwarp();
	}
}

int launch_mul(int field_bits, int layout, const uint32_t* a, const uint32_t* b, uint32_t* o, uint64_t n, cudaStream_t stream) {
	if (n == 0) return 0;
	const unsigned cap = 148 * 16;
	if (layout == 0) {
		unsigned blocks = (unsigned)((n + 255) / 256);
		if (blocks > cap) blocks = cap;
		switch (field_bits) {
			case 8: mul_compact_kernel<3><<<blocks, 256, 0, stream>>>(a, b, o, n); break;
			case 16: mul_compact_kernel<4><<<blocks, 256, 0, stream>>>(a, b, o, n); break;
			case 32:
				blocks = (unsigned)((n + 4095) / 4096);
				if (blocks > cap) blocks = cap;
				mul_compact32_kernel<<<blocks, 128, 0, stream>>>(a, b, o, n);
				break;
			case 128: {
				const uint64_t nb = (n + 31) / 32;
				blocks = (unsigned)((nb + 31) / 32);
				if (blocks > cap) blocks = cap;
				cudaFuncSetAttribute(mul_sliced128_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMulSmemBytes);
				mul_sliced128_kernel<true><<<blocks, kMulThreads, kMulSmemBytes, stream>>>(a, b, o, nb, n);
				break;
			}
			default: return -1;
		}
	} else {
		if (n % 32) return -1;
		const uint64_t nb = n / 32;
		if (field_bits == 32) {
			unsigned blocks = (unsigned)((nb + 127) / 128);
			if (blocks > cap) blocks = cap;
			mul_sliced32_kernel<<<blocks, 128, 0, stream>>>(a, b, o, nb);
		} else if (field_bits == 128) {
			static const int form = getenv("B2NTT_MUL128") ? atoi(getenv("B2NTT_MUL128")) : 1;  // 0: CTA-level product (A/B), 1: warp-owned
			if (form == 1) {
				const uint64_t sets = (nb + 31) / 32;
				unsigned blocks = (unsigned)((sets + kWMulWarps - 1) / kWMulWarps);
				if (blocks > 148 * 4) blocks = 148 * 4;
				cudaFuncSetAttribute(mul_sliced128_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWMulSmemBytes);
				mul_sliced128_warp_kernel<<<blocks, kWMulWarps * 32, kWMulSmemBytes, stream>>>(a, b, o, nb);
			} else {
				unsigned blocks = (unsigned)((nb + 31) / 32);
				if (blocks > cap) blocks = cap;
				cudaFuncSetAttribute(mul_sliced128_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMulSmemBytes);
				mul_sliced128_kernel<false><<<blocks, kMulThreads, kMulSmemBytes, stream>>>(a, b, o, nb, nb * 32);
			}
		} else
			return -1;
	}
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// dependent chain x <- x * y in registers (logic-pipe peak probe)
__global__ void __launch_bounds__(128) mul_repeat32_kernel(uint32_t* io, uint64_t n_threads, int iters) {
	const uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (g >= n_threads) return;
	uint32_t x[32], y[32];
	load32(io + g * 32, x);
#pragma unroll
	for (int i = 0; i < 32; i++) y[i] = x[(i + 1) & 31] ^ 0x9e3779b9u * (i + 1);
#pragma unroll 1
	for (int it = 0; it < iters; it++) Sliced<5>::mul(x, y, x);
	uint4* po = reinterpret_cast<uint4*>(io + g * 32);
#pragma unroll
	for (int i = 0; i < 8; i++) po[i] = make_uint4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]);
}

// GF(2^128) form of the probe: x <- x * y on lane-private rows of shared memory (a 128-word operand does not fit a thread's
// registers), the nine GF(2^32) products of every step through the same single call site as mul_sliced128_warp_kernel
__device__ __forceinline__ void lane_mul128(const uint32_t* xa, const uint32_t* yb, uint32_t* row) {
	uint32_t x[32], y[32], t[32], q[32];
#pragma unroll
	for (int i = 0; i < 32; i++) x[i] = 0, y[i] = 0;
	const uint64_t order = 0x3103213010ull, tm = 0x8C36145AFull, rm = 0x828A00ull, qm = 0x8040000ull, am = 0x80000ull;
#pragma unroll 1
	for (int i = 0; i < 10; i++) {
		const int limb = (int)((order >> (4 * i)) & 15);
		xor_limb(xa + 32 * limb, x), xor_limb(yb + 32 * limb, y);
		if (i == 8) continue;
		const int st = i < 8 ? i : 8;
		Sliced<5>::mul(x, y, t);
		const int mt = (int)((tm >> (4 * st)) & 15), mr = (int)((rm >> (4 * st)) & 15), mq = (int)((qm >> (4 * st)) & 15), ma = (int)((am >> (4 * st)) & 15);
#pragma unroll 1
		for (int l = 0; l < 4; l++)
			if ((mt >> l) & 1) {
				if (st == 0)
					acc_limb<true>(row + 32 * l, t);
				else
					acc_limb<false>(row + 32 * l, t);
			}
		if (mr | mq | ma) {
			Sliced<5>::mul_alpha(t, q);
#pragma unroll 1
			for (int l = 0; l < 4; l++)
				if ((mq >> l) & 1) acc_limb<false>(row + 32 * l, q);
			if (mr) {
#pragma unroll
				for (int k = 0; k < 32; k++) t[k] ^= q[k];
#pragma unroll 1
				for (int l = 0; l < 4; l++)
					if ((mr >> l) & 1) acc_limb<false>(row + 32 * l, t);
			}
			if (ma) {
				Sliced<5>::mul_alpha(q, t);
#pragma unroll 1
				for (int l = 0; l < 4; l++)
					if ((ma >> l) & 1) acc_limb<false>(row + 32 * l, t);
			}
		}
	}
}
constexpr int kRep128Threads = 128;
constexpr size_t kRep128SmemBytes = (size_t)3 * kRep128Threads * kWMulPitch * sizeof(uint32_t);
// thread g owns the 128-word block g of io: x = the block, y = x with its limbs rotated and a constant mixed in
__global__ void __launch_bounds__(kRep128Threads, 1) mul_repeat128_kernel(uint32_t* io, uint64_t n_threads, int iters) {
	extern __shared__ __align__(16) uint32_t rsm[];
	const uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (g >= n_threads) return;  // rows are thread-private: no barrier below
	uint32_t* xr = rsm + (size_t)threadIdx.x * kWMulPitch;
	uint32_t* yr = xr + (size_t)kRep128Threads * kWMulPitch;
	uint32_t* zr = yr + (size_t)kRep128Threads * kWMulPitch;
	for (int i = 0; i < 128; i++) xr[i] = io[g * 128 + i];
	for (int i = 0; i < 128; i++) yr[i] = xr[(i + 33) & 127] ^ 0x9e3779b9u * (i + 1);
#pragma unroll 1
	for (int it = 0; it < iters; it++) {
		lane_mul128(xr, yr, zr);
		uint32_t* sw = xr;
		xr = zr, zr = sw;
	}
	for (int i = 0; i < 128; i++) io[g * 128 + i] = xr[i];
}

int launch_mul_repeat(int field_bits, uint32_t* io, uint64_t n_threads, int iters, cudaStream_t stream) {
	if (field_bits != 32 && field_bits != 128) return -4;
	if (n_threads == 0) return 0;
	if (field_bits == 128) {
		cudaFuncSetAttribute(mul_repeat128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRep128SmemBytes);
		mul_repeat128_kernel<<<(unsigned)((n_threads + kRep128Threads - 1) / kRep128Threads), kRep128Threads, kRep128SmemBytes, stream>>>(io, n_threads, iters);
	} else
		mul_repeat32_kernel<<<(unsigned)((n_threads + 127) / 128), 128, 0, stream>>>(io, n_threads, iters);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace b2
