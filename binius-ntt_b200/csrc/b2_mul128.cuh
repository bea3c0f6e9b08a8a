// b2_mul128.cuh -- CTA-level bit-sliced GF(2^128) product of 32 staged blocks (multiply_unrolled<7> semantics,
// finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh:4-5), shared by the element-wise multiply
// (b2_misc.cu) and the sumcheck kernels (b2_sumcheck.cu).  sm_100a only.
//
// A block is 4 limbs x 32 words (word 32k + b = bit b of limb k of 32 elements).  Two Karatsuba levels above GF(2^32) give
// nine GF(2^32) products per block: warp j of the 9 warps computes product j for the 32 blocks (lane = block), so a
// thread holds one 32-word operand pair and its product.  Operands sit in padded rows (pitch 132 words: conflict-free
// 16-byte reads of 8 different rows), the products are combined through shared memory in two linear steps and the
// result replaces operand a.
//   j = 3u + v:  u = 0: A0 = (a0, a1), 1: A1 = (a2, a3), 2: A0 + A1;  v = 0: low limb, 1: high limb, 2: their sum
//   Z_u = (P_3u + P_3u+1,  P_3u+2 + P_3u + P_3u+1 + alpha5 * P_3u+1)           (GF(2^64) products)
//   out = (Z_0 + Z_1,  Z_2 + Z_0 + Z_1 + alpha6 * Z_1),  alpha6 * (lo, hi) = (hi, lo + alpha5 * hi)
#pragma once
#include <cstdint>

#include "b2_bundle.cuh"

namespace b2 {

constexpr int kMulPitch = 132, kProdPitch = 33, kMulThreads = 288;
constexpr int kMulSmemWords = 2 * 32 * kMulPitch + 15 * 32 * kProdPitch;
constexpr size_t kMulSmemBytes = kMulSmemWords * sizeof(uint32_t);

struct MulSmem {
	uint32_t* sa;  // [32][132] operand a, afterwards the product rows
	uint32_t* sb;  // [32][132] operand b (left intact)
	uint32_t* sp;  // [9][32][33] products P_j
	uint32_t* sz;  // [6][32][33] Z_u low / high halves
};

__device__ __forceinline__ MulSmem mul_smem_view(uint32_t* base) {
	MulSmem m;
	m.sa = base;
	m.sb = base + 32 * kMulPitch;
	m.sp = m.sb + 32 * kMulPitch;
	m.sz = m.sp + 9 * 32 * kProdPitch;
	return m;
}

__device__ __forceinline__ void load32(const uint32_t* p, uint32_t* w) {
	const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const uint4 u = q[i];
		w[4 * i] = u.x, w[4 * i + 1] = u.y, w[4 * i + 2] = u.z, w[4 * i + 3] = u.w;
	}
}

// rows [0, nact) of sa <- sa * sb.  Called by all 288 threads after a __syncthreads() that published the operands; ends
// with a __syncthreads(), after which the product rows are visible in sa.  Keep ONE call site per kernel (code size).
__device__ __forceinline__ void cta_mul128(uint32_t* sa, const uint32_t* sb, uint32_t* sp, uint32_t* sz, int nact) {
	const int j = threadIdx.x >> 5, lb = threadIdx.x & 31, u = j / 3, v = j - 3 * u;
	const bool act = lb < nact;
	if (act) {
		// operand limbs of product j: limbs {l0} (+ {l0 + 1} when v = 2), the same from the other half when u = 2
		uint32_t x[32], y[32], t[32];
		const uint32_t *pa = sa + lb * kMulPitch, *pb = sb + lb * kMulPitch;
		const int base = u == 1 ? 2 : 0, l0 = base + (v == 1 ? 1 : 0);
		load32(pa + 32 * l0, x), load32(pb + 32 * l0, y);
		if (v == 2) {
			load32(pa + 32 * (l0 + 1), t);
#pragma unroll
			for (int i = 0; i < 32; i++) x[i] ^= t[i];
			load32(pb + 32 * (l0 + 1), t);
#pragma unroll
			for (int i = 0; i < 32; i++) y[i] ^= t[i];
		}
		if (u == 2) {
			load32(pa + 32 * (l0 + 2), t);
#pragma unroll
			for (int i = 0; i < 32; i++) x[i] ^= t[i];
			load32(pb + 32 * (l0 + 2), t);
#pragma unroll
			for (int i = 0; i < 32; i++) y[i] ^= t[i];
			if (v == 2) {
				load32(pa + 32 * (l0 + 3), t);
#pragma unroll
				for (int i = 0; i < 32; i++) x[i] ^= t[i];
				load32(pb + 32 * (l0 + 3), t);
#pragma unroll
				for (int i = 0; i < 32; i++) y[i] ^= t[i];
			}
		}
		Sliced<5>::mul(x, y, t);
#pragma unroll
		for (int i = 0; i < 32; i++) sp[(j * 32 + lb) * kProdPitch + i] = t[i];
	}
	__syncthreads();
	if (act && j < 6) {  // Z_u halves: warp 2u -> low, warp 2u + 1 -> high
		const int uu = j >> 1;
		const uint32_t *p0 = sp + ((3 * uu) * 32 + lb) * kProdPitch, *p1 = p0 + 32 * kProdPitch, *p2 = p1 + 32 * kProdPitch;
		uint32_t* z = sz + (j * 32 + lb) * kProdPitch;
		if (!(j & 1)) {
			for (int i = 0; i < 32; i++) z[i] = p0[i] ^ p1[i];
		} else {
			uint32_t q[32], aq[32];
			for (int i = 0; i < 32; i++) q[i] = p1[i];
			Sliced<5>::mul_alpha(q, aq);
			for (int i = 0; i < 32; i++) z[i] = p2[i] ^ p0[i] ^ q[i] ^ aq[i];
		}
	}
	__syncthreads();
	if (act && j < 4) {  // result limb j of the block, into the row that held operand a
		const uint32_t *z0l = sz + lb * kProdPitch, *z0h = z0l + 32 * kProdPitch, *z1l = z0h + 32 * kProdPitch, *z1h = z1l + 32 * kProdPitch,
		               *z2l = z1h + 32 * kProdPitch, *z2h = z2l + 32 * kProdPitch;
		uint32_t* po = sa + lb * kMulPitch + 32 * j;
		if (j == 0) {
			for (int i = 0; i < 32; i++) po[i] = z0l[i] ^ z1l[i];
		} else if (j == 1) {
			for (int i = 0; i < 32; i++) po[i] = z0h[i] ^ z1h[i];
		} else if (j == 2) {
			for (int i = 0; i < 32; i++) po[i] = z2l[i] ^ z0l[i] ^ z1l[i] ^ z1h[i];
		} else {
			uint32_t q[32], aq[32];
			for (int i = 0; i < 32; i++) q[i] = z1h[i];
			Sliced<5>::mul_alpha(q, aq);
			for (int i = 0; i < 32; i++) po[i] = z2h[i] ^ z0h[i] ^ q[i] ^ z1l[i] ^ aq[i];
		}
	}
	__syncthreads();
}

}  // namespace b2
