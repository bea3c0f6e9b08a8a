// b2_bundle.cuh -- register-resident "bundles": one 128-byte line of the data set (32 consecutive
// 32-bit words = 32 GF(2^32) elements, or 8 GF(2^128) elements x 4 limbs) held bit-sliced in 32
// registers, word i = bit i of the 32 words of the line.
//
// Why bundles: every limb of every element of a line meets the SAME twiddle in every butterfly
// layer whose stride is at least one line (twiddles of the reference's domain lie in GF(2^32) and
// act limb-wise on GF(2^128)), so the butterfly u ^= w*v ; v ^= u
// (ntt/additive_ntt.cuh:10-14) becomes one bit-sliced product of 32 lanes by a lane-uniform w.
// The layers below one line (element stride < 32 words) pair lanes inside the words and use
// masked shifts with per-lane twiddle words.
#pragma once
#include "b2_field.cuh"

namespace b2 {

// LSB-first 32x32 bit transpose in registers: out word i, bit y = in word y, bit i.
// (the bit-level content of utils/bitslicing.cuh:14-26; involution)
B2_HD void transpose32(uint32_t* a) {
#if defined(__CUDA_ARCH__)
	// the 16-bit and 8-bit exchange steps are byte permutes (PRMT: 2 instructions per register pair instead of 5)
#pragma unroll
	for (int r = 0; r < 16; r++) {
		const uint32_t x = a[r], y = a[r + 16];
		a[r] = __byte_perm(x, y, 0x5410);
		a[r + 16] = __byte_perm(x, y, 0x7632);
	}
#pragma unroll
	for (int r = 0; r < 32; r++) {
		if (r & 8) continue;
		const uint32_t x = a[r], y = a[r + 8];
		a[r] = __byte_perm(x, y, 0x6240);
		a[r + 8] = __byte_perm(x, y, 0x7351);
	}
	constexpr int kFirstMaskedStep = 2;
#else
	constexpr int kFirstMaskedStep = 4;
#endif
#pragma unroll
	for (int s = kFirstMaskedStep; s >= 0; s--) {
		const int d = 1 << s;
		const uint32_t m = s == 4 ? 0x0000ffffu : s == 3 ? 0x00ff00ffu : s == 2 ? 0x0f0f0f0fu : s == 1 ? 0x33333333u : 0x55555555u;
#pragma unroll
		for (int r = 0; r < 32; r++) {
			if (r & d) continue;
			uint32_t x = ((a[r] >> d) ^ a[r + d]) & m;
			a[r] ^= x << d;
			a[r + d] ^= x;
		}
	}
}

// all-ones / all-zeros word from bit i of w
B2_HD uint32_t bitmask(uint32_t w, int i) { return (uint32_t)0 - ((w >> i) & 1u); }

// Butterfly between two bundles with a lane-uniform twiddle w in GF(2^(2^HW)):
//   U ^= w * V ; V ^= U.   HW == 0: w == 0 (top layer of a log_rate = 0 transform).
template <int HW>
B2_HD void butterfly_uniform(uint32_t* U, uint32_t* V, uint32_t w) {
	if constexpr (HW > 0) {
		constexpr int C = 1 << HW;
		uint32_t mw[C];
#pragma unroll
		for (int i = 0; i < C; i++) mw[i] = bitmask(w, i);
#pragma unroll
		for (int c = 0; c < 32 / C; c++) Sliced<HW>::mul_acc(mw, V + c * C, U + c * C);
	}
#pragma unroll
	for (int i = 0; i < 32; i++) V[i] ^= U[i];
}

// Inverse butterfly: V ^= U ; U ^= w * V.
template <int HW>
B2_HD void butterfly_uniform_inv(uint32_t* U, uint32_t* V, uint32_t w) {
#pragma unroll
	for (int i = 0; i < 32; i++) V[i] ^= U[i];
	if constexpr (HW > 0) {
		constexpr int C = 1 << HW;
		uint32_t mw[C];
#pragma unroll
		for (int i = 0; i < C; i++) mw[i] = bitmask(w, i);
#pragma unroll
		for (int c = 0; c < 32 / C; c++) Sliced<HW>::mul_acc(mw, V + c * C, U + c * C);
	}
}

// lanes whose bit p is clear
B2_HD uint32_t lane_mask_lo(int p) {
	return p == 0 ? 0x55555555u : p == 1 ? 0x33333333u : p == 2 ? 0x0f0f0f0fu : p == 3 ? 0x00ff00ffu : 0x0000ffffu;
}

// Butterfly layer INSIDE the lines of two bundles A and B at lane bit p (partner lanes x, x + 2^p).
// To keep all 32 lanes of the product busy, the v-lanes of A are shifted down onto A's u-lane
// positions while the v-lanes of B stay where they are; one product serves both bundles.
//   lane_tw : 2^HW words, per-lane part of the twiddle, the same value at lane x and x + 2^p
//   wa, wb  : line-uniform parts of the twiddle for A and for B (XORed onto lane_tw)
template <int HW, bool INVERSE>
B2_HD void butterfly_in_line(uint32_t* A, uint32_t* B, int p, const uint32_t* lane_tw, uint32_t wa, uint32_t wb) {
	const int sh = 1 << p;
	const uint32_t mu = lane_mask_lo(p);
	if (INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			A[i] ^= (A[i] & mu) << sh;
			B[i] ^= (B[i] & mu) << sh;
		}
	}
	if constexpr (HW > 0) {
		constexpr int C = 1 << HW;
		uint32_t W[C];
#pragma unroll
		for (int i = 0; i < C; i++) W[i] = lane_tw[i] ^ (bitmask(wa, i) & mu) ^ (bitmask(wb, i) & ~mu);
#pragma unroll
		for (int c = 0; c < 32 / C; c++) {
			uint32_t X[C], Pr[C];
#pragma unroll
			for (int i = 0; i < C; i++) X[i] = ((A[c * C + i] >> sh) & mu) | (B[c * C + i] & ~mu);
			Sliced<HW>::mul(W, X, Pr);
#pragma unroll
			for (int i = 0; i < C; i++) {
				A[c * C + i] ^= Pr[i] & mu;
				B[c * C + i] ^= (Pr[i] >> sh) & mu;
			}
		}
	}
	if (!INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			A[i] ^= (A[i] & mu) << sh;
			B[i] ^= (B[i] & mu) << sh;
		}
	}
}

// XOR of table[b] over the set bits b of `bits` (twiddle of additive_ntt.cuh:59-77 for one layer)
B2_HD uint32_t xor_select(const uint32_t* table, uint32_t bits) {
	uint32_t w = 0;
#pragma unroll 1
	for (int b = 0; bits; b++, bits >>= 1)
		if (bits & 1u) w ^= table[b];
	return w;
}

}  // namespace b2
