/*
 * bitslice.c -- little-endian bit-slicing and the bit-sliced tower product, CPU oracle.
 * TEST INFRASTRUCTURE ONLY.
 *
 * Layout (utils/bitslicing.cuh:32-47, README.md:41-45): a block is 32 consecutive W-bit elements
 * (W = 32*k words).  After bitslice, word 32*limb + b holds bit b of limb `limb` of the 32
 * elements, element e in bit position e.
 */
#include "oracle.h"

#include <string.h>

/* LSB-first 32x32 bit transpose: out word b, bit e  =  in word e, bit b.
 * Same data movement as utils/bitslicing.cuh:14-26, written as five masked swap rounds. */
void orc_transpose32(uint32_t a[32]) {
	static const uint32_t low_half[5] = {0x55555555u, 0x33333333u, 0x0f0f0f0fu, 0x00ff00ffu, 0x0000ffffu};
	for (int s = 4; s >= 0; s--) {
		const int d = 1 << s;
		for (int r = 0; r < 32; r++) {
			if (r & d) continue;
			/* swap the high d-bit blocks of row r with the low d-bit blocks of row r+d */
			uint32_t x = ((a[r] >> d) ^ a[r + d]) & low_half[s];
			a[r] ^= x << d;
			a[r + d] ^= x;
		}
	}
}

void orc_bitslice(uint32_t* blk, int width_bits) { /* bitslicing.cuh:32-47 */
	int k = width_bits / 32;
	uint32_t tmp[128];
	if (k < 1 || k > 4) return;
	memcpy(tmp, blk, sizeof(uint32_t) * width_bits);
	for (int i = 0; i < width_bits; i++) blk[32 * (i % k) + i / k] = tmp[i]; /* limb gather */
	for (int c = 0; c < k; c++) orc_transpose32(blk + 32 * c);
}

void orc_unbitslice(uint32_t* blk, int width_bits) { /* bitslicing.cuh:49-64 */
	int k = width_bits / 32;
	uint32_t tmp[128];
	if (k < 1 || k > 4) return;
	memcpy(tmp, blk, sizeof(uint32_t) * width_bits);
	for (int c = 0; c < k; c++) orc_transpose32(tmp + 32 * c);
	for (int i = 0; i < width_bits; i++) blk[k * (i % 32) + i / 32] = tmp[i];
}

/* multiply_alpha on a bit-sliced element of 2^h words (multiply_and_generate_circuit.cpp:62-84):
 * (a0, a1) -> (a1, a0 + alpha*a1), recursively; h = 0 is the identity. */
static void bs_mul_alpha(const uint32_t* a, uint32_t* out, int h) {
	if (h == 0) {
		out[0] = a[0];
		return;
	}
	int half = 1 << (h - 1);
	uint32_t t[64];
	bs_mul_alpha(a + half, t, h - 1);
	for (int i = 0; i < half; i++) {
		uint32_t lo = a[half + i];
		uint32_t hi = a[i] ^ t[i];
		out[i] = lo;
		out[half + i] = hi;
	}
}

/* Karatsuba over the tower on word vectors (multiply_and_generate_circuit.cpp:86-155); the
 * generated multiply_unrolled<h> is this recursion flattened.  out may alias a or b. */
void orc_mul_bitsliced(const uint32_t* a, const uint32_t* b, uint32_t* out, int h) {
	if (h == 0) {
		out[0] = a[0] & b[0];
		return;
	}
	int half = 1 << (h - 1);
	uint32_t sa[64], sb[64], z0[64], z2[64], z1[64], z2a[64];
	for (int i = 0; i < half; i++) {
		sa[i] = a[i] ^ a[half + i];
		sb[i] = b[i] ^ b[half + i];
	}
	orc_mul_bitsliced(a, b, z0, h - 1);
	orc_mul_bitsliced(a + half, b + half, z2, h - 1);
	orc_mul_bitsliced(sa, sb, z1, h - 1);
	bs_mul_alpha(z2, z2a, h - 1);
	for (int i = 0; i < half; i++) {
		uint32_t lo = z0[i] ^ z2[i];
		uint32_t hi = z1[i] ^ z0[i] ^ z2[i] ^ z2a[i];
		out[i] = lo;
		out[half + i] = hi;
	}
}
