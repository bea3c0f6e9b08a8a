/*
 * tower.c -- compact Fan-Paar (Wiedemann) binary tower field, CPU oracle.  TEST INFRASTRUCTURE ONLY.
 *
 * GF(2^(2^h)) = GF(2^(2^(h-1)))[X_{h-1}] / (X_{h-1}^2 + alpha_{h-1} X_{h-1} + 1), alpha_{h-1} = X_{h-2},
 * element = lo | hi << 2^(h-1).   Follows finite_fields/binary_tower.cuh:
 *   multiply        generic_multiply        :35-50   (Karatsuba, z2*alpha folded into the high half)
 *   square          generic_square          :52-61
 *   inverse         generic_inverse         :63-81   (norm to the subfield)
 *   multiply_alpha  generic_multiply_alpha  :83-93
 *   base case h=0   FanPaarTowerField<0>    :252-262
 * The reference tabulates h=2 and parts of h=3 (:130-250); those tables are the values of this
 * same recursion, which tests/test_oracle_field.py checks against oracle/_ref and the KATs.
 * h = 6, 7 extend the recursion unchanged, as the reference's own test-side code does
 * (sumcheck/test/utils/unbitsliced_mul.cuh:35-93, sumcheck/test/utils/tower_7_mul.cu:4-20).
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

static inline orc_u128 low_mask(int bits) {
	return bits >= 128 ? ~(orc_u128)0 : (((orc_u128)1 << bits) - 1);
}

orc_u128 orc_tower_mul_alpha(orc_u128 a, int h) {
	if (h == 0) return a & 1; /* binary_tower.cuh:261 */
	int half = 1 << (h - 1);
	orc_u128 m = low_mask(half);
	orc_u128 a0 = a & m, a1 = (a >> half) & m;
	orc_u128 z1 = orc_tower_mul_alpha(a1, h - 1);
	return a1 | ((a0 ^ z1) << half); /* :88-92 */
}

orc_u128 orc_tower_mul(orc_u128 a, orc_u128 b, int h) {
	if (h == 0) return a & b & 1; /* :255 */
	int half = 1 << (h - 1);
	orc_u128 m = low_mask(half);
	orc_u128 a0 = a & m, a1 = (a >> half) & m;
	orc_u128 b0 = b & m, b1 = (b >> half) & m;
	orc_u128 z0 = orc_tower_mul(a0, b0, h - 1);
	orc_u128 z2 = orc_tower_mul(a1, b1, h - 1);
	orc_u128 z1 = orc_tower_mul(a0 ^ a1, b0 ^ b1, h - 1) ^ z0 ^ z2;
	orc_u128 z2a = orc_tower_mul_alpha(z2, h - 1);
	return (z0 ^ z2) | ((z1 ^ z2a) << half); /* :44-49 */
}

orc_u128 orc_tower_square(orc_u128 a, int h) {
	if (h == 0) return a & 1; /* :257 */
	int half = 1 << (h - 1);
	orc_u128 m = low_mask(half);
	orc_u128 a0 = a & m, a1 = (a >> half) & m;
	orc_u128 z0 = orc_tower_square(a0, h - 1);
	orc_u128 z2 = orc_tower_square(a1, h - 1);
	orc_u128 z2a = orc_tower_mul_alpha(z2, h - 1);
	return (z0 ^ z2) | (z2a << half); /* :57-60 */
}

orc_u128 orc_tower_inverse(orc_u128 a, int h) {
	if (h == 0) return a & 1; /* :259 */
	if (a == 0) return 0;     /* the reference asserts; the oracle returns 0 */
	int half = 1 << (h - 1);
	orc_u128 m = low_mask(half);
	if ((a >> half) == 0) return orc_tower_inverse(a, h - 1); /* :69 */
	orc_u128 a0 = a & m, a1 = (a >> half) & m;
	orc_u128 inter = a0 ^ orc_tower_mul_alpha(a1, h - 1);
	orc_u128 delta = orc_tower_mul(a0, inter, h - 1) ^ orc_tower_square(a1, h - 1);
	orc_u128 dinv = orc_tower_inverse(delta, h - 1);
	orc_u128 inv0 = orc_tower_mul(dinv, inter, h - 1);
	orc_u128 inv1 = orc_tower_mul(dinv, a1, h - 1);
	return inv0 | (inv1 << half); /* :70-80 */
}

uint32_t orc_mul32(uint32_t a, uint32_t b) { return (uint32_t)orc_tower_mul(a, b, 5); }
uint32_t orc_square32(uint32_t a) { return (uint32_t)orc_tower_square(a, 5); }
uint32_t orc_inverse32(uint32_t a) { return (uint32_t)orc_tower_inverse(a, 5); }

void orc_mul128_words(const uint32_t a[4], const uint32_t b[4], uint32_t out[4]) {
	orc_u128 x = 0, y = 0;
	for (int i = 3; i >= 0; i--) {
		x = (x << 32) | a[i];
		y = (y << 32) | b[i];
	}
	orc_u128 z = orc_tower_mul(x, y, 7);
	for (int i = 0; i < 4; i++) out[i] = (uint32_t)(z >> (32 * i));
}

/* ---- fast 32-bit product: same recursion, h = 3 leaves from a 64 KiB table built once from
 *      orc_tower_mul (so it is the tabulation the reference itself uses at h=2, :133-151, one level up) ---- */
static uint8_t* g_mul8;      /* [a][b] */
static uint8_t g_alpha8[256]; /* multiply_alpha at h = 3 */

static void fast_init(void) {
	uint8_t* t = (uint8_t*)malloc(65536);
	for (int a = 0; a < 256; a++)
		for (int b = a; b < 256; b++) {
			uint8_t p = (uint8_t)orc_tower_mul((orc_u128)a, (orc_u128)b, 3);
			t[a * 256 + b] = p;
			t[b * 256 + a] = p;
		}
	for (int a = 0; a < 256; a++) g_alpha8[a] = (uint8_t)orc_tower_mul_alpha((orc_u128)a, 3);
	__atomic_store_n(&g_mul8, t, __ATOMIC_RELEASE);
}

static inline const uint8_t* mul8_table(void) {
	const uint8_t* t = __atomic_load_n(&g_mul8, __ATOMIC_ACQUIRE);
	if (!t) {
#pragma omp critical(orc_fast_init)
		{
			if (!__atomic_load_n(&g_mul8, __ATOMIC_ACQUIRE)) fast_init();
		}
		t = __atomic_load_n(&g_mul8, __ATOMIC_ACQUIRE);
	}
	return t;
}

static inline uint32_t alpha16(uint32_t a) { /* multiply_alpha at h = 4 */
	uint32_t a0 = a & 0xff, a1 = a >> 8;
	return a1 | ((a0 ^ g_alpha8[a1]) << 8);
}

static inline uint32_t mul16_fast(const uint8_t* T, uint32_t a, uint32_t b) {
	uint32_t a0 = a & 0xff, a1 = a >> 8, b0 = b & 0xff, b1 = b >> 8;
	uint32_t z0 = T[a0 * 256 + b0], z2 = T[a1 * 256 + b1];
	uint32_t z1 = T[(a0 ^ a1) * 256 + (b0 ^ b1)] ^ z0 ^ z2;
	return (z0 ^ z2) | ((z1 ^ g_alpha8[z2]) << 8);
}

uint32_t orc_mul32_fast(uint32_t a, uint32_t b) {
	const uint8_t* T = mul8_table();
	uint32_t a0 = a & 0xffff, a1 = a >> 16, b0 = b & 0xffff, b1 = b >> 16;
	uint32_t z0 = mul16_fast(T, a0, b0), z2 = mul16_fast(T, a1, b1);
	uint32_t z1 = mul16_fast(T, a0 ^ a1, b0 ^ b1) ^ z0 ^ z2;
	return (z0 ^ z2) | ((z1 ^ alpha16(z2)) << 16);
}

/* ---- packed-SIMD semantics (finite_fields/binary_tower_simd.cuh:82-127): every 2^h-bit lane of
 *      the word is an independent GF(2^(2^h)) element ---- */
uint32_t orc_mul_packed32(uint32_t a, uint32_t b, int h) {
	if (h < 0 || h > 5) return 0;
	int w = 1 << h;
	uint32_t out = 0;
	for (int off = 0; off < 32; off += w) {
		uint32_t m = (w == 32) ? 0xffffffffu : ((1u << w) - 1);
		uint32_t x = (a >> off) & m, y = (b >> off) & m;
		out |= (uint32_t)orc_tower_mul(x, y, h) << off;
	}
	return out;
}

void orc_interleave32(uint32_t* a, uint32_t* b, int h) { /* binary_tower_simd.cuh:129-139 */
	static const uint32_t masks[5] = {0x55555555u, 0x33333333u, 0x0f0f0f0fu, 0x00ff00ffu, 0x0000ffffu};
	if (h < 0 || h > 4) return;
	int bl = 1 << h;
	uint32_t t = ((*a >> bl) ^ *b) & masks[h];
	*a ^= t << bl;
	*b ^= t;
}
