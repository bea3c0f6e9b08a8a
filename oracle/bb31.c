/*
 * bb31.c -- CPU restatement of the reference's BabyBear radix-2 NTT (TEST INFRASTRUCTURE ONLY, see oracle.h).
 *
 *   field        BB31 = risc0::Fp, p = 15 * 2^27 + 1       finite_fields/risc0_baby_bear.h:40-187 (values here are CANONICAL
 *                                                            residues 0..p-1; the reference keeps them in Montgomery form)
 *   bit reversal bit_reverse_in_place_ker                    ntt/gpuntt.cuh:12-34
 *   butterfly    U = u + v ; V = (u - v) * w                 ntt/gpuntt.cuh:36-44
 *   indexing     u = (t mod 2^s) | (t >> s) << (s + 1), v = u + 2^s, w = tw_br[t >> s]     ntt/gpuntt.cuh:46-63, 107-118
 *   twiddles     tw[i] = (g^(2^(G - L)))^i, stored bit-reversed over L - 1 bits              ntt/gpuntt.cuh:132-143, 190-207
 *   driver       bit-reverse an IN_ORDER input, stages 0 .. L-1                             ntt/gpuntt.cuh:158-183
 *
 * Pinned: reproduces bb31_ntt_hashes[1..27] (ntt/tests/test_ntt.cu:21-50; tests/golden/bb31_ntt_md5.json).
 */
#include <stdlib.h>

#include "oracle.h"

#define BB31_P 2013265921u

uint32_t orc_bb31_mul(uint32_t a, uint32_t b) { return (uint32_t)((uint64_t)a * b % BB31_P); }
uint32_t orc_bb31_pow(uint32_t x, uint64_t n) {
	uint32_t r = 1;
	while (n) {
		if (n & 1) r = orc_bb31_mul(r, x);
		x = orc_bb31_mul(x, x);
		n >>= 1;
	}
	return r;
}
static uint32_t rev_bits(uint32_t x, int bits) {
	uint32_t r = 0;
	for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
	return r;
}

/* in: n = 2^log_n canonical residues (any 32-bit word is reduced first, as BB31(uint32_t) does); out: canonical residues */
int orc_bb31_ntt(const uint32_t* in, uint32_t* out, int log_n, uint32_t generator, int log_group_order, int input_bit_reversed) {
	if (log_n < 1 || log_n > 27 || log_group_order < log_n) return -1;
	const uint32_t n = 1u << log_n, half = n >> 1;
	for (uint32_t i = 0; i < n; i++) out[input_bit_reversed ? i : rev_bits(i, log_n)] = in[i] % BB31_P;
	uint32_t* tw = (uint32_t*)malloc(sizeof(uint32_t) * half);
	if (!tw) return -2;
	const uint32_t w = orc_bb31_pow(generator % BB31_P, (uint64_t)1 << (log_group_order - log_n));
	uint32_t cur = 1;
	for (uint32_t i = 0; i < half; i++) {
		tw[rev_bits(i, log_n - 1)] = cur;
		cur = orc_bb31_mul(cur, w);
	}
	for (int s = 0; s < log_n; s++) {
#pragma omp parallel for schedule(static)
		for (uint32_t t = 0; t < half; t++) {
			const uint32_t u = (t & ((1u << s) - 1)) | ((t >> s) << (s + 1)), v = u + (1u << s);
			const uint32_t a = out[u], b = out[v];
			const uint32_t sum = a + b >= BB31_P ? a + b - BB31_P : a + b;
			const uint32_t dif = a >= b ? a - b : a + BB31_P - b;
			out[u] = sum;
			out[v] = orc_bb31_mul(dif, tw[t >> s]);
		}
	}
	free(tw);
	return 0;
}
