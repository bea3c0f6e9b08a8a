/* antt_general.c -- TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
 *
 * Additive NTT over GF(2^128) on an arbitrary GF(2)-independent basis beta_0..beta_{log_h-1}.
 * The reference only has the basis beta_i = 1 << i (ntt/additive_ntt.cuh:281-283); this restates the same
 * recurrence and butterfly with full-width constants:
 *   constants      ntt/additive_ntt.cuh:273-309 (precompute_subspace_evals), subspace_map :16-19
 *   twiddle        :59-77 (XOR of the constants selected by the bits of the block index)
 *   butterfly      :10-14, stage order high -> low :233-247
 * PARITY: unpinned by the reference (it has no full-width transform).  Pinned instead (tests/test_general.py) to
 *   (1) orc_antt_forward_full128 -- itself pinned to the reference's MD5 digests -- on the reference basis, and
 *   (2) orc_antt_general_direct below: the definition (evaluate sum_i c_i X_i at every domain point, X_i = product of the
 *       normalised subspace polynomials W^_b over the set bits b of i, W_b(x) = prod_{u in span(beta_0..beta_{b-1})} (x + u)).
 */
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

static orc_u128 ld(const uint32_t* w) {
	return (orc_u128)w[0] | (orc_u128)w[1] << 32 | (orc_u128)w[2] << 64 | (orc_u128)w[3] << 96;
}
static void st(uint32_t* w, orc_u128 x) {
	for (int l = 0; l < 4; l++) w[l] = (uint32_t)(x >> (32 * l));
}
static orc_u128 mul(orc_u128 a, orc_u128 b) { return orc_tower_mul(a, b, 7); }

/* C: [log_h][log_h] elements (4 words each), row i holds log_h - 1 - i entries.  basis NULL = 1 << i.
 * Returns 0, or -1 when the basis is linearly dependent. */
int orc_antt_general_constants(int log_h, const uint32_t* basis, uint32_t* C) {
	orc_u128* cur = (orc_u128*)malloc(sizeof(orc_u128) * log_h);
	memset(C, 0, sizeof(uint32_t) * 4 * (size_t)log_h * log_h);
	for (int k = 0; k < log_h; k++) cur[k] = basis ? ld(basis + 4 * k) : (orc_u128)1 << k;
	for (int i = 0; i < log_h; i++) {
		if (i > 0) { /* :286-298: E_i(x) = E_{i-1}(x)^2 + E_{i-1}(beta_{i-1}) E_{i-1}(x) */
			orc_u128 np = cur[i - 1];
			for (int k = i; k < log_h; k++) cur[k] = orc_tower_square(cur[k], 7) ^ mul(np, cur[k]);
		}
		if (cur[i] == 0) {
			free(cur);
			return -1;
		}
		orc_u128 inv = orc_tower_inverse(cur[i], 7); /* :300-306 */
		for (int k = i + 1; k < log_h; k++) st(C + 4 * ((size_t)i * log_h + (k - i - 1)), mul(inv, cur[k]));
	}
	free(cur);
	return 0;
}

int orc_antt_general(const uint32_t* in, uint32_t* out, int log_h, const uint32_t* basis, int inverse) {
	const uint64_t n = (uint64_t)1 << log_h;
	uint32_t* C = (uint32_t*)malloc(sizeof(uint32_t) * 4 * (size_t)log_h * log_h);
	if (orc_antt_general_constants(log_h, basis, C)) {
		free(C);
		return -1;
	}
	if (in != out) memcpy(out, in, n * 16);
	for (int s = 0; s < log_h; s++) {
		const int stage = inverse ? s : log_h - 1 - s;
		const int64_t blocks = (int64_t)(n >> (stage + 1));
#pragma omp parallel for schedule(static)
		for (int64_t j = 0; j < blocks; j++) {
			orc_u128 w = 0;
			for (int b = 0; b < log_h - 1 - stage; b++)
				if ((j >> b) & 1) w ^= ld(C + 4 * ((size_t)stage * log_h + b));
			for (uint64_t k = 0; k < ((uint64_t)1 << stage); k++) {
				uint32_t* pu = out + (((uint64_t)j << (stage + 1)) + k) * 4;
				uint32_t* pv = pu + ((uint64_t)4 << stage);
				orc_u128 u = ld(pu), v = ld(pv);
				if (inverse) {
					v ^= u;
					u ^= mul(w, v);
				} else {
					u ^= mul(w, v);
					v ^= u;
				}
				st(pu, u), st(pv, v);
			}
		}
	}
	free(C);
	return 0;
}

/* the definition, O(4^log_h): for log_h <= 8 */
int orc_antt_general_direct(const uint32_t* in, uint32_t* out, int log_h, const uint32_t* basis) {
	if (log_h > 8) return -1;
	const int n = 1 << log_h;
	orc_u128 beta[8], What[8]; /* What[b] at the current point */
	for (int k = 0; k < log_h; k++) beta[k] = basis ? ld(basis + 4 * k) : (orc_u128)1 << k;
	for (int p = 0; p < n; p++) {
		orc_u128 x = 0;
		for (int b = 0; b < log_h; b++)
			if ((p >> b) & 1) x ^= beta[b];
		for (int b = 0; b < log_h; b++) {
			orc_u128 num = 1, den = 1;
			for (int m = 0; m < (1 << b); m++) {
				orc_u128 u = 0;
				for (int c = 0; c < b; c++)
					if ((m >> c) & 1) u ^= beta[c];
				num = mul(num, x ^ u);
				den = mul(den, beta[b] ^ u);
			}
			if (den == 0) return -1;
			What[b] = mul(num, orc_tower_inverse(den, 7));
		}
		orc_u128 acc = 0;
		for (int i = 0; i < n; i++) {
			orc_u128 X = 1;
			for (int b = 0; b < log_h; b++)
				if ((i >> b) & 1) X = mul(X, What[b]);
			acc ^= mul(ld(in + 4 * i), X);
		}
		st(out + 4 * p, acc);
	}
	return 0;
}
