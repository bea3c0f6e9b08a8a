/*
 * sumcheck.c -- CPU restatement, in COMPACT form, of the reference's bit-sliced GF(2^128) sumcheck prover
 * (sumcheck/sumcheck.cuh:10-300, sumcheck/core/core.cu:9-96, sumcheck/core/kernels.cuh:5-102) and of its test-side
 * verifier (sumcheck/test/verifier.cu:9-107, sumcheck/test/kernel/verifier_kernel.cu:5-37).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * The reference works on bit-sliced batches of 32 evaluations; every step is position-wise in the batch, so the
 * messages are functions of the compact values alone:
 *   this_round_messages (sumcheck.cuh:136-246), n = 2^num_vars >> round evaluations left per multilinear:
 *     sum       = XOR_{x < n}   PROD_c a_c[x]                                 (compute_compositions, kernels.cuh:24-40;
 *                                                                               compute_sum, core.cu:81-95)
 *     points[t] = XOR_{x < n/2} PROD_c (a_c[x] ^ t * (a_c[x] ^ a_c[x + n/2]))   t = 0..C   (fold_batch with
 *                 is_interpolation, core.cu:24-55: t is the field element whose low bits are the integer t)
 *   move_to_next_round (sumcheck.cuh:248-299): a_c[x] <- a_c[x] ^ r * (a_c[x] ^ a_c[x + n/2]), x < n/2
 *     (fold_large_list_halves, kernels.cu:9-30; fold_small, core.cu:57-79): the FIRST challenge binds the TOP index bit.
 * Parity status of this file: the reference has no golden vectors for the sumcheck; its test is the verifier's
 * consistency checks with random data (sumcheck/test/test.cu:16-96).  The restatement is pinned by passing exactly those
 * checks (tests/test_oracle.py) on top of the KAT-pinned GF(2^128) product.
 */
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

struct orc_sc {
	int num_vars, comp, round;
	size_t n0; /* column stride = 2^num_vars */
	orc_u128* a;
};
typedef struct orc_sc orc_sc;

static orc_u128 ld128(const uint32_t* w) {
	orc_u128 x = 0;
	for (int i = 3; i >= 0; i--) x = (x << 32) | w[i];
	return x;
}
static void st128(uint32_t* w, orc_u128 x) {
	for (int i = 0; i < 4; i++) w[i] = (uint32_t)(x >> (32 * i));
}

orc_sc* orc_sc_create(const uint32_t* evals, int num_vars, int comp) {
	if (num_vars < 0 || num_vars > 30 || comp < 1 || comp > 8) return NULL;
	orc_sc* s = (orc_sc*)calloc(1, sizeof(orc_sc));
	s->num_vars = num_vars, s->comp = comp, s->n0 = (size_t)1 << num_vars;
	s->a = (orc_u128*)malloc(sizeof(orc_u128) * s->n0 * comp);
	for (size_t i = 0; i < s->n0 * comp; i++) s->a[i] = ld128(evals + 4 * i);
	return s;
}

void orc_sc_destroy(orc_sc* s) {
	if (s) free(s->a), free(s);
}

void orc_sc_round_messages(const orc_sc* s, uint32_t sum[4], uint32_t* points) {
	const size_t n = s->n0 >> s->round, half = n / 2;
	const int C = s->comp;
	orc_u128 acc_sum = 0, acc_pt[9] = {0};
#pragma omp parallel
	{
		orc_u128 ls = 0, lp[9] = {0};
#pragma omp for schedule(static) nowait
		for (size_t x = 0; x < n; x++) {
			orc_u128 p = s->a[x];
			for (int c = 1; c < C; c++) p = orc_tower_mul(p, s->a[c * s->n0 + x], 7);
			ls ^= p;
		}
#pragma omp for schedule(static) nowait
		for (size_t x = 0; x < half; x++) {
			for (int t = 0; t <= C; t++) {
				orc_u128 p = 0;
				for (int c = 0; c < C; c++) {
					const orc_u128 lo = s->a[c * s->n0 + x], hi = s->a[c * s->n0 + x + half];
					const orc_u128 f = lo ^ orc_tower_mul(lo ^ hi, (orc_u128)t, 7);
					p = c ? orc_tower_mul(p, f, 7) : f;
				}
				lp[t] ^= p;
			}
		}
#pragma omp critical
		{
			acc_sum ^= ls;
			for (int t = 0; t <= C; t++) acc_pt[t] ^= lp[t];
		}
	}
	st128(sum, acc_sum);
	for (int t = 0; t <= C; t++) st128(points + 4 * t, acc_pt[t]);
}

void orc_sc_next_round(orc_sc* s, const uint32_t challenge[4]) {
	const size_t n = s->n0 >> s->round, half = n / 2;
	const orc_u128 r = ld128(challenge);
	for (int c = 0; c < s->comp; c++) {
		orc_u128* a = s->a + c * s->n0;
#pragma omp parallel for schedule(static)
		for (size_t x = 0; x < half; x++) a[x] ^= orc_tower_mul(a[x] ^ a[x + half], r, 7);
	}
	s->round++;
}

/* evaluate_univariate_given_points (verifier.cu:9-31): Lagrange interpolation through (t, points[t]), t = 0..np-1 */
void orc_sc_interpolate(const uint32_t challenge[4], const uint32_t* points, int num_points, uint32_t out[4]) {
	const orc_u128 r = ld128(challenge);
	orc_u128 e = 0;
	for (int i = 0; i < num_points; i++) {
		orc_u128 p = ld128(points + 4 * i);
		for (int j = 0; j < num_points; j++)
			if (j != i) {
				p = orc_tower_mul(p, r ^ (orc_u128)j, 7);
				p = orc_tower_mul(p, orc_tower_inverse((orc_u128)(i ^ j), 7), 7);
			}
		e ^= p;
	}
	st128(out, e);
}

/* evaluate_multilinear_composition (verifier.cu:92-107) with lagrange_basis_eval (verifier_kernel.cu:5-37): brute force,
 * index bit v pairs with challenge num_vars-1-v */
void orc_sc_eval_composition(const uint32_t* evals, const uint32_t* challenges, int num_vars, int comp, uint32_t out[4]) {
	const size_t n = (size_t)1 << num_vars;
	orc_u128 prod = 1;
	for (int c = 0; c < comp; c++) {
		orc_u128 acc = 0;
#pragma omp parallel
		{
			orc_u128 la = 0;
#pragma omp for schedule(static) nowait
			for (size_t x = 0; x < n; x++) {
				orc_u128 p = ld128(evals + 4 * (c * n + x));
				for (int v = 0; v < num_vars; v++) {
					const orc_u128 ch = ld128(challenges + 4 * (num_vars - 1 - v));
					p = orc_tower_mul(p, ((x >> v) & 1) ? ch : ch ^ 1, 7);
				}
				la ^= p;
			}
#pragma omp critical
			acc ^= la;
		}
		prod = orc_tower_mul(prod, acc, 7);
	}
	st128(out, prod);
}
