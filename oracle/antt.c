/*
 * antt.c -- additive (novel-polynomial-basis) NTT over the binary tower, CPU oracle.
 * TEST INFRASTRUCTURE ONLY.
 *
 * The reference has no CPU transform (its tests compare MD5 digests "pulled from the python
 * model", ntt/tests/test_ntt.cu:20); this file restates the arithmetic of its one GPU kernel and
 * host driver, ntt/additive_ntt.cuh:
 *   butterfly            antt_butterfly           :10-14     u ^= w*v ; v ^= u
 *   subspace map         subspace_map             :16-19     e^2 + c*e
 *   index algebra        get_*_offset/_block      :21-33
 *   twiddle              calculate_twiddle        :59-77     XOR of constants over the set bits of
 *                                                            (coset << (log_h-1-stage)) | block
 *   stage loop           additive_ntt_kernel      :127-159   high stage first (:233-247)
 *   domain constants     precompute_subspace_evals :273-309  basis beta_i = 1 << i
 *   replication / order  apply                    :213-215, :250-259  out[c * 2^log_h + i]
 * and reproduces the golden digests of test_ntt.cu:52-124 (tests/test_oracle_ntt.py).
 *
 * GF(2^128) elements are four little-endian 32-bit limbs.  All twiddles of a domain with
 * log_h + log_rate <= 32 lie in the GF(2^32) subfield, where the tower product acts limb-wise;
 * orc_antt_forward uses that, orc_antt_forward_full128 does not (it multiplies in GF(2^128)) and
 * the two are compared in the tests.
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int orc_num_threads(void) {
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}

static int args_ok(int log_h, int log_rate) {
	/* nttconf.cuh:55-60 */
	return log_h >= 1 && log_rate >= 0 && log_rate <= 4 && log_h + log_rate <= 32;
}

int orc_antt_constants(int log_h, int log_rate, uint32_t* C) {
	int pitch = log_h + log_rate - 1;
	if (!args_ok(log_h, log_rate)) return -1;
	if (pitch == 0) return 0;
	uint32_t* norm = (uint32_t*)malloc(sizeof(uint32_t) * log_h);
	memset(C, 0, sizeof(uint32_t) * (size_t)log_h * pitch);
	for (int i = 1; i < log_h + log_rate; i++) C[i - 1] = (uint32_t)1 << i; /* :281-283 */
	norm[0] = 1;
	for (int i = 1; i < log_h; i++) { /* :286-298 */
		const uint32_t* prev = C + (size_t)(i - 1) * pitch;
		uint32_t* row = C + (size_t)i * pitch;
		uint32_t np = norm[i - 1];
		norm[i] = orc_square32(prev[0]) ^ orc_mul32(np, prev[0]);
		for (int j = 1; j < log_h + log_rate - i; j++) row[j - 1] = orc_square32(prev[j]) ^ orc_mul32(np, prev[j]);
	}
	for (int i = 0; i < log_h; i++) { /* :300-306 */
		uint32_t inv = orc_inverse32(norm[i]);
		uint32_t* row = C + (size_t)i * pitch;
		for (int j = 0; j < log_h + log_rate - i - 1; j++) row[j] = orc_mul32(inv, row[j]);
	}
	free(norm);
	return pitch;
}

uint32_t orc_antt_twiddle(const uint32_t* C, int pitch, int log_h, int log_rate, int coset, int stage,
                          uint64_t block) {
	uint64_t ind = ((uint64_t)coset << (log_h - 1 - stage)) | block; /* :71 */
	uint32_t w = 0;
	for (int k = 0; k < log_h + log_rate - 1 - stage; k++)
		if ((ind >> k) & 1) w ^= C[(size_t)stage * pitch + k];
	return w;
}

/* w * x for many x: four 256-entry window tables of w*(byte << 8k).  Pure speed-up of the scalar
 * product (linearity of x -> w*x over GF(2)); tests compare it with orc_mul32. */
typedef struct {
	uint32_t t[4][256];
} wtab;

static void wtab_build(wtab* T, uint32_t w) {
	for (int k = 0; k < 4; k++) {
		T->t[k][0] = 0;
		for (int b = 0; b < 8; b++) {
			uint32_t p = orc_mul32_fast(w, (uint32_t)1 << (8 * k + b));
			int span = 1 << b;
			for (int i = 0; i < span; i++) T->t[k][span + i] = T->t[k][i] ^ p;
		}
	}
}

static inline uint32_t wtab_mul(const wtab* T, uint32_t x) {
	return T->t[0][x & 0xff] ^ T->t[1][(x >> 8) & 0xff] ^ T->t[2][(x >> 16) & 0xff] ^ T->t[3][x >> 24];
}

static int g_plain = 0; /* 1: every product through orc_mul32 (no tables, no fast path) */
void orc_antt_set_plain(int on) { g_plain = on; }

/* one stage over rows x 2^log_h elements; dir = +1 forward butterfly, -1 inverse butterfly.
 * coset0 = first coset index of `data` rows. */
static void run_stage(uint32_t* data, const uint32_t* C, int pitch, int log_h, int log_rate, int rows, int coset0,
                      int stage, int limbs, int dir, int threads) {
	const uint64_t n = (uint64_t)1 << log_h;
	const uint64_t half = n >> 1;
	const uint64_t total = (uint64_t)rows * half; /* butterflies */
	const uint64_t run = (uint64_t)1 << stage;    /* butterflies sharing one twiddle */
	const uint64_t chunk = run < 4096 ? 4096 : (run < 65536 ? run : 65536);
	const int64_t nchunks = (int64_t)((total + chunk - 1) / chunk);
	const int use_tab = !g_plain && run * (uint64_t)limbs >= 1024;
	(void)threads;
#pragma omp parallel num_threads(threads)
	{
		wtab* T = use_tab ? (wtab*)malloc(sizeof(wtab)) : 0;
#pragma omp for schedule(static)
		for (int64_t ci = 0; ci < nchunks; ci++) {
			uint64_t t0 = (uint64_t)ci * chunk, t1 = t0 + chunk;
			if (t1 > total) t1 = total;
			uint64_t t = t0;
			while (t < t1) {
				uint64_t row = t / half, rem = t % half;
				uint64_t j = rem >> stage, k0 = rem & (run - 1);
				uint64_t cnt = run - k0;
				if (cnt > t1 - t) cnt = t1 - t;
				uint32_t w = orc_antt_twiddle(C, pitch, log_h, log_rate, coset0 + (int)row, stage, j);
				uint32_t* base = data + (row * n + (j << (stage + 1)) + k0) * limbs;
				uint32_t* u = base;
				uint32_t* v = base + (run * limbs);
				uint64_t words = cnt * limbs;
				if (use_tab) wtab_build(T, w);
				for (uint64_t x = 0; x < words; x++) {
					if (dir > 0) { /* :10-14 */
						uint32_t p = use_tab ? wtab_mul(T, v[x]) : (g_plain ? orc_mul32(w, v[x]) : orc_mul32_fast(w, v[x]));
						u[x] ^= p;
						v[x] ^= u[x];
					} else { /* exact inverse of the above */
						v[x] ^= u[x];
						uint32_t p = use_tab ? wtab_mul(T, v[x]) : (g_plain ? orc_mul32(w, v[x]) : orc_mul32_fast(w, v[x]));
						u[x] ^= p;
					}
				}
				t += cnt;
			}
		}
		free(T);
	}
}

int orc_antt_forward(const uint32_t* in, uint32_t* out, int log_h, int log_rate, int limbs, int threads) {
	if (!args_ok(log_h, log_rate) || (limbs != 1 && limbs != 4)) return -1;
	if (threads <= 0) threads = orc_num_threads();
	const uint64_t n = (uint64_t)1 << log_h;
	const int rows = 1 << log_rate;
	int pitch = log_h + log_rate - 1;
	uint32_t* C = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)log_h * (pitch ? pitch : 1));
	orc_antt_constants(log_h, log_rate, C);
	orc_mul32_fast(1, 1); /* build the base table outside the parallel region */
	for (int c = 0; c < rows; c++) memcpy(out + (uint64_t)c * n * limbs, in, n * limbs * sizeof(uint32_t)); /* :213-215 */
	for (int stage = log_h - 1; stage >= 0; stage--) /* :138, :233-247 */
		run_stage(out, C, pitch, log_h, log_rate, rows, 0, stage, limbs, +1, threads);
	free(C);
	return 0;
}

int orc_antt_inverse(const uint32_t* in, uint32_t* out, int log_h, int log_rate, int coset, int limbs, int threads) {
	if (!args_ok(log_h, log_rate) || (limbs != 1 && limbs != 4) || coset < 0 || coset >= (1 << log_rate)) return -1;
	if (threads <= 0) threads = orc_num_threads();
	const uint64_t n = (uint64_t)1 << log_h;
	int pitch = log_h + log_rate - 1;
	uint32_t* C = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)log_h * (pitch ? pitch : 1));
	orc_antt_constants(log_h, log_rate, C);
	orc_mul32_fast(1, 1);
	if (out != in) memcpy(out, in, n * limbs * sizeof(uint32_t));
	for (int stage = 0; stage < log_h; stage++)
		run_stage(out, C, pitch, log_h, log_rate, 1, coset, stage, limbs, -1, threads);
	free(C);
	return 0;
}

int orc_antt_forward_full128(const uint32_t* in, uint32_t* out, int log_h, int log_rate) {
	if (!args_ok(log_h, log_rate)) return -1;
	const uint64_t n = (uint64_t)1 << log_h;
	const int rows = 1 << log_rate;
	int pitch = log_h + log_rate - 1;
	uint32_t* C = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)log_h * (pitch ? pitch : 1));
	orc_antt_constants(log_h, log_rate, C);
	for (int c = 0; c < rows; c++) memcpy(out + (uint64_t)c * n * 4, in, n * 4 * sizeof(uint32_t));
	for (int stage = log_h - 1; stage >= 0; stage--)
		for (int c = 0; c < rows; c++)
			for (uint64_t j = 0; j < (n >> (stage + 1)); j++) {
				uint32_t w[4] = {orc_antt_twiddle(C, pitch, log_h, log_rate, c, stage, j), 0, 0, 0};
				for (uint64_t k = 0; k < ((uint64_t)1 << stage); k++) {
					uint32_t* u = out + ((uint64_t)c * n + (j << (stage + 1)) + k) * 4;
					uint32_t* v = u + ((uint64_t)4 << stage);
					uint32_t p[4];
					orc_mul128_words(w, v, p);
					for (int l = 0; l < 4; l++) {
						u[l] ^= p[l];
						v[l] ^= u[l];
					}
				}
			}
	free(C);
	return 0;
}
