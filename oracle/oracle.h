/*
 * oracle.h -- CPU restatement of the reference's additive-NTT / tower-field hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and
 * there only as the checker or the timed CPU baseline.  The product (libb2ntt.so) never links,
 * loads or calls this code.
 *
 * Parity status: PINNED.  The NTT restatement reproduces the reference's golden MD5 table
 * (/root/reference/src/ulvt/ntt/tests/test_ntt.cu:52-124) and the field restatement reproduces
 * every known-answer vector of finite_fields/tests/{tests.cu,test_fanpaartower.cu}; both are also
 * cross-checked against the reference's own host-callable field code compiled into
 * oracle/_ref/libulvt_ref.so (see oracle/Makefile, oracle/ref_shim.cu).
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/src/ulvt/).
 */
#ifndef B2NTT_ORACLE_H
#define B2NTT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef unsigned __int128 orc_u128;

/* ---- compact Fan-Paar tower field GF(2^(2^h)), h = 0..7 (finite_fields/binary_tower.cuh:35-93,
 *      extended to h = 6,7 the way sumcheck/test/utils/unbitsliced_mul.cuh:35-93 and
 *      sumcheck/test/utils/tower_7_mul.cu:4-20 do) ---- */
orc_u128 orc_tower_mul(orc_u128 a, orc_u128 b, int h);
orc_u128 orc_tower_mul_alpha(orc_u128 a, int h);
orc_u128 orc_tower_square(orc_u128 a, int h);
orc_u128 orc_tower_inverse(orc_u128 a, int h);

/* 32-bit conveniences (h = 5) -- the fast variants use a GF(2^8) product table as recursion base. */
uint32_t orc_mul32(uint32_t a, uint32_t b);
uint32_t orc_square32(uint32_t a);
uint32_t orc_inverse32(uint32_t a);
uint32_t orc_mul32_fast(uint32_t a, uint32_t b);
/* word-pointer form of the 128-bit product: 4 little-endian limbs each */
void orc_mul128_words(const uint32_t a[4], const uint32_t b[4], uint32_t out[4]);

/* ---- packed-SIMD multiply (finite_fields/binary_tower_simd.cuh:77-149) ---- */
/* 32/2^h independent GF(2^(2^h)) lanes packed in one word, multiplied lane-wise */
uint32_t orc_mul_packed32(uint32_t a, uint32_t b, int h);
/* interleave_32b<h>: swap the odd 2^h-bit blocks of a with the even blocks of b (:129-139) */
void orc_interleave32(uint32_t* a, uint32_t* b, int h);

/* ---- little-endian bit-slicing (utils/bitslicing.cuh:14-64) ---- */
void orc_transpose32(uint32_t a[32]);
void orc_bitslice(uint32_t* block, int width_bits);   /* in place, width_bits words = 32 elements */
void orc_unbitslice(uint32_t* block, int width_bits); /* inverse */
/* bit-sliced tower product of 32 lanes: 2^h words each (multiply_unrolled<h> semantics,
 * finite_fields/circuit_generator/multiply_and_generate_circuit.cpp:62-155) */
void orc_mul_bitsliced(const uint32_t* a, const uint32_t* b, uint32_t* out, int h);

/* ---- BabyBear radix-2 NTT (ntt/gpuntt.cuh:12-209), canonical residues, bb31.c ---- */
uint32_t orc_bb31_mul(uint32_t a, uint32_t b);
uint32_t orc_bb31_pow(uint32_t x, uint64_t n);
int orc_bb31_ntt(const uint32_t* in, uint32_t* out, int log_n, uint32_t generator, int log_group_order, int input_bit_reversed);

/* ---- additive NTT (ntt/additive_ntt.cuh) ---- */
/* normalised subspace-polynomial constants C[i][j], i < log_h, j < log_h+log_rate-1-i, row pitch
 * = log_h+log_rate-1 words (additive_ntt.cuh:273-309).  Returns the pitch. */
int orc_antt_constants(int log_h, int log_rate, uint32_t* out /* log_h * pitch */);
/* twiddle of (stage, coset, butterfly block)  (additive_ntt.cuh:59-77) */
uint32_t orc_antt_twiddle(const uint32_t* constants, int pitch, int log_h, int log_rate, int coset, int stage,
                          uint64_t block);
/* forward transform.  limbs = 1 (GF(2^32)) or 4 (GF(2^128), element = 4 LE words).
 * in: 2^log_h elements, out: 2^(log_h+log_rate) elements, coset-major (additive_ntt.cuh:213-215,
 * 250-259).  threads <= 0: use all cores (OpenMP).  Returns 0, or -1 on bad arguments. */
int orc_antt_forward(const uint32_t* in, uint32_t* out, int log_h, int log_rate, int limbs, int threads);
/* 1: route every product of orc_antt_forward/inverse through the definitional orc_mul32 (slow) */
void orc_antt_set_plain(int on);
/* same transform, but each butterfly multiplies with the full GF(2^128) tower product
 * (definitional form; slow; limbs must be 4). */
int orc_antt_forward_full128(const uint32_t* in, uint32_t* out, int log_h, int log_rate);
/* inverse of orc_antt_forward on ONE coset row: out (2^log_h) from evaluations on
 * coset `coset`. */
int orc_antt_inverse(const uint32_t* in, uint32_t* out, int log_h, int log_rate, int coset, int limbs,
                     int threads);


/* ---- general-basis GF(2^128) additive NTT (antt_general.c) ---- */
int orc_antt_general_constants(int log_h, const uint32_t* basis /* log_h * 4 or NULL */, uint32_t* C /* log_h*log_h*4 */);
int orc_antt_general(const uint32_t* in, uint32_t* out, int log_h, const uint32_t* basis, int inverse);
int orc_antt_general_direct(const uint32_t* in, uint32_t* out, int log_h, const uint32_t* basis);

/* ---- MD5 (ntt/tests/md5.cu; RFC 1321) and the reference's test-vector stream ---- */
void orc_md5(const void* data, size_t len, uint8_t digest[16]);
void orc_md5_hex(const void* data, size_t len, char hex[33]);
/* std::mt19937 stream, one 32-bit draw per word (ntt/tests/test_ntt.cu:192-199) */
void orc_mt19937_fill(uint32_t seed, uint32_t* out, size_t n);
/* golden digest for (log_rate, log_h) or NULL (ntt/tests/test_ntt.cu:52-124) */
const char* orc_golden_md5(int log_rate, int log_h);

int orc_num_threads(void);

/* ---- bit-sliced GF(2^128) sumcheck prover + verifier, compact restatement (sumcheck.c; sumcheck/sumcheck.cuh:10-300,
 *      sumcheck/core/core.cu:9-96, sumcheck/test/verifier.cu:9-107) ---- */
struct orc_sc;
struct orc_sc* orc_sc_create(const uint32_t* evals /* comp columns of 2^num_vars compact elements */, int num_vars, int comp);
void orc_sc_destroy(struct orc_sc* s);
void orc_sc_round_messages(const struct orc_sc* s, uint32_t sum[4], uint32_t* points /* (comp + 1) * 4 */);
void orc_sc_next_round(struct orc_sc* s, const uint32_t challenge[4]);
void orc_sc_interpolate(const uint32_t challenge[4], const uint32_t* points, int num_points, uint32_t out[4]);
void orc_sc_eval_composition(const uint32_t* evals, const uint32_t* challenges, int num_vars, int comp, uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
