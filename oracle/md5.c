/*
 * md5.c -- RFC 1321 MD5 and the std::mt19937 stream, CPU oracle.  TEST INFRASTRUCTURE ONLY.
 *
 * The reference pins its NTT outputs by MD5 digests (ntt/tests/test_ntt.cu:208-214, MD5 code in
 * ntt/tests/md5.cu) over inputs drawn from std::mt19937 (test_ntt.cu:192-199).  Both algorithms
 * are public standards; this is an independent implementation from the RFC / the MT paper.
 */
#include "oracle.h"

#include <stdio.h>
#include <string.h>

#include "golden_md5.h"

static inline uint32_t rotl32(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }

static const uint8_t MD5_SHIFT[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                                      5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                      4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                      6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
static uint32_t MD5_K[64];
static int md5_k_ready;

static void md5_init_k(void) {
	/* K[i] = floor(2^32 * |sin(i+1)|); computed with a small fixed-point-safe long double sine */
	extern long double sinl(long double);
	for (int i = 0; i < 64; i++) {
		long double s = sinl((long double)(i + 1));
		if (s < 0) s = -s;
		MD5_K[i] = (uint32_t)(uint64_t)(s * 4294967296.0L);
	}
	md5_k_ready = 1;
}

static void md5_block(uint32_t st[4], const uint8_t* p) {
	uint32_t w[16];
	for (int i = 0; i < 16; i++)
		w[i] = (uint32_t)p[4 * i] | (uint32_t)p[4 * i + 1] << 8 | (uint32_t)p[4 * i + 2] << 16 |
		       (uint32_t)p[4 * i + 3] << 24;
	uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
	for (int i = 0; i < 64; i++) {
		uint32_t f;
		int g;
		if (i < 16) {
			f = (b & c) | (~b & d);
			g = i;
		} else if (i < 32) {
			f = (d & b) | (~d & c);
			g = (5 * i + 1) & 15;
		} else if (i < 48) {
			f = b ^ c ^ d;
			g = (3 * i + 5) & 15;
		} else {
			f = c ^ (b | ~d);
			g = (7 * i) & 15;
		}
		uint32_t t = d;
		d = c;
		c = b;
		b = b + rotl32(a + f + MD5_K[i] + w[g], MD5_SHIFT[i]);
		a = t;
	}
	st[0] += a;
	st[1] += b;
	st[2] += c;
	st[3] += d;
}

void orc_md5(const void* data, size_t len, uint8_t digest[16]) {
	if (!md5_k_ready) md5_init_k();
	uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
	const uint8_t* p = (const uint8_t*)data;
	size_t n = len;
	while (n >= 64) {
		md5_block(st, p);
		p += 64;
		n -= 64;
	}
	uint8_t tail[128];
	memset(tail, 0, sizeof tail);
	memcpy(tail, p, n);
	tail[n] = 0x80;
	size_t tl = (n < 56) ? 64 : 128;
	uint64_t bits = (uint64_t)len * 8;
	for (int i = 0; i < 8; i++) tail[tl - 8 + i] = (uint8_t)(bits >> (8 * i));
	md5_block(st, tail);
	if (tl == 128) md5_block(st, tail + 64);
	for (int i = 0; i < 4; i++)
		for (int j = 0; j < 4; j++) digest[4 * i + j] = (uint8_t)(st[i] >> (8 * j));
}

void orc_md5_hex(const void* data, size_t len, char hex[33]) {
	uint8_t d[16];
	orc_md5(data, len, d);
	for (int i = 0; i < 16; i++) sprintf(hex + 2 * i, "%02x", d[i]);
	hex[32] = 0;
}

/* std::mt19937 (Matsumoto-Nishimura MT19937, 32-bit), seeded as std::mt19937(seed). */
void orc_mt19937_fill(uint32_t seed, uint32_t* out, size_t n) {
	enum { N = 624, M = 397 };
	static const uint32_t MAG = 0x9908b0dfu;
	uint32_t mt[N];
	mt[0] = seed;
	for (int i = 1; i < N; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
	int idx = N;
	for (size_t k = 0; k < n; k++) {
		if (idx >= N) {
			for (int i = 0; i < N; i++) {
				uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % N] & 0x7fffffffu);
				mt[i] = mt[(i + M) % N] ^ (y >> 1) ^ ((y & 1) ? MAG : 0);
			}
			idx = 0;
		}
		uint32_t y = mt[idx++];
		y ^= y >> 11;
		y ^= (y << 7) & 0x9d2c5680u;
		y ^= (y << 15) & 0xefc60000u;
		y ^= y >> 18;
		out[k] = y;
	}
}

const char* orc_golden_md5(int log_rate, int log_h) {
	if (log_rate < 0 || log_rate > 2 || log_h < 0 || log_h > 30) return 0;
	return ORC_GOLDEN_MD5[log_rate][log_h];
}
