// b2_field.cuh -- binary tower field arithmetic for the B200 additive-NTT library.
//
// Two forms of the Fan-Paar / Wiedemann tower GF(2^(2^h)) = GF(2^(2^(h-1)))[X]/(X^2 + a*X + 1),
// a = generator of the level below, element = lo | hi << 2^(h-1)
// (reference semantics: finite_fields/binary_tower.cuh:35-93):
//
//  * compact scalar   -- one element per machine word; a table-free Karatsuba recursion in
//                        registers.  Used for plan building on the host (domain constants,
//                        additive_ntt.cuh:273-309), for tiny transforms and for the element-wise
//                        compact multiply entry point.
//  * bit-sliced       -- 32 independent elements per call, 2^h words, word i = bit i of all 32
//                        (layout of utils/bitslicing.cuh:32-47).  Straight-line XOR/AND networks
//                        written as fully unrolled templates so that ptxas maps them onto LOP3
//                        (three-input logic) and keeps every word in a register.  This is the
//                        multiply of the butterfly kernels and of multiply_unrolled<h>
//                        (circuit_generator/unrolled/*.cu), re-derived, not generated.
//
// Everything is __host__ __device__ and header-only; the same source is compiled by nvcc for
// sm_100a and by g++ for the plan builder and the CPU-side unit tests.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline __attribute__((always_inline))
#endif

namespace b2 {

// ------------------------------------------------------------------------------------------------
// compact scalar tower arithmetic (table-free)
// ------------------------------------------------------------------------------------------------
template <int H>
struct Tower {
	static constexpr int HALF = 1 << (H - 1);
	static constexpr uint64_t MASK = (uint64_t(1) << HALF) - 1;
	using Sub = Tower<H - 1>;

	static B2_HD uint64_t mul_alpha(uint64_t a) {
		uint64_t a0 = a & MASK, a1 = (a >> HALF) & MASK;
		return a1 | ((a0 ^ Sub::mul_alpha(a1)) << HALF);
	}
	static B2_HD uint64_t mul(uint64_t a, uint64_t b) {
		uint64_t a0 = a & MASK, a1 = (a >> HALF) & MASK, b0 = b & MASK, b1 = (b >> HALF) & MASK;
		uint64_t z0 = Sub::mul(a0, b0), z2 = Sub::mul(a1, b1);
		uint64_t zz = Sub::mul(a0 ^ a1, b0 ^ b1);
		return (z0 ^ z2) | ((zz ^ z0 ^ z2 ^ Sub::mul_alpha(z2)) << HALF);
	}
	static B2_HD uint64_t square(uint64_t a) {
		uint64_t a0 = a & MASK, a1 = (a >> HALF) & MASK;
		uint64_t z0 = Sub::square(a0), z2 = Sub::square(a1);
		return (z0 ^ z2) | (Sub::mul_alpha(z2) << HALF);
	}
	// inverse through the norm to the subfield; inverse(0) = 0
	static B2_HD uint64_t inverse(uint64_t a) {
		uint64_t a0 = a & MASK, a1 = (a >> HALF) & MASK;
		uint64_t inter = a0 ^ Sub::mul_alpha(a1);
		uint64_t delta = Sub::mul(a0, inter) ^ Sub::square(a1);
		uint64_t dinv = Sub::inverse(delta);
		return Sub::mul(dinv, inter) | (Sub::mul(dinv, a1) << HALF);
	}
};
template <>
struct Tower<0> {
	static B2_HD uint64_t mul_alpha(uint64_t a) { return a & 1; }
	static B2_HD uint64_t mul(uint64_t a, uint64_t b) { return a & b & 1; }
	static B2_HD uint64_t square(uint64_t a) { return a & 1; }
	static B2_HD uint64_t inverse(uint64_t a) { return a & 1; }
};

B2_HD uint32_t gf32_mul(uint32_t a, uint32_t b) { return (uint32_t)Tower<5>::mul(a, b); }
B2_HD uint32_t gf32_square(uint32_t a) { return (uint32_t)Tower<5>::square(a); }
B2_HD uint32_t gf32_inverse(uint32_t a) { return (uint32_t)Tower<5>::inverse(a); }

// GF(2^128) element: four little-endian 32-bit limbs.  Product by two Karatsuba levels over GF(2^32).
struct gf128 {
	uint32_t w[4];
};
B2_HD uint64_t gf64_mul(uint64_t a, uint64_t b) { return Tower<6>::mul(a, b); }
B2_HD gf128 gf128_mul(const gf128& a, const gf128& b) {
	uint64_t a0 = a.w[0] | (uint64_t)a.w[1] << 32, a1 = a.w[2] | (uint64_t)a.w[3] << 32;
	uint64_t b0 = b.w[0] | (uint64_t)b.w[1] << 32, b1 = b.w[2] | (uint64_t)b.w[3] << 32;
	uint64_t z0 = gf64_mul(a0, b0), z2 = gf64_mul(a1, b1), zz = gf64_mul(a0 ^ a1, b0 ^ b1);
	uint64_t lo = z0 ^ z2, hi = zz ^ z0 ^ z2 ^ Tower<6>::mul_alpha(z2);
	gf128 r;
	r.w[0] = (uint32_t)lo;
	r.w[1] = (uint32_t)(lo >> 32);
	r.w[2] = (uint32_t)hi;
	r.w[3] = (uint32_t)(hi >> 32);
	return r;
}

// ------------------------------------------------------------------------------------------------
// bit-sliced networks: N = 2^H words per operand, word i = bit i of 32 lanes
// ------------------------------------------------------------------------------------------------
template <int H>
struct Sliced {
	static constexpr int N = 1 << H;
	static constexpr int M = N / 2;
	using Sub = Sliced<H - 1>;

	// out = alpha_H * a, (a0, a1) -> (a1, a0 + alpha_{H-1} * a1); out must not alias a
	static B2_HD void mul_alpha(const uint32_t* a, uint32_t* out) {
		uint32_t t[M];
		Sub::mul_alpha(a + M, t);
#pragma unroll
		for (int i = 0; i < M; i++) {
			out[i] = a[M + i];
			out[M + i] = a[i] ^ t[i];
		}
	}

	// out = a * b (Karatsuba); out may alias a or b
	static B2_HD void mul(const uint32_t* a, const uint32_t* b, uint32_t* out) {
		uint32_t sa[M], sb[M], z0[M], z2[M], zz[M], za[M];
#pragma unroll
		for (int i = 0; i < M; i++) {
			sa[i] = a[i] ^ a[M + i];
			sb[i] = b[i] ^ b[M + i];
		}
		Sub::mul(a, b, z0);
		Sub::mul(a + M, b + M, z2);
		Sub::mul(sa, sb, zz);
		Sub::mul_alpha(z2, za);
#pragma unroll
		for (int i = 0; i < M; i++) {
			uint32_t lo = z0[i] ^ z2[i];
			out[i] = lo;
			out[M + i] = zz[i] ^ lo ^ za[i];
		}
	}

	// acc ^= a * b
	static B2_HD void mul_acc(const uint32_t* a, const uint32_t* b, uint32_t* acc) {
		uint32_t p[N];
		mul(a, b, p);
#pragma unroll
		for (int i = 0; i < N; i++) acc[i] ^= p[i];
	}
};
template <>
struct Sliced<0> {
	static B2_HD void mul_alpha(const uint32_t* a, uint32_t* out) { out[0] = a[0]; }
	static B2_HD void mul(const uint32_t* a, const uint32_t* b, uint32_t* out) { out[0] = a[0] & b[0]; }
};

// 32-word operand v (GF(2^32) lanes) times a twiddle w that lies in the subfield GF(2^(2^HW)):
// the product acts independently on every 2^HW-word chunk of v (tower limbs).
// w has 2^HW words.  HW = 5 is the full product.
template <int HW>
B2_HD void sliced_mul_subfield(const uint32_t* w, const uint32_t* v, uint32_t* out) {
	constexpr int C = 1 << HW;
#pragma unroll
	for (int c = 0; c < 32 / C; c++) Sliced<HW>::mul(w, v + c * C, out + c * C);
}

}  // namespace b2
