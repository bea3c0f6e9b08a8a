// b2_constmat.cuh -- constant multiplication of bit-sliced tower-field operands, two-pipe form.
//
// out = a * w for a twiddle w that is uniform over the 32 bit positions of the operand words (every limb
// of every element of a 128-byte line meets the same twiddle).  The w side of the product is
// a table of 0/1 WORDS ("coefficients"), so the AND of a data word with a twiddle bit is an integer
// multiply by 0 or 1: IMAD, which issues on the FMA pipe of the SM, while the XOR network stays on the
// ALU pipe (LOP3).  b2_upper.cuh's first form (ConstMul, AND masks, GF(4) leaves) keeps every operation
// on the ALU pipe, which round 1 measured as the wall (ALU pipe 80 % busy, issue slots 46 %).
//
// Circuit: Karatsuba over the tower levels H = 5, 4, 3 (data-side pre-additions only), and at H = 2 the
// GF(16) constant product as a direct 4 x 4 bit matrix: out_i = XOR_k a_k * c[i][k].  A deeper Karatsuba
// recursion would save AND terms (free here) at the price of more XORs (not free).
// Per 32-word butterfly of class 5 (SASS): 470 IMAD + 528 LOP3 + 108 LDS.128, against 876 LOP3 + 81 LDS.128 for ConstMul<5>.
//
// Replaces, like ConstMul, P::multiply in antt_butterfly (ntt/additive_ntt.cuh:10-14) for FanPaarTowerField<5>
// (finite_fields/binary_tower.cuh:35-50).
#pragma once
#include "b2_field.cuh"

namespace b2 {

struct alignas(16) Coef4 {
	uint32_t x, y, z, w;
};

// a * m for m in {0, 1}: kept an integer multiply (FMA pipe) on the device
B2_HD uint32_t sel01(uint32_t a, uint32_t m) {
#if defined(__CUDA_ARCH__)
	uint32_t r;
	asm("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(m));
	return r;
#else
	return a * m;
#endif
}

constexpr int kCoefLeafWords = 16;  // words per GF(16) leaf: row i = coefficients of output bit i

// NI = how many of the four terms of an output bit go through IMAD (coefficient word 0 / 1); the other 4 - NI terms are
// ANDs with a mask word (0 / ~0) fused into the XOR (LOP3).  Measured on B200 (tools/ubench/pipes.cu, profiles/r2c_pipes.jsonl),
// cycles per warp butterfly per SM sub-partition: ConstMul 1793, NI = 0: 1767, 1: 1684, 2: 1573, 3: 1487, 4: 1490.  With both
// pipes in use the sub-partition issues about 0.76 instructions per clock (register-file operand bandwidth), so NI = 3 and 4
// tie; 4 leaves the ALU pipe the most room for the other warps of the SM.
constexpr int kMatImadTerms = 4;

template <int H, int NI = kMatImadTerms>
struct MatMul {
	static constexpr int N = 1 << H, M = N / 2;
	using Sub = MatMul<H - 1, NI>;
	static constexpr int L = 3 * Sub::L;  // GF(16) leaves, ternary order, top level most significant (0 = low, 1 = high, 2 = low ^ high)
	static B2_HD void mul(const uint32_t* a, const Coef4* c, uint32_t* out) {
		uint32_t sa[M], z0[M], z2[M], zz[M], za[M];
#pragma unroll
		for (int i = 0; i < M; i++) sa[i] = a[i] ^ a[M + i];
		Sub::mul(a, c, z0);
		Sub::mul(a + M, c + 4 * Sub::L, z2);
		Sub::mul(sa, c + 8 * Sub::L, zz);
		Sliced<H - 1>::mul_alpha(z2, za);
#pragma unroll
		for (int i = 0; i < M; i++) {
			const uint32_t lo = z0[i] ^ z2[i];
			out[i] = lo;
			out[M + i] = zz[i] ^ lo ^ za[i];
		}
	}
};
template <int NI>
struct MatMul<2, NI> {
	static constexpr int L = 1;
	static B2_HD uint32_t term(uint32_t a, uint32_t q, int k) { return k >= 4 - NI ? sel01(a, q) : (a & q); }
	static B2_HD void mul(const uint32_t* a, const Coef4* c, uint32_t* out) {
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const Coef4 q = c[i];
			out[i] = term(a[0], q.x, 0) ^ term(a[1], q.y, 1) ^ term(a[2], q.z, 2) ^ term(a[3], q.w, 3);
		}
	}
};

// number of GF(16) leaves of a twiddle of class hw (2..5)
B2_HD int coef_leaves(int hw) { return hw == 5 ? 27 : hw == 4 ? 9 : hw == 3 ? 3 : hw == 2 ? 1 : 0; }

// GF(16) value at leaf g of the w-side Karatsuba tree of a class-hw twiddle (ternary digits of g, top level first)
B2_HD uint32_t coef_leaf_value(uint32_t w, int hw, int g) {
	uint32_t v = w;
	int div = hw == 5 ? 9 : hw == 4 ? 3 : 1;  // 3^(hw-3): weight of the top digit
	for (int half = 1 << (hw - 1); half >= 4; half >>= 1) {
		const int d = (g >= div) + (g >= 2 * div);  // no integer division on the device
		g -= d * div;
		div = div == 9 ? 3 : 1;
		const uint32_t lo = v & ((1u << half) - 1), hi = (v >> half) & ((1u << half) - 1);
		v = d == 0 ? lo : d == 1 ? hi : lo ^ hi;
	}
	return v & 15u;
}
// row i (coefficients of output bit i) of the 4 x 4 bit matrix of "multiply by lam" in GF(16)
template <int NI = kMatImadTerms>
B2_HD Coef4 coef_row(uint32_t lam, int i) {
	Coef4 q;
	q.x = (uint32_t)(Tower<2>::mul(lam, 1) >> i) & 1u;
	q.y = (uint32_t)(Tower<2>::mul(lam, 2) >> i) & 1u;
	q.z = (uint32_t)(Tower<2>::mul(lam, 4) >> i) & 1u;
	q.w = (uint32_t)(Tower<2>::mul(lam, 8) >> i) & 1u;
	if (NI < 4) q.x = 0u - q.x;  // mask terms: all-ones / all-zeros words
	if (NI < 3) q.y = 0u - q.y;
	if (NI < 2) q.z = 0u - q.z;
	if (NI < 1) q.w = 0u - q.w;
	return q;
}

// U ^= w * V ; V ^= U  (forward) or  V ^= U ; U ^= w * V  (inverse); w of class HW given by its coefficient rows
// NP = planes of the operands: 32 (a whole bundle) or 16 (one half of it -- a product of class <= 4 never mixes the halves)
template <int HW, bool INV, int NI = kMatImadTerms, int NP = 32>
B2_HD void butterfly_mat(uint32_t* U, uint32_t* V, const Coef4* c) {
	if (INV) {
#pragma unroll
		for (int i = 0; i < NP; i++) V[i] ^= U[i];
	}
	if constexpr (HW >= 2) {
		constexpr int C = 1 << HW;
#pragma unroll
		for (int ch = 0; ch < NP / C; ch++) {
			uint32_t p[C];
			MatMul<HW, NI>::mul(V + ch * C, c, p);
#pragma unroll
			for (int i = 0; i < C; i++) U[ch * C + i] ^= p[i];
		}
	}
	if (!INV) {
#pragma unroll
		for (int i = 0; i < NP; i++) V[i] ^= U[i];
	}
}
// two butterflies of one class (<= 4) on half bundles behind one dispatch
template <bool INV>
B2_HD void butterfly_half_dispatch2(int cls, uint32_t* U0, uint32_t* V0, const Coef4* c0, uint32_t* U1, uint32_t* V1, const Coef4* c1) {
	switch (cls) {
		case 0: butterfly_mat<0, INV, kMatImadTerms, 16>(U0, V0, c0), butterfly_mat<0, INV, kMatImadTerms, 16>(U1, V1, c1); break;
		case 2: butterfly_mat<2, INV, kMatImadTerms, 16>(U0, V0, c0), butterfly_mat<2, INV, kMatImadTerms, 16>(U1, V1, c1); break;
		case 3: butterfly_mat<3, INV, kMatImadTerms, 16>(U0, V0, c0), butterfly_mat<3, INV, kMatImadTerms, 16>(U1, V1, c1); break;
		default: butterfly_mat<4, INV, kMatImadTerms, 16>(U0, V0, c0), butterfly_mat<4, INV, kMatImadTerms, 16>(U1, V1, c1); break;
	}
}
// ---- the same product for a twiddle that is private to the THREAD (the per-lane layers of the tile pass) -------------------------
// A per-thread table of 108 coefficient rows would cost more to build than the product.  Instead the thread computes the 27
// GF(16) leaf values of its twiddle (packed arithmetic on the 32-bit scalar: about 70 integer operations) and fetches row i of
// leaf g from a STATIC table of the 16 possible 4 x 4 matrices, rows[i * 16 + lambda] (1 KB of shared memory, filled once per
// CTA): the 32 lanes of a warp read at most 16 different 16-byte chunks inside one 256-byte span per instruction.
struct LaneTwiddle {
	uint32_t y[9];       // y[3 * d3 + d4]: bits 0..3 / 16..19 = the value for d5 = 0 / 1 (d = 0: low half, 1: high half, 2: their sum)
	const Coef4* rows;   // [4][16]
	B2_HD explicit LaneTwiddle(uint32_t w, const Coef4* table) : rows(table) {
		const uint32_t x0 = w & 0x0f0f0f0fu, x1 = (w >> 4) & 0x0f0f0f0fu, x[3] = {x0, x1, x0 ^ x1};  // by nibble parity (d3)
#pragma unroll
		for (int d3 = 0; d3 < 3; d3++) {
			const uint32_t lo = x[d3] & 0x00ff00ffu, hi = (x[d3] >> 8) & 0x00ff00ffu;  // by byte parity (d4)
			y[3 * d3] = lo, y[3 * d3 + 1] = hi, y[3 * d3 + 2] = lo ^ hi;
		}
	}
	template <int G>
	B2_HD const Coef4* leaf() const {  // rows of leaf G = 9 d5 + 3 d4 + d3 (the order of coef_leaf_value)
		constexpr int d5 = G / 9, d4 = (G / 3) % 3, d3 = G % 3;
		const uint32_t v = y[3 * d3 + d4];
		const uint32_t lam = d5 == 0 ? (v & 15u) : d5 == 1 ? (v >> 16) : ((v ^ (v >> 16)) & 15u);
		return rows + lam;
	}
};
template <int H, int G0>
struct LaneMul {
	static constexpr int N = 1 << H, M = N / 2;
	static constexpr int L = MatMul<H - 1>::L;
	static B2_HD void mul(const uint32_t* a, const LaneTwiddle& tw, uint32_t* out) {
		uint32_t sa[M], z0[M], z2[M], zz[M], za[M];
#pragma unroll
		for (int i = 0; i < M; i++) sa[i] = a[i] ^ a[M + i];
		LaneMul<H - 1, G0>::mul(a, tw, z0);
		LaneMul<H - 1, G0 + L>::mul(a + M, tw, z2);
		LaneMul<H - 1, G0 + 2 * L>::mul(sa, tw, zz);
		Sliced<H - 1>::mul_alpha(z2, za);
#pragma unroll
		for (int i = 0; i < M; i++) {
			const uint32_t lo = z0[i] ^ z2[i];
			out[i] = lo;
			out[M + i] = zz[i] ^ lo ^ za[i];
		}
	}
};
template <int G0>
struct LaneMul<2, G0> {
	static B2_HD void mul(const uint32_t* a, const LaneTwiddle& tw, uint32_t* out) {
		const Coef4* c = tw.template leaf<G0>();
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const Coef4 q = c[16 * i];
			out[i] = sel01(a[0], q.x) ^ sel01(a[1], q.y) ^ sel01(a[2], q.z) ^ sel01(a[3], q.w);
		}
	}
};
// row table of LaneTwiddle: entry i * 16 + lam = row i of "multiply by lam" (all four terms as 0 / 1 coefficients)
B2_HD Coef4 lane_row_entry(int idx) { return coef_row<4>((uint32_t)idx & 15u, idx >> 4); }

// U ^= w * V ; V ^= U (forward) or V ^= U ; U ^= w * V (inverse) for a thread-private w in GF(2^32)
template <bool INV>
B2_HD void butterfly_lane(uint32_t* U, uint32_t* V, uint32_t w, const Coef4* rows) {
	if (INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
	{
		const LaneTwiddle tw(w, rows);
		uint32_t p[32];
		LaneMul<5, 0>::mul(V, tw, p);
#pragma unroll
		for (int i = 0; i < 32; i++) U[i] ^= p[i];
	}
	if (!INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
}

// the same for a twiddle of class HW (w < 2^(2^HW)): chunks of 2^HW planes, the leaves of the low digits only
template <int HW, bool INV>
B2_HD void butterfly_lane_cls(uint32_t* U, uint32_t* V, uint32_t w, const Coef4* rows) {
	if (INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
	if constexpr (HW >= 2) {
		constexpr int C = 1 << HW;
		const LaneTwiddle tw(w, rows);
#pragma unroll
		for (int ch = 0; ch < 32 / C; ch++) {
			uint32_t p[C];
			LaneMul<HW, 0>::mul(V + ch * C, tw, p);
#pragma unroll
			for (int i = 0; i < C; i++) U[ch * C + i] ^= p[i];
		}
	}
	if (!INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
}
template <bool INV>
B2_HD void butterfly_lane_dispatch(int cls, uint32_t* U, uint32_t* V, uint32_t w, const Coef4* rows) {
	switch (cls) {
		case 0: butterfly_lane_cls<0, INV>(U, V, w, rows); break;
		case 2: butterfly_lane_cls<2, INV>(U, V, w, rows); break;
		case 3: butterfly_lane_cls<3, INV>(U, V, w, rows); break;
		case 4: butterfly_lane_cls<4, INV>(U, V, w, rows); break;
		default: butterfly_lane<INV>(U, V, w, rows); break;
	}
}

template <bool INV>
B2_HD void butterfly_mat_dispatch(int cls, uint32_t* U, uint32_t* V, const Coef4* c) {
	switch (cls) {
		case 0: butterfly_mat<0, INV>(U, V, c); break;
		case 2: butterfly_mat<2, INV>(U, V, c); break;
		case 3: butterfly_mat<3, INV>(U, V, c); break;
		case 4: butterfly_mat<4, INV>(U, V, c); break;
		default: butterfly_mat<5, INV>(U, V, c); break;
	}
}

}  // namespace b2
