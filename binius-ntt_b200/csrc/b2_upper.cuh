// b2_upper.cuh -- "upper" passes of the additive NTT: butterfly layers whose stride is at least
// 32 lines (4 KB), fused two or four per pass, with WARP-UNIFORM twiddles.
//
// Index split of a pass that owns the line bits [q0, q0+k), q0 >= 5, k in {2, 4}:
//     line = hi << (q0+k) | x << q0 | lo_hi << 5 | t          t = lane of the warp (5 bits)
// A "super-line" is the 32 consecutive lines (4 KB) of one (hi, x, lo_hi); lane t owns line t of
// every super-line it meets, held bit-sliced in 32 registers (a bundle, b2_bundle.cuh).  All 32 x
// 32 words of a super-line meet the same twiddle in every layer of the pass (the twiddle of
// ntt/additive_ntt.cuh:59-77 depends only on the index bits above the butterfly bit), so:
//   * the product w * V is a CONSTANT multiplication: everything on the w side of the Karatsuba
//     circuit (bit masks, the pre-additions) is computed once per (coset row, hi) into a small
//     shared-memory table of "leaf masks" and fetched by warp-broadcast 16-byte loads, and the
//     butterfly costs only the V-side network (about 23 % fewer logic operations);
//   * lanes never exchange data: a warp runs two layers (a radix-4 block over 4 super-lines) on
//     registers alone, loading its lines straight from global memory.  With k = 4 the 4 warps of a
//     CTA swap super-lines once through shared memory (one __syncthreads per group); with k = 2
//     there is no exchange at all and every warp works on its own group.
//
// Replaces, for these layers, additive_ntt_kernel (ntt/additive_ntt.cuh:91-160).  The phase
// functions take the thread identity as arguments so tests/emul can drive them on the CPU.
#pragma once
#include "b2_constmat.cuh"
#include "b2_pass.cuh"

namespace b2 {

constexpr int kUpperWarps = 4;        // warps per CTA
constexpr int kLeafRows5 = 108;       // coefficient rows (Coef4) per twiddle of class 5: 27 GF(16) leaves x 4 output bits (b2_constmat.cuh)
constexpr int kLeafRows4 = 36;        // ... when no layer of the pass is above class 4: 9 leaves
constexpr int kLeafPitch = kLeafRows5 * 4;  // words per twiddle in the coefficient table
constexpr int kUpperMaxTw = 15;       // twiddles of a k = 4 group: 8 + 4 + 2 + 1
constexpr int kLeafTableWords = 63 * kLeafRows4 * 4;  // the larger of 15 x 108 (k = 4, class 5) and 63 x 36 (k = 6, classes <= 4) rows

struct alignas(16) Mask4 {
	uint32_t x, y, z, w;
};

// ---- constant multiplication, V side only ---------------------------------------------------------
// out = a * w for a lane-uniform w in GF(2^(2^H)) given by its 3^(H-1) leaf groups m[] (ternary order,
// top level most significant: 0 = low half, 1 = high half, 2 = low ^ high).  a: 2^H words.
template <int H>
struct ConstMul {
	static constexpr int N = 1 << H, M = N / 2;
	using Sub = ConstMul<H - 1>;
	static constexpr int G = 3 * Sub::G;
	static B2_HD void mul(const uint32_t* a, const Mask4* m, uint32_t* out) {
		uint32_t sa[M], z0[M], z2[M], zz[M], za[M];
#pragma unroll
		for (int i = 0; i < M; i++) sa[i] = a[i] ^ a[M + i];
		Sub::mul(a, m, z0);
		Sub::mul(a + M, m + Sub::G, z2);
		Sub::mul(sa, m + 2 * Sub::G, zz);
		Sliced<H - 1>::mul_alpha(z2, za);
#pragma unroll
		for (int i = 0; i < M; i++) {
			const uint32_t lo = z0[i] ^ z2[i];
			out[i] = lo;
			out[M + i] = zz[i] ^ lo ^ za[i];
		}
	}
};
// GF(4): (a0 + a1 X)(w0 + w1 X), X^2 = X + 1  ->  (a0 w0 + a1 w1) + (a0 w1 + a1 (w0 + w1)) X
template <>
struct ConstMul<1> {
	static constexpr int G = 1;
	static B2_HD void mul(const uint32_t* a, const Mask4* m, uint32_t* out) {
		const Mask4 q = m[0];
		out[0] = (a[0] & q.x) ^ (a[1] & q.y);
		out[1] = (a[0] & q.y) ^ (a[1] & q.z);
	}
};

// leaf group g (ternary digits, top level first) of a twiddle w of class HW: the two bits of the
// GF(4) value at that node of the w-side Karatsuba tree, as all-ones / all-zeros words
B2_HD Mask4 leaf_group(uint32_t w, int hw, int g) {
	uint32_t v = w;
	int div = hw == 5 ? 27 : hw == 4 ? 9 : hw == 3 ? 3 : 1;  // 3^(hw-2): weight of the top digit
	for (int half = 1 << (hw - 1); half > 1; half >>= 1) {
		const int d = (g >= div) + (g >= 2 * div);  // no integer division on the device
		g -= d * div;
		div = div == 27 ? 9 : div == 9 ? 3 : 1;
		const uint32_t lo = v & ((1u << half) - 1), hi = (v >> half) & ((1u << half) - 1);
		v = d == 0 ? lo : d == 1 ? hi : lo ^ hi;
	}
	Mask4 q;
	q.x = 0u - (v & 1u), q.y = 0u - ((v >> 1) & 1u), q.z = q.x ^ q.y, q.w = 0;
	return q;
}
B2_HD int leaf_groups(int hw) { return hw == 5 ? 81 : hw == 4 ? 27 : hw == 3 ? 9 : hw == 2 ? 3 : 0; }

// U ^= w * V ; V ^= U  (forward) or  V ^= U ; U ^= w * V  (inverse); w of class HW given by its leaf masks
template <int HW, bool INV>
B2_HD void butterfly_const(uint32_t* U, uint32_t* V, const Mask4* m) {
	if (INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
	if constexpr (HW > 0) {
		constexpr int C = 1 << HW;
#pragma unroll
		for (int c = 0; c < 32 / C; c++) {
			uint32_t p[C];
			ConstMul<HW>::mul(V + c * C, m, p);
#pragma unroll
			for (int i = 0; i < C; i++) U[c * C + i] ^= p[i];
		}
	}
	if (!INV) {
#pragma unroll
		for (int i = 0; i < 32; i++) V[i] ^= U[i];
	}
}
template <bool INV>
B2_HD void butterfly_const_dispatch(int cls, uint32_t* U, uint32_t* V, const Mask4* m) {
	switch (cls) {
		case 0: butterfly_const<0, INV>(U, V, m); break;
		case 2: butterfly_const<2, INV>(U, V, m); break;
		case 3: butterfly_const<3, INV>(U, V, m); break;
		case 4: butterfly_const<4, INV>(U, V, m); break;
		default: butterfly_const<5, INV>(U, V, m); break;
	}
}

// ---- pass description ---------------------------------------------------------------------------
struct UpperParams {
	const uint32_t* src;   // row r reads words [(r >> src_row_shift) * src_row_words, ...)
	uint32_t* dst;         // row r at r * dst_row_words
	uint64_t src_row_words, dst_row_words;
	int src_row_shift;     // log_rate when the first pass fans one column out to its coset rows
	uint32_t coset_mask;   // row & coset_mask = coset of the row
	int coset0;            // coset of row 0 (single-coset transforms)
	const uint32_t* consts;  // [32][kConstPitch] normalised subspace constants (additive_ntt.cuh:273-309)
	int log_h;             // of the whole transform (twiddle index shift)
	int nb;                // line bits per row
	int q0, k;             // this pass owns line bits [q0, q0+k), q0 >= 5, k = 2, 4 or 6
	int ls;                // log2 of the lines per super-line: 5 (k = 2, 4), 3 (k = 6)
	int lpitch;            // coefficient rows per twiddle in the table: 108, or 36 when no layer of the pass has class 5
	int stage0;            // butterfly stage of local bit 0
	int src_sliced, dst_sliced;  // lines are bundles already / stay bundles (no 32x32 transpose)
	int inverse;
	uint64_t n_groups;     // rows << (nb - k - 5)
	uint8_t cls[8];        // twiddle class of local bit i (0, 2..5)
};

constexpr int kUSlotWords = 36;  // padded line slot in the exchange buffer (as b2_pass.cuh)
B2_HD size_t upper_smem_words(int k) {
	(void)k;
	return (size_t)kLeafTableWords + 64 + 1024 + 256 + (size_t)512 * kUSlotWords;
}
struct UpperSmem {
	Coef4* leaves;     // [2^k - 1][lpitch] coefficient rows (b2_constmat.cuh)
	uint32_t* tw;      // [64] twiddle words
	uint32_t* consts;  // [32][32] copy of the subspace constants
	uint32_t* slots;   // [512 lines][kUSlotWords]: super-line x, line l at (x << ls) + l; k = 2: 4 super-lines per warp
	const Coef4* rows16;  // [4][16] static: row i of the GF(16) matrix of lam at i * 16 + lam (lane_row_entry), filled once per CTA
};
B2_HD UpperSmem upper_smem_view(uint32_t* base) {
	UpperSmem sm;
	sm.leaves = reinterpret_cast<Coef4*>(base);
	sm.tw = base + kLeafTableWords;
	sm.consts = sm.tw + 64;
	sm.rows16 = reinterpret_cast<const Coef4*>(sm.consts + 1024);
	sm.slots = sm.consts + 1024 + 256;
	return sm;
}

struct UGroup {
	uint32_t row;
	uint64_t hi, lo_hi;
};
B2_HD UGroup upper_group(const UpperParams& P, uint64_t g) {
	UGroup c;
	const int gb = P.nb - P.k - P.ls;  // group bits per row
	c.row = (uint32_t)(g >> gb);
	const uint64_t rem = g & ((uint64_t(1) << gb) - 1);
	c.lo_hi = rem & ((uint64_t(1) << (P.q0 - P.ls)) - 1);
	c.hi = rem >> (P.q0 - P.ls);
	return c;
}
B2_HD uint64_t upper_line(const UpperParams& P, const UGroup& c, int x, int lane) {
	return (c.hi << (P.q0 + P.k)) | ((uint64_t)x << P.q0) | (c.lo_hi << P.ls) | (uint64_t)lane;
}
// table index of the twiddle of layer i for the super-line x (depends on x >> (i+1))
B2_HD int upper_tw_index(int k, int i, int x) { return (1 << (k - 1 - i)) - 1 + (x >> (i + 1)); }

// once per CTA, before the first upper_leaves
B2_HD void upper_rows16_fill(const UpperSmem& sm, int tid, int nthreads) {
	Coef4* t = const_cast<Coef4*>(sm.rows16);
	for (int i = tid; i < 64; i += nthreads) t[i] = lane_row_entry(i);
}

// twiddle words of (row, hi): threads tid < 2^k - 1; consts = the [32][32] constants (a shared-memory copy on the GPU)
B2_HD void upper_twiddles(const UpperParams& P, UpperSmem& sm, const UGroup& c, int tid, const uint32_t* consts) {
	if (tid >= (1 << P.k) - 1) return;
	int lg = 0;
	while ((tid + 1) >> (lg + 1)) lg++;
	const int i = P.k - 1 - lg, j = tid + 1 - (1 << lg);
	const int stage = P.stage0 + i;
	const uint64_t ind = ((uint64_t)(P.coset0 + (c.row & P.coset_mask)) << (P.log_h - 1 - stage)) | (c.hi << (P.k - 1 - i)) | (uint64_t)j;
	uint32_t w = 0;
	const uint32_t* t = consts + stage * 32;
	for (int b = 0; (ind >> b) != 0; b++)
		if ((ind >> b) & 1) w ^= t[b];
	sm.tw[tid] = w;
}
// coefficient rows of every twiddle: item = twiddle * lpitch + leaf * 4 + output bit, strided over nthreads
B2_HD void upper_leaves(const UpperParams& P, UpperSmem& sm, int tid, int nthreads) {
	const int ntw = (1 << P.k) - 1, lp = P.lpitch;
	for (int item = tid; item < ntw * lp; item += nthreads) {
		const int tw = lp == kLeafRows5 ? (item * 607) >> 16 : (item * 1821) >> 16;  // item / 108 (item < 1620), item / 36 (item < 2268)
		const int r = item - tw * lp;
		int lg = 0;
		while ((tw + 1) >> (lg + 1)) lg++;
		const int hw = P.cls[P.k - 1 - lg];
		// a row of a twiddle's table is a row of one of the 16 GF(16) matrices: copied from the static table
		if ((r >> 2) < coef_leaves(hw)) sm.leaves[tw * lp + r] = sm.rows16[(r & 3) * 16 + coef_leaf_value(sm.tw[tw], hw, r >> 2)];
	}
}

// ---- one line (32 words) <-> registers -----------------------------------------------------------
B2_HD void line_load(const uint32_t* p, uint32_t* w) {
#if defined(__CUDA_ARCH__)
#pragma unroll
	for (int g = 0; g < 8; g++) {
		const uint4 v = reinterpret_cast<const uint4*>(p)[g];
		w[4 * g] = v.x, w[4 * g + 1] = v.y, w[4 * g + 2] = v.z, w[4 * g + 3] = v.w;
	}
#else
	for (int j = 0; j < 32; j++) w[j] = p[j];
#endif
}
B2_HD void line_store(uint32_t* p, const uint32_t* w) {
#if defined(__CUDA_ARCH__)
#pragma unroll
	for (int g = 0; g < 8; g++) reinterpret_cast<uint4*>(p)[g] = make_uint4(w[4 * g], w[4 * g + 1], w[4 * g + 2], w[4 * g + 3]);
#else
	for (int j = 0; j < 32; j++) p[j] = w[j];
#endif
}

// Two butterfly layers (local bits b1 and b1 - 1) of one lane over the 4 super-lines x(a) = xbase | a << xshift,
// a = a1 a0, held in the slots slot_off + x(a).  The lane's lines are private to it, so the two layers need
// no synchronisation.  The loops are deliberately NOT unrolled: one copy of each butterfly circuit serves
// all 4 butterflies, which keeps the executed code inside the instruction cache.
//   first: the slots hold lines as copied from the pass input (transpose compact lines on first touch)
//   last : the slots are copied to the pass output afterwards (transpose back for a compact output)
//   NT   : "no table": every layer is of class 5 and the butterfly takes the twiddle WORD (sm.tw) through the static GF(16)
//          rows (butterfly_lane), so the caller does not build sm.leaves at all
// A layer pair whose twiddles are both of class <= 2 (GF(16) scalars or 0 / 1: the top stages of a large transform): the
// circuits are a few dozen instructions, so the lane's 4 lines stay in registers across both layers and cross the shared
// memory once instead of twice.  (With class 4 / 5 circuits the same form spills: profiles/README.md, upper_quad.)
template <bool INV>
B2_HD void small_butterfly(int cls, uint32_t* U, uint32_t* V, const Coef4* c) {
	if (cls == 2)
		butterfly_mat<2, INV>(U, V, c);
	else
		butterfly_mat<0, INV>(U, V, c);
}
template <bool INV>
B2_HD void upper_quad_small(const UpperParams& P, const UpperSmem& sm, int lane, int xbase, int xshift, int slot_off, int b1, bool first, bool last) {
	const int k = P.k, ls = P.ls, l = lane & ((1 << ls) - 1);
	uint32_t* s0 = sm.slots + ((size_t)((slot_off + xbase) << ls) + l) * kUSlotWords;
	const size_t sa = ((size_t)1 << (xshift + ls)) * kUSlotWords;  // slot stride of a
	uint32_t X0[32], X1[32], X2[32], X3[32];
	line_load(s0, X0);
	line_load(s0 + sa, X1);
	line_load(s0 + 2 * sa, X2);
	line_load(s0 + 3 * sa, X3);
	if (first && !P.src_sliced) {
		transpose32(X0);
		transpose32(X1);
		transpose32(X2);
		transpose32(X3);
	}
	const int ct = P.cls[b1], cl = P.cls[b1 - 1];
	const Coef4* tt = sm.leaves + P.lpitch * upper_tw_index(k, b1, xbase);  // one twiddle for both pairs of the top layer
	const Coef4* t0 = sm.leaves + P.lpitch * upper_tw_index(k, b1 - 1, xbase);
	const Coef4* t1 = sm.leaves + P.lpitch * upper_tw_index(k, b1 - 1, xbase | (2 << xshift));
	if (!INV) {
		small_butterfly<false>(ct, X0, X2, tt);
		small_butterfly<false>(ct, X1, X3, tt);
		small_butterfly<false>(cl, X0, X1, t0);
		small_butterfly<false>(cl, X2, X3, t1);
	} else {
		small_butterfly<true>(cl, X0, X1, t0);
		small_butterfly<true>(cl, X2, X3, t1);
		small_butterfly<true>(ct, X0, X2, tt);
		small_butterfly<true>(ct, X1, X3, tt);
	}
	if (last && !P.dst_sliced) {
		transpose32(X0);
		transpose32(X1);
		transpose32(X2);
		transpose32(X3);
	}
	line_store(s0, X0);
	line_store(s0 + sa, X1);
	line_store(s0 + 2 * sa, X2);
	line_store(s0 + 3 * sa, X3);
}

// A layer pair of class <= 4 on bundles that need no transpose in this call: the product never mixes planes 0..15 with
// planes 16..31, so the pair runs on the lower halves of the lane's 4 lines, then on the upper halves -- 64 data registers at
// a time, and every line crosses the shared memory once per layer PAIR.
B2_HD void half_load(const uint32_t* p, uint32_t* w) {
#if defined(__CUDA_ARCH__)
#pragma unroll
	for (int g = 0; g < 4; g++) {
		const uint4 v = reinterpret_cast<const uint4*>(p)[g];
		w[4 * g] = v.x, w[4 * g + 1] = v.y, w[4 * g + 2] = v.z, w[4 * g + 3] = v.w;
	}
#else
	for (int j = 0; j < 16; j++) w[j] = p[j];
#endif
}
B2_HD void half_store(uint32_t* p, const uint32_t* w) {
#if defined(__CUDA_ARCH__)
#pragma unroll
	for (int g = 0; g < 4; g++) reinterpret_cast<uint4*>(p)[g] = make_uint4(w[4 * g], w[4 * g + 1], w[4 * g + 2], w[4 * g + 3]);
#else
	for (int j = 0; j < 16; j++) p[j] = w[j];
#endif
}
template <bool INV, int CLS = 0>
B2_HD void upper_quad_halves(const UpperParams& P, const UpperSmem& sm, int lane, int xbase, int xshift, int slot_off, int b1) {
	const int k = P.k, ls = P.ls, l = lane & ((1 << ls) - 1);
	uint32_t* s0 = sm.slots + ((size_t)((slot_off + xbase) << ls) + l) * kUSlotWords;
	const size_t sa = ((size_t)1 << (xshift + ls)) * kUSlotWords;  // slot stride of a
	const int ct = P.cls[b1], cl = P.cls[b1 - 1];
	const Coef4* tt = sm.leaves + P.lpitch * upper_tw_index(k, b1, xbase);  // one twiddle for both pairs of the top layer
	const Coef4* t0 = sm.leaves + P.lpitch * upper_tw_index(k, b1 - 1, xbase);
	const Coef4* t1 = sm.leaves + P.lpitch * upper_tw_index(k, b1 - 1, xbase | (2 << xshift));
#pragma unroll 1
	for (int h = 0; h < 2; h++) {
		uint32_t* s = s0 + 16 * h;
		uint32_t X0[16], X1[16], X2[16], X3[16];
		half_load(s, X0);
		half_load(s + sa, X1);
		half_load(s + 2 * sa, X2);
		half_load(s + 3 * sa, X3);
		if constexpr (CLS == 4) {  // every layer of the pass has class 4: no dispatch, one circuit
			constexpr int NI = kMatImadTerms;
			if (!INV) {
				butterfly_mat<4, false, NI, 16>(X0, X2, tt), butterfly_mat<4, false, NI, 16>(X1, X3, tt);
				butterfly_mat<4, false, NI, 16>(X0, X1, t0), butterfly_mat<4, false, NI, 16>(X2, X3, t1);
			} else {
				butterfly_mat<4, true, NI, 16>(X0, X1, t0), butterfly_mat<4, true, NI, 16>(X2, X3, t1);
				butterfly_mat<4, true, NI, 16>(X0, X2, tt), butterfly_mat<4, true, NI, 16>(X1, X3, tt);
			}
		} else if (!INV) {
			butterfly_half_dispatch2<false>(ct, X0, X2, tt, X1, X3, tt);
			butterfly_half_dispatch2<false>(cl, X0, X1, t0, X2, X3, t1);
		} else {
			butterfly_half_dispatch2<true>(cl, X0, X1, t0, X2, X3, t1);
			butterfly_half_dispatch2<true>(ct, X0, X2, tt, X1, X3, tt);
		}
		half_store(s, X0);
		half_store(s + sa, X1);
		half_store(s + 2 * sa, X2);
		half_store(s + 3 * sa, X3);
	}
}

// CLS: 0 = the classes of the layers are read from P.cls; 4 / 5 = every layer of the pass has that class (lean instantiations)
template <bool INV, int NT = 0, int CLS = 0>
B2_HD void upper_pair(const UpperParams& P, const UpperSmem& sm, int lane, int xbase, int xshift, int slot_off, int b1, bool first, bool last) {
	const int k = P.k, ls = P.ls, l = lane & ((1 << ls) - 1);  // the lane's line inside every super-line it meets
	if (NT == 0 && CLS == 0 && P.cls[b1] <= 2 && P.cls[b1 - 1] <= 2) {
		upper_quad_small<INV>(P, sm, lane, xbase, xshift, slot_off, b1, first, last);
		return;
	}
	if (NT == 0 && CLS != 5 && (CLS == 4 || (P.cls[b1] <= 4 && P.cls[b1 - 1] <= 4)) && !(first && !P.src_sliced) && !(last && !P.dst_sliced)) {
		upper_quad_halves<INV, CLS>(P, sm, lane, xbase, xshift, slot_off, b1);
		return;
	}
#pragma unroll 1
	for (int step = 0; step < 2; step++) {
		const bool top = INV ? step == 1 : step == 0;  // which of the two layers
		const int bit = top ? b1 : b1 - 1;
		const int cls = P.cls[bit];
#pragma unroll 1
		for (int j = 0; j < 2; j++) {
			const int au = top ? j : 2 * j, av = top ? j + 2 : 2 * j + 1;
			const int xu = xbase | (au << xshift), xv = xbase | (av << xshift);
			uint32_t* su = sm.slots + ((size_t)((slot_off + xu) << ls) + l) * kUSlotWords;
			uint32_t* sv = sm.slots + ((size_t)((slot_off + xv) << ls) + l) * kUSlotWords;
			uint32_t U[32], V[32];
			line_load(su, U);
			line_load(sv, V);
			if (first && step == 0 && !P.src_sliced) {
				transpose32(U);
				transpose32(V);
			}
			if (NT == 2)  // no table, any class (the pipelined upper pass keeps no coefficient rows)
				butterfly_lane_dispatch<INV>(cls, U, V, sm.tw[upper_tw_index(k, bit, xu)], sm.rows16);
			else if (NT)
				butterfly_lane<INV>(U, V, sm.tw[upper_tw_index(k, bit, xu)], sm.rows16);
			else if constexpr (CLS == 5)
				butterfly_mat<5, INV>(U, V, sm.leaves + P.lpitch * upper_tw_index(k, bit, xu));
			else if constexpr (CLS == 4)
				butterfly_mat<4, INV>(U, V, sm.leaves + P.lpitch * upper_tw_index(k, bit, xu));
			else
				butterfly_mat_dispatch<INV>(cls, U, V, sm.leaves + P.lpitch * upper_tw_index(k, bit, xu));
			if (last && step == 1 && !P.dst_sliced) {
				transpose32(U);
				transpose32(V);
			}
			line_store(su, U);
			line_store(sv, V);
		}
	}
}

// ---- coalesced copies of super-lines: global <-> slots, 16-byte chunks ---------------------------------
// One warp copies the super-lines x = x0 + i1 * s1 + i2 * s2 (i1 < n1, i2 < n2) of the group between global memory and
// the slots (slot_off + x).  A super-line is 2^ls lines = 2^(ls+3) 16-byte chunks, contiguous in global memory; the
// warp moves 32 consecutive chunks (512 bytes) per step.
template <bool STORE>
B2_HD void upper_copy(const UpperParams& P, const UpperSmem& sm, const UGroup& c, int slot_off, int x0, int s1, int n1, int s2, int n2, int lane) {
	const int ls = P.ls, steps = 1 << (ls - 2);  // steps per super-line
	const int piece = lane & 7, l0 = lane >> 3;
	const uint32_t* srow = P.src + (uint64_t)(c.row >> P.src_row_shift) * P.src_row_words;
	uint32_t* drow = P.dst + (uint64_t)c.row * P.dst_row_words;
#pragma unroll 1
	for (int i2 = 0; i2 < n2; i2++) {
#pragma unroll 1
		for (int i1 = 0; i1 < n1; i1++) {
			const int x = x0 + i1 * s1 + i2 * s2;
			const uint64_t g = upper_line(P, c, x, l0) * 32 + piece * 4;
			uint32_t* s = sm.slots + ((size_t)((slot_off + x) << ls) + l0) * kUSlotWords + piece * 4;
#pragma unroll 2
			for (int st = 0; st < steps; st++) {
				if (STORE)
					copy16_s2g(drow + g + (size_t)st * 128, s + (size_t)st * 4 * kUSlotWords);
				else
					copy16_g2s(s + (size_t)st * 4 * kUSlotWords, srow + g + (size_t)st * 128);
			}
		}
	}
}

}  // namespace b2
