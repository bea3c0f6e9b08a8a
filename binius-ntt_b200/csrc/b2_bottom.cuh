// b2_bottom.cuh -- the lowest pass of the additive NTT: a tile of 512 consecutive lines (64 KB) per CTA.
//
//   line bits 5..8 of the tile ("phase A") -- strides of 32..256 lines: exactly an upper pass with q0 = 5
//       (b2_upper.cuh): warp-uniform twiddles, leaf-mask constant multiply, lane-private lines, one
//       __syncthreads between its two layer pairs;
//   line bits bb..4 ("phase B", bb = batch_bits) -- pairs of lines inside one super-line, and the
//       butterfly layers INSIDE a line: per-lane twiddles, bit-sliced Karatsuba product.  After phase A
//       warp w owns the super-lines 4w..4w+3, so phase B needs only __syncwarp between its layers.
//
// Replaces additive_ntt_kernel (ntt/additive_ntt.cuh:91-160) for the stages below 2^(9-bb) lines and the
// 32x32 bit transposes between the compact image and the bundle form.
#pragma once
#include "b2_upper.cuh"

namespace b2 {

struct BottomParams {
	UpperParams up;          // phase A: k = 4, q0 = 5 (n_groups = tiles)
	const uint32_t* lane_tw; // [5][32] per-lane twiddle words of the in-line layers
	int n_inline;            // butterfly layers inside a line
	int lane_shift;          // in-line layer s pairs lanes differing in bit s + lane_shift
	int batch_bits;          // low line bits that are not butterfly bits
	uint8_t cls[32];         // twiddle class per stage
	int swap_inline;         // 1: the in-line layers run in the exchanged form (in_line_stage_swapped); decided on the host (b2_plan.h)
};

struct BottomSmem {
	UpperSmem us;
	uint32_t* TL;   // line-layer twiddle, tile-independent part: heap order, bit b owns [2^(8-b), 2^(9-b)) - index (1 << (8-b)) + (l >> (b+1))
	uint32_t* TH;   // [8] line-layer twiddle, per tile and coset: by line bit
	uint32_t* IH;   // [8] in-line twiddle, per tile and coset: by stage
	uint32_t* IL;   // [n_inline][512] in-line twiddle by local line
	uint32_t* LT;   // [5][32] lane_tw copy
	Coef4* rows;    // [4][16] the GF(16) coefficient rows of butterfly_lane (b2_constmat.cuh)
	uint32_t* CW;   // [5][16] position masks of the GF(16) correction of butterfly_in_line_split; [80 + s] = 1: stage s uses it;
	                // [96 + 16 s]: masks of the exchanged form (in_line_stage_swapped), when B.swap_inline
};
// One CTA runs kBottomTeams teams of 4 warps, each on its own tile.  Warps w and w + 4 share an SM sub-partition and
// meet at the same CTA barriers, so they execute the same stretch of straight-line butterfly code at the same time:
// the instruction cache serves both (two independent CTAs per SM ran out of phase and stalled on instruction fetch).
// layout: [consts 1024][TL 512][IL 5*512][LT 160][rows 256][CW 160]  then per team  [coefficient rows 15*108*4][tw 16][TH 8][IH 8][slots 16*32*36]
constexpr int kBottomTeams = 2;
B2_HD size_t bottom_team_words() { return (size_t)kUpperMaxTw * kLeafPitch + 16 + 8 + 8 + (size_t)16 * 32 * kUSlotWords; }
B2_HD size_t bottom_smem_words() { return 1024 + 512 + 5 * 512 + 160 + 256 + 160 + kBottomTeams * bottom_team_words(); }
B2_HD BottomSmem bottom_smem_view(uint32_t* base, int team = 0) {
	BottomSmem b;
	b.us.consts = base;
	b.TL = base + 1024;
	b.IL = b.TL + 512;
	b.LT = b.IL + 5 * 512;
	b.rows = reinterpret_cast<Coef4*>(b.LT + 160);  // 16-byte aligned: 1024 + 512 + 2560 + 160 words before it
	b.us.rows16 = b.rows;                           // the same table serves the coefficient-table build of phase A
	b.CW = b.LT + 160 + 256;
	uint32_t* t = b.CW + 160 + (size_t)team * bottom_team_words();
	b.us.leaves = reinterpret_cast<Coef4*>(t);
	b.us.tw = t + kUpperMaxTw * kLeafPitch;
	b.TH = b.us.tw + 16;
	b.IH = b.TH + 8;
	b.us.slots = b.IH + 8;
	return b;
}

B2_HD uint32_t xsel(const uint32_t* table, uint64_t bits) {
	uint32_t w = 0;
	for (int b = 0; bits; b++, bits >>= 1)
		if (bits & 1u) w ^= table[b];
	return w;
}
B2_HD int bottom_stage(const BottomParams& B, int bit) { return B.n_inline + bit - B.batch_bits; }

// ---- in-line layer with the twiddle split into a thread-uniform word and a small per-position correction -----------------------
// butterfly_in_line (b2_bundle.cuh) multiplies by 32 arbitrary twiddle WORDS (the twiddle differs from position to position
// inside a line), a two-sided product that lives on the ALU pipe alone (1087 LOP3).  But the per-position part is small:
//     W(pos) = wa ^ D(pos),   D = lane_tw ^ (delta on the positions that hold the second line's lanes),
// wa = the twiddle of the thread's first line, delta = the (kernel-wide) twiddle difference of the two lines; lane_tw collects
// constants C[s][j] of the lowest index bits.  For the reference basis these are all below 16 -- elements of GF(16), which act
// on every 4-plane group of the sliced operand separately -- so
//     X * W = X * wa  (the thread-uniform two-pipe product, LaneMul)  ^  XOR_k X[4 g + k] & cw[i][k]  (16 LOP3 per group),
// cw[i][k] = the positions whose D has coefficient (i, k) of its GF(16) matrix set.  The stage falls back to butterfly_in_line
// when some D(pos) >= 16 or the stage's twiddles are not of class 5.
B2_HD void in_line_split_tables(const BottomParams& B, BottomSmem& bs, int s, const uint32_t* consts) {
	const int p = s + B.lane_shift;
	const uint32_t delta = consts[s * 32 + B.n_inline - 1 - s], hi = ~lane_mask_lo(p);
	bool small = B.cls[s] == 5 && delta < 16u;
	uint32_t D[4];
	for (int r = 0; r < 32; r++) {
		const uint32_t d = B.lane_tw[s * 32 + r] ^ (bitmask(delta, r) & hi);
		if (r < 4)
			D[r] = d;
		else if (d)
			small = false;
	}
	for (int i = 0; i < 4; i++)
		for (int k = 0; k < 4; k++) {
			uint32_t m = 0;
			for (int r = 0; r < 4; r++)
				if ((Tower<2>::mul(1u << r, 1u << k) >> i) & 1u) m ^= D[r];
			bs.CW[s * 16 + 4 * i + k] = m;
		}
	bs.CW[80 + s] = small ? 1u : 0u;
}

template <bool INVERSE>
B2_HD void butterfly_in_line_split(uint32_t* A, uint32_t* B, int p, const uint32_t* cw, uint32_t wa, const Coef4* rows) {
	const int sh = 1 << p;
	const uint32_t mu = lane_mask_lo(p);
	if (INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			A[i] ^= (A[i] & mu) << sh;
			B[i] ^= (B[i] & mu) << sh;
		}
	}
	{
		uint32_t X[32], Pr[32];
#pragma unroll
		for (int i = 0; i < 32; i++) X[i] = ((A[i] >> sh) & mu) | (B[i] & ~mu);
		const LaneTwiddle tw(wa, rows);
		LaneMul<5, 0>::mul(X, tw, Pr);
		uint32_t c[16];
#pragma unroll
		for (int i = 0; i < 16; i++) c[i] = cw[i];
#pragma unroll
		for (int g = 0; g < 8; g++) {
#pragma unroll
			for (int i = 0; i < 4; i++) {
				const uint32_t q = Pr[4 * g + i] ^ (X[4 * g] & c[4 * i]) ^ (X[4 * g + 1] & c[4 * i + 1]) ^ (X[4 * g + 2] & c[4 * i + 2]) ^ (X[4 * g + 3] & c[4 * i + 3]);
				A[4 * g + i] ^= q & mu;
				B[4 * g + i] ^= (q >> sh) & mu;
			}
		}
	}
	if (!INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			A[i] ^= (A[i] & mu) << sh;
			B[i] ^= (B[i] & mu) << sh;
		}
	}
}

// ---- the in-line layers of the compact GF(2^128) image in EXCHANGED form -------------------------------------------------------
// butterfly_in_line_split still pays for the layout: the partner of a position sits 2^p positions higher in the SAME word, so every
// layer shifts the upper halves of one bundle down, scatters the product back and shifts the sum up again (about 290 ALU
// instructions per layer).  Instead, before layer p the two bundles exchange halves (an involution: 1 PRMT per word for the
// byte-sized layers):  A' = A[bit p clear] | B[bit p clear] << 2^p,  B' = A[bit p set] >> 2^p | B[bit p set].
// A' then holds every u position of both lines and B' every v position, in the same places: the layer is the plain butterfly
// A' ^= W * B' ; B' ^= A'.  The exchanges are not undone between the layers -- layer p - 1 exchanges (A', B') on its own bit --
// so after the three layers position bit 4 of a bundle tells the LINE, bits 3 and 2 the element bits e2 and e1, and the bundle
// (A or B) the element bit e0.  That permutation moves whole elements (16-byte chunks), so it costs nothing: the transposed words
// are stored chunk by chunk to where they belong (swapped_store), and the inverse transform loads them the same way.
// The twiddle of position pos at stage s (p = s + 2) is wa ^ D_s(pos), D_s = XOR_j (bit p + j of pos) * C[s][j], j = 0 .. 2 - s:
// the element bits above the stage and the line bit, exactly the index bits the thread-uniform wa (line A, element 0) leaves out.
B2_HD void exchange_halves(uint32_t* A, uint32_t* B, int p) {
#if defined(__CUDA_ARCH__)
	if (p == 4) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint32_t a = A[i], b = B[i];
			A[i] = __byte_perm(a, b, 0x5410), B[i] = __byte_perm(a, b, 0x7632);
		}
		return;
	}
	if (p == 3) {
#pragma unroll
		for (int i = 0; i < 32; i++) {
			const uint32_t a = A[i], b = B[i];
			A[i] = __byte_perm(a, b, 0x6240), B[i] = __byte_perm(a, b, 0x7351);
		}
		return;
	}
#endif
	const int sh = 1 << p;
	const uint32_t mu = lane_mask_lo(p);
#pragma unroll
	for (int i = 0; i < 32; i++) {
		const uint32_t a = A[i], b = B[i];
		A[i] = (a & mu) | ((b & mu) << sh), B[i] = ((a >> sh) & mu) | (b & ~mu);
	}
}
template <bool INVERSE>
B2_HD void in_line_stage_swapped(uint32_t* A, uint32_t* B, int p, const uint32_t* cw, uint32_t wa, const Coef4* rows) {
	if (!INVERSE) exchange_halves(A, B, p);
	if (INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) B[i] ^= A[i];
	}
	{
		uint32_t Pr[32], c[16];
		const LaneTwiddle tw(wa, rows);
		LaneMul<5, 0>::mul(B, tw, Pr);
#pragma unroll
		for (int i = 0; i < 16; i++) c[i] = cw[i];
#pragma unroll
		for (int g = 0; g < 8; g++) {
#pragma unroll
			for (int i = 0; i < 4; i++)
				A[4 * g + i] ^= Pr[4 * g + i] ^ (B[4 * g] & c[4 * i]) ^ (B[4 * g + 1] & c[4 * i + 1]) ^ (B[4 * g + 2] & c[4 * i + 2]) ^ (B[4 * g + 3] & c[4 * i + 3]);
		}
	}
	if (!INVERSE) {
#pragma unroll
		for (int i = 0; i < 32; i++) B[i] ^= A[i];
	}
	if (INVERSE) exchange_halves(A, B, p);
}
// chunk c = (line, e2, e1) of bundle A / B <-> element (e2, e1, 0 / 1) of line lu / lv
B2_HD void swapped_store(uint32_t* su, uint32_t* sv, const uint32_t* A, const uint32_t* B) {
#pragma unroll
	for (int c = 0; c < 8; c++) {
		uint32_t* d = (c & 4 ? sv : su) + 8 * (c & 3);
#if defined(__CUDA_ARCH__)
		reinterpret_cast<uint4*>(d)[0] = make_uint4(A[4 * c], A[4 * c + 1], A[4 * c + 2], A[4 * c + 3]);
		reinterpret_cast<uint4*>(d)[1] = make_uint4(B[4 * c], B[4 * c + 1], B[4 * c + 2], B[4 * c + 3]);
#else
		for (int j = 0; j < 4; j++) d[j] = A[4 * c + j], d[4 + j] = B[4 * c + j];
#endif
	}
}
B2_HD void swapped_load(const uint32_t* su, const uint32_t* sv, uint32_t* A, uint32_t* B) {
#pragma unroll
	for (int c = 0; c < 8; c++) {
		const uint32_t* d = (c & 4 ? sv : su) + 8 * (c & 3);
#if defined(__CUDA_ARCH__)
		const uint4 x = reinterpret_cast<const uint4*>(d)[0], y = reinterpret_cast<const uint4*>(d)[1];
		A[4 * c] = x.x, A[4 * c + 1] = x.y, A[4 * c + 2] = x.z, A[4 * c + 3] = x.w;
		B[4 * c] = y.x, B[4 * c + 1] = y.y, B[4 * c + 2] = y.z, B[4 * c + 3] = y.w;
#else
		for (int j = 0; j < 4; j++) A[4 * c + j] = d[j], B[4 * c + j] = d[4 + j];
#endif
	}
}
// can the exchanged form be used: compact GF(2^128) image, class-5 stages, GF(16) low constants (host: b2_plan.h)
B2_HD bool in_line_swap_ok(int n_inline, int lane_shift, int batch_bits, const uint8_t* cls, const uint32_t* consts, int pitch) {
	bool ok = n_inline == 3 && lane_shift == 2 && batch_bits == 0;
	for (int s = 0; s < 3 && ok; s++) {
		if (cls[s] != 5) ok = false;
		for (int j = 0; j <= 2 - s; j++)
			if (consts[s * pitch + j] >= 16u) ok = false;
	}
	return ok;
}
// one thread: the masks of the exchanged form
B2_HD void in_line_swap_tables(const BottomParams& B, BottomSmem& bs, const uint32_t* consts) {
	if (!B.swap_inline) return;
	for (int s = 0; s < 3; s++) {
		uint32_t D[4] = {0, 0, 0, 0};
		for (int j = 0; j <= 2 - s; j++) {
			const uint32_t c = consts[s * 32 + j], m = ~lane_mask_lo(s + 2 + j);
			for (int r = 0; r < 4; r++) D[r] ^= bitmask(c, r) & m;
		}
		for (int i = 0; i < 4; i++)
			for (int k = 0; k < 4; k++) {
				uint32_t m = 0;
				for (int r = 0; r < 4; r++)
					if ((Tower<2>::mul(1u << r, 1u << k) >> i) & 1u) m ^= D[r];
				bs.CW[96 + s * 16 + 4 * i + k] = m;
			}
	}
}

// once per CTA: tile-independent tables
B2_HD void bottom_prologue(const BottomParams& B, BottomSmem& bs, int tid, int nthreads, const uint32_t* consts) {
	for (int i = tid; i < 160; i += nthreads) bs.LT[i] = B.lane_tw[i];
	for (int i = tid; i < 64; i += nthreads) bs.rows[i] = lane_row_entry(i);
	if (tid < B.n_inline) in_line_split_tables(B, bs, tid, consts);
	if (tid == 8) in_line_swap_tables(B, bs, consts);
	for (int idx = tid; idx < 512; idx += nthreads) {
		if (idx == 0) {
			bs.TL[0] = 0;
			continue;
		}
		int lg = 0;
		while (idx >> (lg + 1)) lg++;
		const int bit = 8 - lg;  // bit b owns [2^(8-b), 2^(9-b))
		bs.TL[idx] = bit >= B.batch_bits ? xsel(consts + bottom_stage(B, bit) * 32, (uint64_t)(idx - (1 << lg))) : 0u;
	}
	// in-line stage s, local line l: block index (butterfly-line part of l) << (n_inline-1-s)
	for (int idx = tid; idx < B.n_inline * 512; idx += nthreads) {
		const int s = idx >> 9, l = idx & 511;
		bs.IL[idx] = xsel(consts + s * 32, (uint64_t)(l >> B.batch_bits) << (B.n_inline - 1 - s));
	}
}
// once per tile: the tile- and coset-dependent parts (threads 0..15)
B2_HD void bottom_tile_tables(const BottomParams& B, BottomSmem& bs, const UGroup& c, int tid, const uint32_t* consts) {
	const UpperParams& P = B.up;
	const uint64_t coset = (uint64_t)(P.coset0 + (c.row & P.coset_mask));
	if (tid < 0) return;
	if (tid < 8) {  // line bit tid (only batch_bits..4 are used here; 5..8 go through the leaf masks)
		const int bit = tid;
		if (bit >= B.batch_bits && bit < 5) {
			const int stage = bottom_stage(B, bit);
			bs.TH[bit] = xsel(consts + stage * 32, (coset << (P.log_h - 1 - stage)) | (c.hi << (8 - bit)));
		}
	} else if (tid < 16) {
		const int s = tid - 8;
		if (s < B.n_inline) bs.IH[s] = xsel(consts + s * 32, (coset << (P.log_h - 1 - s)) | (c.hi << (9 - B.batch_bits + B.n_inline - 1 - s)));
	}
}

// Phase B for one lane: the warp's super-lines 4*warp .. 4*warp+3, two at a time (iter; 16 lanes x 2 lines each).
// n-th line layer in execution order, n < 5 - batch_bits; callers put __syncwarp() between layers.  The in-line
// layers and the transposes ride on the layer that touches a line last (forward) or first (inverse).
template <bool INV, bool SWAP = false, bool ALL5 = false>
B2_HD void bottom_phase_b(const BottomParams& B, const BottomSmem& bs, int warp, int lane, int iter, int n) {
	const UpperParams& P = B.up;
	const int bb = B.batch_bits, L0 = B.n_inline;
	const int sl = 4 * warp + 2 * iter + (lane >> 4), tt = lane & 15;
	uint32_t* slot0 = bs.us.slots + (size_t)sl * 32 * kUSlotWords;
	const int nline = 5 - bb;
	uint32_t U[32], V[32];
	{
		const int b = INV ? bb + n : 4 - n;  // forward: 4 .. bb, inverse: bb .. 4
		const int lu = ((tt >> b) << (b + 1)) | (tt & ((1 << b) - 1)), lv = lu | (1 << b);
		uint32_t* su = slot0 + (size_t)lu * kUSlotWords;
		uint32_t* sv = slot0 + (size_t)lv * kUSlotWords;
		// exchanged form of the in-line layers (in_line_stage_swapped): compact GF(2^128) images with GF(16) low constants
		constexpr bool swap = SWAP;  // chosen by the launch: B.swap_inline and a compact image on the in-line side
		if (INV && n == 0 && swap)
			swapped_load(su, sv, U, V);
		else {
			line_load(su, U);
			line_load(sv, V);
		}
		const int l_u = (sl << 5) | lu, l_v = (sl << 5) | lv;
		if (INV && n == 0) {  // first touch: back from the compact image, then the in-line layers upwards
			if (!P.src_sliced) {
				transpose32(U);
				transpose32(V);
			}
			if constexpr (swap) {
#pragma unroll 1
				for (int s = 0; s < 3; s++) in_line_stage_swapped<true>(U, V, s + 2, bs.CW + 96 + s * 16, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.rows);
			} else
#pragma unroll 1
			for (int s = 0; s < L0; s++) {
				if (bs.CW[80 + s])
					butterfly_in_line_split<true>(U, V, s + B.lane_shift, bs.CW + s * 16, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.rows);
				else
					in_line_dispatch<true>(B.cls[s], U, V, s + B.lane_shift, bs.LT + s * 32, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.IH[s] ^ bs.IL[s * 512 + l_v]);
			}
		}
		const uint32_t w = bs.TH[b] ^ bs.TL[(1 << (8 - b)) + (l_u >> (b + 1))];
		// twiddles differ per lane: class 5 through the static GF(16) row table (IMAD + LOP3), smaller classes two-sided
		const int cls = B.cls[bottom_stage(B, b)];
		if (ALL5 || cls == 5)
			butterfly_lane<INV>(U, V, w, bs.rows);
		else if constexpr (!ALL5)
			uniform_dispatch<INV>(cls, U, V, w);
		if (!INV && n == nline - 1) {  // last touch: the in-line layers downwards, then the compact image
			if constexpr (swap) {
#pragma unroll 1
				for (int s = 2; s >= 0; s--) in_line_stage_swapped<false>(U, V, s + 2, bs.CW + 96 + s * 16, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.rows);
			} else
#pragma unroll 1
			for (int s = L0 - 1; s >= 0; s--) {
				if (bs.CW[80 + s])
					butterfly_in_line_split<false>(U, V, s + B.lane_shift, bs.CW + s * 16, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.rows);
				else
					in_line_dispatch<false>(B.cls[s], U, V, s + B.lane_shift, bs.LT + s * 32, bs.IH[s] ^ bs.IL[s * 512 + l_u], bs.IH[s] ^ bs.IL[s * 512 + l_v]);
			}
			if (!P.dst_sliced) {
				transpose32(U);
				transpose32(V);
			}
		}
		if (!INV && n == nline - 1 && swap)
			swapped_store(su, sv, U, V);
		else {
			line_store(su, U);
			line_store(sv, V);
		}
	}
}

}  // namespace b2
