// b2_bsl.cuh -- the lowest pass of the additive NTT in BLOCK-SLICED form: the stages whose stride is below one 4 KB block
// (GF(2^128): stages 0..7, GF(2^32): 0..9), on tiles of 32 consecutive blocks (128 KB contiguous).
//
// Why another slicing.  In the line-sliced bundles of b2_bundle.cuh a register holds one bit of the 32 WORDS OF A LINE, so the
// twiddle of a low stage differs from thread to thread (line layers) or from bit position to bit position (layers inside a
// line), and the butterfly needs the two-sided product (1087 LOP3 + masked shifts; b2_bottom.cuh).  Here a register set holds
// word `wi` of 32 DIFFERENT BLOCKS (bit position p = block p of the tile, register i = bit i):
//   * the partner of a butterfly (word wi ^ 2^(s + limbs_log) of the same blocks) is another register set: no shifts;
//   * the twiddle differs per bit position, but the SAME 32 twiddle words serve every butterfly of the tile whose element
//     offset agrees above the stage bit ("class" c), for all limbs: by linearity of calculate_twiddle (ntt/additive_ntt.cuh:59-77)
//         W[i] = Pat_s[i] ^ (bit i of u ? ~0 : 0),    u = T_tile[s] ^ T_c[s][c]   (scalar),
//     Pat_s = the bit-sliced contribution of the block number (5 constants per stage), so the w side of the Karatsuba circuit
//     (leaf masks, b2_upper.cuh) is evaluated once per (stage, class) into shared memory by a few threads and every butterfly
//     runs the V-side-only circuit ConstMul<5> with pattern masks: 876 LOP3 instead of 1087 + shifts.
// Stages with more than 64 classes per tile (GF(2^128): stage 0; GF(2^32): stages 0..2) keep the two-sided product with the
// thread's own W words.  The 32 x 32 bit transposes between the compact image and the sliced form happen once per tile on
// the way in and once on the way out; the pass reads and writes compact words (the pass above it stores compact lines).
//
// Replaces additive_ntt_kernel (ntt/additive_ntt.cuh:91-160) for these stages.  Phase functions take the thread id so that
// tests/emul drives the same code on the CPU.
#pragma once
#include "b2_upper.cuh"

namespace b2 {

constexpr int kBslThreads = 256;
constexpr int kBslBlockBits = 10;                   // words per block: 1024 (4 KB)
constexpr int kBslTileWords = 32 << kBslBlockBits;  // 32 blocks
constexpr int kBslMaxClasses = 64;                  // classes per stage that get a leaf table
constexpr int kBslMaxStages = 10;

struct BslParams {
	const uint32_t* src;   // row r reads words [(r >> src_row_shift) * src_row_words, ...)
	uint32_t* dst;         // row r at r * dst_row_words
	uint64_t src_row_words, dst_row_words;
	int src_row_shift;
	uint32_t coset_mask;
	int coset0;
	const uint32_t* consts;  // [32][kConstPitch]
	int log_h, limbs_log;
	int inverse;
	int tiles_per_row_log;   // log2(row words) - 15
	uint64_t n_tiles;        // rows << tiles_per_row_log
};

struct BslSmem {
	uint32_t* T;       // [32][1024]: compact: block p, word wi at p * 1024 + wi; sliced: bit i of register set wi at i * 1024 + wi
	Mask4* tab;        // [kBslMaxClasses][81] leaf masks of the current stage
	uint32_t* consts;  // [32][32]
	uint32_t* pat;     // [kBslMaxStages][32] bit-sliced block-number part of the twiddle
	uint32_t* ttile;   // [kBslMaxStages] tile- and coset-dependent part
};
B2_HD size_t bsl_smem_words() { return (size_t)kBslTileWords + (size_t)kBslMaxClasses * 81 * 4 + 1024 + kBslMaxStages * 32 + 16; }
B2_HD BslSmem bsl_smem_view(uint32_t* base) {
	BslSmem sm;
	sm.T = base;
	sm.tab = reinterpret_cast<Mask4*>(base + kBslTileWords);
	sm.consts = base + kBslTileWords + kBslMaxClasses * 81 * 4;
	sm.pat = sm.consts + 1024;
	sm.ttile = sm.pat + kBslMaxStages * 32;
	return sm;
}

B2_HD int bsl_stages(const BslParams& P) { return kBslBlockBits - P.limbs_log; }
// index bits of the twiddle of stage s that come from the element offset inside the block, above the stage bit
B2_HD int bsl_class_bits(const BslParams& P, int s) { return kBslBlockBits - 1 - P.limbs_log - s; }
B2_HD bool bsl_has_table(const BslParams& P, int s) { return (1 << bsl_class_bits(P, s)) <= kBslMaxClasses; }

// once per CTA: Pat_s[i] bit p = bit i of XOR_{k : bit k of p} C[s][class_bits + k]
B2_HD void bsl_prologue(const BslParams& P, BslSmem& sm, int tid, int nthreads) {
	const int S = bsl_stages(P);
	for (int idx = tid; idx < S * 32; idx += nthreads) {
		const int s = idx >> 5, i = idx & 31, cb = bsl_class_bits(P, s);
		uint32_t w = 0;
		for (int k = 0; k < 5; k++) {
			const uint32_t lanes = k == 0 ? 0xaaaaaaaau : k == 1 ? 0xccccccccu : k == 2 ? 0xf0f0f0f0u : k == 3 ? 0xff00ff00u : 0xffff0000u;
			if (cb + k < 32 && ((sm.consts[s * 32 + cb + k] >> i) & 1u)) w ^= lanes;
		}
		sm.pat[idx] = w;
	}
}

struct BslTile {
	uint32_t row;
	uint64_t tw;  // tile of the row
};
B2_HD BslTile bsl_tile_of(const BslParams& P, uint64_t tile) {
	BslTile t;
	t.row = (uint32_t)(tile >> P.tiles_per_row_log);
	t.tw = tile & ((uint64_t(1) << P.tiles_per_row_log) - 1);
	return t;
}

// once per tile: threads tid < stages
B2_HD void bsl_tile_twiddles(const BslParams& P, BslSmem& sm, const BslTile& t, int tid) {
	if (tid >= bsl_stages(P)) return;
	const int s = tid, cb = bsl_class_bits(P, s);
	const uint64_t coset = (uint64_t)(P.coset0 + (t.row & P.coset_mask));
	const uint64_t ind = (coset << (P.log_h - 1 - s)) | (t.tw << (cb + 5));
	sm.ttile[s] = xor_select64(sm.consts + s * 32, ind);
}

// tile <-> global, 16-byte chunks, 128 KB contiguous
template <bool STORE>
B2_HD void bsl_copy(const BslParams& P, BslSmem& sm, const BslTile& t, int tid) {
	const uint32_t* s = P.src + (uint64_t)(t.row >> P.src_row_shift) * P.src_row_words + t.tw * kBslTileWords;
	uint32_t* d = P.dst + (uint64_t)t.row * P.dst_row_words + t.tw * kBslTileWords;
#pragma unroll 4
	for (int it = 0; it < kBslTileWords / 4 / kBslThreads; it++) {
		const int off = (it * kBslThreads + tid) * 4;
		if (STORE)
			copy16_s2g(d + off, sm.T + off);
		else
			copy16_g2s(sm.T + off, s + off);
	}
}

// compact <-> sliced, in place: thread tid owns the columns wi = tid + 256 j (involution)
B2_HD void bsl_transpose(BslSmem& sm, int tid) {
#pragma unroll 1
	for (int j = 0; j < (1 << kBslBlockBits) / kBslThreads; j++) {
		const int wi = tid + kBslThreads * j;
		uint32_t r[32];
#pragma unroll
		for (int p = 0; p < 32; p++) r[p] = sm.T[(p << kBslBlockBits) + wi];
		transpose32(r);
#pragma unroll
		for (int i = 0; i < 32; i++) sm.T[(i << kBslBlockBits) + wi] = r[i];
	}
}

// leaf masks of a bit-sliced (per position) twiddle: the w side of the Karatsuba tree on WORDS
template <int H>
struct WordLeaves {
	static constexpr int M = 1 << (H - 1);
	static constexpr int G = 3 * WordLeaves<H - 1>::G;
	static B2_HD void build(const uint32_t* w, Mask4* out) {
		uint32_t s[M];
#pragma unroll
		for (int i = 0; i < M; i++) s[i] = w[i] ^ w[M + i];
		WordLeaves<H - 1>::build(w, out);
		WordLeaves<H - 1>::build(w + M, out + G / 3);
		WordLeaves<H - 1>::build(s, out + 2 * (G / 3));
	}
};
template <>
struct WordLeaves<1> {
	static constexpr int G = 1;
	static B2_HD void build(const uint32_t* w, Mask4* out) {
		Mask4 q;
		q.x = w[0], q.y = w[1], q.z = w[0] ^ w[1], q.w = 0;
		out[0] = q;
	}
};

// the 32 twiddle words of (stage s, class c)
B2_HD void bsl_twiddle_words(const BslSmem& sm, int s, int c, uint32_t* W) {
	const uint32_t u = sm.ttile[s] ^ xor_select(sm.consts + s * 32, (uint32_t)c);
#pragma unroll
	for (int i = 0; i < 32; i++) W[i] = sm.pat[s * 32 + i] ^ bitmask(u, i);
}

// before a table stage: item = class * 3 + top-level digit (low half, high half, their sum), one item per thread
B2_HD void bsl_build_tables(const BslParams& P, BslSmem& sm, int s, int tid) {
	const int classes = 1 << bsl_class_bits(P, s);
	if (tid >= classes * 3) return;
	const int c = tid / 3, d = tid - 3 * c;
	uint32_t W[32], h[16];
	bsl_twiddle_words(sm, s, c, W);
#pragma unroll
	for (int i = 0; i < 16; i++) h[i] = d == 0 ? W[i] : d == 1 ? W[16 + i] : (W[i] ^ W[16 + i]);
	WordLeaves<4>::build(h, sm.tab + c * 81 + d * 27);
}

// one stage: thread tid runs the butterflies tid and tid + 256 of the tile's 512
template <bool INV>
B2_HD void bsl_stage(const BslParams& P, BslSmem& sm, int s, int tid) {
	const int bp = s + P.limbs_log;  // bit of the word index that the stage pairs
	const bool table = bsl_has_table(P, s);
#pragma unroll 1
	for (int b = tid; b < (1 << (kBslBlockBits - 1)); b += kBslThreads) {
		const int wu = ((b >> bp) << (bp + 1)) | (b & ((1 << bp) - 1)), wv = wu | (1 << bp);
		const int c = wu >> (bp + 1);
		uint32_t U[32], V[32];
#pragma unroll
		for (int i = 0; i < 32; i++) U[i] = sm.T[(i << kBslBlockBits) + wu], V[i] = sm.T[(i << kBslBlockBits) + wv];
		if (table) {
			butterfly_const<5, INV>(U, V, sm.tab + c * 81);
		} else {
			uint32_t W[32];
			bsl_twiddle_words(sm, s, c, W);
			if (INV) {
#pragma unroll
				for (int i = 0; i < 32; i++) V[i] ^= U[i];
			}
			Sliced<5>::mul_acc(W, V, U);
			if (!INV) {
#pragma unroll
				for (int i = 0; i < 32; i++) V[i] ^= U[i];
			}
		}
#pragma unroll
		for (int i = 0; i < 32; i++) sm.T[(i << kBslBlockBits) + wu] = U[i], sm.T[(i << kBslBlockBits) + wv] = V[i];
	}
}

}  // namespace b2
