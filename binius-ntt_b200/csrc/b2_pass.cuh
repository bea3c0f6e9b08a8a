// b2_pass.cuh -- one fused pass of the additive NTT over bundles.
//
// The transform of ntt/additive_ntt.cuh:127-159 runs its layers from the highest stride down.
// A pass fuses `k` consecutive line-stride layers (and, in the last pass, the layers inside a
// line) so that the data set crosses HBM once per pass:
//
//   element index  e  = line * EPL + lane_elem          EPL = 32 / limbs elements per 128-B line
//   line index     l  = (hi << (q0+k)) | (x << q0) | lo   x = the k line bits this pass owns
//   a tile             = the 2^k lines of one (coset row, hi, lo)
//   a group            = the 2T lines a CTA of T threads holds in shared memory at a time:
//                        2T >> k tiles with consecutive tile numbers (neighbouring `lo`)
//   thread t of a tile = one butterfly per layer: layer i (line bit q0+i) pairs the lines
//                        x_u = t with a 0 inserted at bit i, x_v = x_u | 1 << i
//
// Lines travel global -> shared memory as coalesced 16-byte chunks (cp.async on the GPU), are
// held bit-sliced ("bundles") while the layers run, and go back the same way.  Between passes
// the data set stays bit-sliced in HBM; only the first load and the last store of a transform on
// compact images transpose.
//
// Twiddles are never recomputed per butterfly: a layer's twiddle is linear in the butterfly-block
// index (additive_ntt.cuh:59-77), so w = tw_hi[tile][layer] ^ tw_lo[layer][t >> i], two tables in
// shared memory filled once per CTA (tw_lo) and once per group (tw_hi).
//
// The per-thread work is written as phase functions that take the thread identity as arguments:
// the __global__ wrapper in b2_kernels.cu calls them with __syncthreads() between phases, and
// tests/emul drives the very same functions thread by thread on the CPU, which is how the
// indexing is checked in a container without a GPU.
#pragma once
#include "b2_bundle.cuh"
#include "b2_constmat.cuh"

namespace b2 {

constexpr int kMaxPassBits = 9;      // up to 512 lines (64 KiB) per group
constexpr int kMaxCtaThreads = 256;  // 2^(kMaxPassBits-1)
constexpr int kConstPitch = 32;      // words per layer in the constants table
constexpr int kSlotWords = 36;       // padded words per 32-word line slot in shared memory

struct PassParams {
	const uint32_t* src;   // pass input: row r reads words [ (r >> src_row_shift) * src_row_words, ... )
	uint32_t* dst;         // pass output: row r at r * dst_row_words
	uint64_t src_row_words;
	uint64_t dst_row_words;
	int src_row_shift;     // log_rate when the first pass fans one input column out to its coset rows, else 0
	uint32_t coset_mask;   // row & coset_mask = coset index of the row (rows = columns x cosets)
	const uint32_t* consts;   // [32][kConstPitch] normalised subspace constants (additive_ntt.cuh:273-309)
	const uint32_t* lane_tw;  // [5][32] per-lane twiddle words of the in-line layers (both lane halves)
	int log_h;
	int log_rate;
	int limbs_log;   // 0: GF(2^32), 2: GF(2^128)
	int n_inline;    // butterfly layers inside a line: 5 - limbs_log (compact), 5 (bit-sliced images)
	int lane_shift;  // in-line layer s pairs lanes differing in bit s + lane_shift
	int batch_bits;  // low line bits that are not butterfly bits (the limb index of a bit-sliced image)
	int src_sliced;  // 1: lines of the pass input are bit-sliced words (no transpose on load)
	int dst_sliced;  // 1: lines of the pass output stay bit-sliced (no transpose on store)
	int nb;          // line bits per row = log_h + limbs_log - 5
	int q0;          // first line bit of this pass (>= batch_bits)
	int k;           // line bits fused in this pass (1..log2(2T))
	int in_line;     // 1: also run the layers inside a line (the pass with q0 == batch_bits)
	int inverse;     // 0: forward, 1: inverse transform (layers in the opposite order)
	int coset0;      // coset index of row 0 (inverse transform of one coset, forward transform of one coset)
	int stage_shift; // added to the local stage number when looking up constants and classes (top phase of a split transform)
	int cta_threads; // T
	int group_lines; // G = 2 * T * (butterflies per thread per layer): lines a CTA holds at a time
	int log_tpg;     // log2 of tiles per group = log2(2T) - k
	uint64_t n_tiles;   // rows << (nb - k)
	uint64_t n_groups;  // ceil(n_tiles / (2T >> k))
	uint8_t cls[32];    // per layer: 0 = twiddle is always 0, else log2 of the twiddle subfield bits (2..5)
	// Top phase of a transform split over 2^k devices with the exchange fused into the store: line (x, lo) of the
	// pass output goes to device x's block buffer, slot scatter_rank, line lo -- 16-byte stores into peer memory over NVLink.
	int scatter_n;         // 0: ordinary pass; else 2^k = number of destination buffers
	int scatter_rank;      // this device's slot in every destination block
	uint32_t* scatter[16]; // destination block of device x (mapped in this process: peer access or CUDA IPC)
};

struct Bundles {
	uint32_t A[32];
	uint32_t B[32];
};

// shared-memory image (32-bit words), sized by T = cta_threads
struct PassSmem {
	uint32_t* lane_tw;  // [5][32]
	uint32_t* tw_lo;    // [2T]   heap order: layer i owns [2^(k-1-i), 2^(k-i))
	uint32_t* tw_hi;    // [T]    [tile_in_group * k + i]
	uint32_t* slots;    // [2T][32]
	uint32_t* inl_lo;   // [5][2T] in-line layers, by line x of the tile
	uint32_t* inl_hi;   // [5][T]  in-line layers, by tile of the group
	Coef4* rows;        // [4][16] static GF(16) coefficient rows of butterfly_lane_dispatch (b2_constmat.cuh)
};
B2_HD size_t pass_smem_words(int G, int in_line) { return 256 + 160 + G + G / 2 + kSlotWords * (size_t)G + (in_line ? 5 * (size_t)G + 5 * (size_t)G / 2 : 0); }
B2_HD PassSmem pass_smem_view(uint32_t* base, int G) {
	PassSmem sm;
	sm.slots = base;  // first: 16-byte aligned for the vector accesses
	sm.rows = reinterpret_cast<Coef4*>(sm.slots + kSlotWords * (size_t)G);  // 16-byte aligned: kSlotWords is a multiple of 4
	sm.lane_tw = sm.slots + kSlotWords * (size_t)G + 256;
	sm.tw_lo = sm.lane_tw + 160;
	sm.tw_hi = sm.tw_lo + G;
	sm.inl_lo = sm.tw_hi + G / 2;
	sm.inl_hi = sm.inl_lo + 5 * G;
	return sm;
}

// Slots are padded to 36 words: a quarter warp reading the same 16-byte group of eight
// consecutive slots then touches eight different bank groups, and every access of a thread to its
// slot is a fixed offset from one base address (no per-access index arithmetic).
B2_HD int slot_word(int slot, int word) { return slot * kSlotWords + word; }

// XOR of table[b] over the set bits b of `bits`, table in global or shared memory
B2_HD uint32_t xor_select64(const uint32_t* table, uint64_t bits) {
	uint32_t w = 0;
#pragma unroll 1
	for (int b = 0; bits; b++, bits >>= 1)
		if (bits & 1u) w ^= table[b];
	return w;
}

B2_HD int pass_stage(const PassParams& P, int i) { return P.n_inline + P.q0 - P.batch_bits + i + P.stage_shift; }

// once per CTA: tile-independent twiddle tables
B2_HD void pass_prologue(const PassParams& P, PassSmem& sm, int tid) {
	const int T = P.cta_threads;
	for (int i = tid; i < 5 * 32; i += T) sm.lane_tw[i] = P.lane_tw[i];
	for (int i = tid; i < 64; i += T) sm.rows[i] = lane_row_entry(i);
	for (int idx = tid; idx < (1 << P.k); idx += T) {
		if (idx == 0) {
			sm.tw_lo[0] = 0;
			continue;
		}
		int lg = 31;
		while (!((uint32_t)idx >> lg)) lg--;
		const int i = P.k - 1 - lg;  // layer
		const uint32_t j = (uint32_t)idx - (1u << lg);
		sm.tw_lo[idx] = xor_select64(P.consts + pass_stage(P, i) * kConstPitch, j);
	}
	if (P.in_line) {
		const int L0 = P.n_inline;
		for (int idx = tid; idx < L0 << P.k; idx += T) {
			const int s = idx >> P.k, x = idx & ((1 << P.k) - 1);
			sm.inl_lo[s * P.group_lines + x] = xor_select64(P.consts + s * kConstPitch, (uint64_t)x << (L0 - 1 - s));
		}
	}
}

struct TileCoord {
	bool active;
	int tile_in_group;
	int slot_base;     // first slot of this tile inside the group
	int t;             // thread index inside the tile
	uint32_t row;      // coset row
	uint64_t hi, lo;   // line index parts
};

B2_HD TileCoord tile_coord_of(const PassParams& P, uint64_t group, int tile_in_group) {
	TileCoord c;
	const uint64_t tile = group * (uint64_t)(P.group_lines >> P.k) + tile_in_group;  // tiles per group = G >> k
	c.tile_in_group = tile_in_group;
	c.t = 0;
	c.active = tile < P.n_tiles;
	c.slot_base = tile_in_group << P.k;
	const int tb = P.nb - P.k;  // tile bits per row
	c.row = (uint32_t)(tile >> tb);
	const uint64_t tau = tile & ((uint64_t(1) << tb) - 1);
	c.lo = tau & ((uint64_t(1) << P.q0) - 1);
	c.hi = tau >> P.q0;
	return c;
}
B2_HD TileCoord tile_coord(const PassParams& P, uint64_t group, int tid) {
	TileCoord c = tile_coord_of(P, group, tid >> (P.k - 1));  // 2^(k-1) threads per tile
	c.t = tid & ((1 << (P.k - 1)) - 1);
	return c;
}

B2_HD uint64_t line_of(const PassParams& P, const TileCoord& c, int x) {
	return (c.hi << (P.q0 + P.k)) | ((uint64_t)x << P.q0) | c.lo;
}

// once per group: the tile-dependent part of every layer's twiddle
B2_HD void group_prologue(const PassParams& P, PassSmem& sm, uint64_t group, int tid) {
	const int T = P.cta_threads;
	const int tpg = P.group_lines >> P.k;
	for (int idx = tid; idx < tpg * P.k; idx += T) {
		const int tig = idx / P.k, i = idx % P.k;
		const TileCoord c = tile_coord_of(P, group, tig);
		const int stage = pass_stage(P, i);
		const uint64_t ind = ((uint64_t)(P.coset0 + (c.row & P.coset_mask)) << (P.log_h - 1 - stage)) | (c.hi << (P.k - 1 - i));
		sm.tw_hi[idx] = c.active ? xor_select64(P.consts + stage * kConstPitch, ind) : 0u;
	}
	if (P.in_line) {
		const int L0 = P.n_inline;
		for (int idx = tid; idx < tpg * L0; idx += T) {
			const int tig = idx / L0, s = idx % L0;
			const TileCoord c = tile_coord_of(P, group, tig);
			const uint64_t ind = ((uint64_t)(P.coset0 + (c.row & P.coset_mask)) << (P.log_h - 1 - s)) | (c.hi << (P.k + L0 - 1 - s));
			sm.inl_hi[s * (P.group_lines / 2) + tig] = c.active ? xor_select64(P.consts + s * kConstPitch, ind) : 0u;
		}
	}
}

// ---- global <-> shared memory, 16-byte chunks, coalesced over the group's lines -----------------
B2_HD void copy16_g2s(uint32_t* s, const uint32_t* g) {
#if defined(__CUDA_ARCH__)
	const unsigned sa = (unsigned)__cvta_generic_to_shared(s);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(g) : "memory");
#else
	for (int j = 0; j < 4; j++) s[j] = g[j];
#endif
}
B2_HD void copy_wait_all() {
#if defined(__CUDA_ARCH__)
	asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
#endif
}
B2_HD void copy16_s2g(uint32_t* g, const uint32_t* s) {
#if defined(__CUDA_ARCH__)
	*reinterpret_cast<uint4*>(g) = *reinterpret_cast<const uint4*>(s);
#else
	for (int j = 0; j < 4; j++) g[j] = s[j];
#endif
}

// chunk c of a group: piece = c & 7 (16 bytes of a line), then the tile (neighbouring lines in
// memory), then the line x of the tile.  With k >= 4 a thread keeps its (piece, tile) and only
// walks x, so the addresses advance by constant strides.
template <bool STORE>
B2_HD void group_copy(const PassParams& P, PassSmem& sm, uint64_t group, int tid) {
	const int T = P.cta_threads;
	const int ltpg = P.log_tpg, tpg = 1 << ltpg;
	if (STORE && P.scatter_n) {
		// fused exchange: the tile's line x belongs to device x (the pass owns the top line bits of the local buffer, hi = 0)
#pragma unroll 1
		for (int c = tid; c < 8 * P.group_lines; c += T) {
			const int piece = c & 7, r = c >> 3;
			const int tig = r & (tpg - 1), x = r >> ltpg;
			const TileCoord tc = tile_coord_of(P, group, tig);
			if (!tc.active) continue;
			const uint64_t off = ((((uint64_t)P.scatter_rank << P.q0) | tc.lo) << 5) + piece * 4;
			copy16_s2g(P.scatter[x] + (uint64_t)tc.row * P.dst_row_words + off, &sm.slots[slot_word(tc.slot_base + x, piece * 4)]);
		}
		return;
	}
	if (8 * tpg <= T) {
		const int piece = tid & 7, r0 = tid >> 3;
		const TileCoord tc = tile_coord_of(P, group, r0 & (tpg - 1));
		if (tc.active) {
			const int x0 = r0 >> ltpg, dx = T >> (3 + ltpg), nit = (8 * P.group_lines) / T;
			uint32_t* s = sm.slots + (tc.slot_base + x0) * kSlotWords + piece * 4;
			const uint64_t off = line_of(P, tc, x0) * 32 + piece * 4;
			const uint64_t gstep = ((uint64_t)dx << P.q0) * 32;
			uint32_t* gd = P.dst + (uint64_t)tc.row * P.dst_row_words + off;
			const uint32_t* gs = P.src + (uint64_t)(tc.row >> P.src_row_shift) * P.src_row_words + off;
#pragma unroll 4
			for (int it = 0; it < nit; it++) {
				if (STORE)
					copy16_s2g(gd, s);
				else
					copy16_g2s(s, gs);
				s += dx * kSlotWords, gd += gstep, gs += gstep;
			}
		}
	} else {
#pragma unroll 1
		for (int c = tid; c < 8 * P.group_lines; c += T) {
			const int piece = c & 7, r = c >> 3;
			const int tig = r & (tpg - 1), x = r >> ltpg;
			const TileCoord tc = tile_coord_of(P, group, tig);
			if (!tc.active) continue;
			uint32_t* s = &sm.slots[slot_word(tc.slot_base + x, piece * 4)];
			const uint64_t off = line_of(P, tc, x) * 32 + piece * 4;
			if (STORE)
				copy16_s2g(P.dst + (uint64_t)tc.row * P.dst_row_words + off, s);
			else
				copy16_g2s(s, P.src + (uint64_t)(tc.row >> P.src_row_shift) * P.src_row_words + off);
		}
	}
	if (!STORE) copy_wait_all();
}

B2_HD void slot_load(const PassSmem& sm, int slot, uint32_t* w) {
#pragma unroll
	for (int g = 0; g < 8; g++) {
		const int base = slot_word(slot, 4 * g);
#if defined(__CUDA_ARCH__)
		uint4 v = *reinterpret_cast<const uint4*>(&sm.slots[base]);
		w[4 * g] = v.x, w[4 * g + 1] = v.y, w[4 * g + 2] = v.z, w[4 * g + 3] = v.w;
#else
		for (int j = 0; j < 4; j++) w[4 * g + j] = sm.slots[base + j];
#endif
	}
}
B2_HD void slot_store(PassSmem& sm, int slot, const uint32_t* w) {
#pragma unroll
	for (int g = 0; g < 8; g++) {
		const int base = slot_word(slot, 4 * g);
#if defined(__CUDA_ARCH__)
		*reinterpret_cast<uint4*>(&sm.slots[base]) = make_uint4(w[4 * g], w[4 * g + 1], w[4 * g + 2], w[4 * g + 3]);
#else
		for (int j = 0; j < 4; j++) sm.slots[base + j] = w[4 * g + j];
#endif
	}
}

template <bool INV>
B2_HD void uniform_dispatch(int cls, uint32_t* U, uint32_t* V, uint32_t w) {
	switch (cls) {
		case 0: INV ? butterfly_uniform_inv<0>(U, V, w) : butterfly_uniform<0>(U, V, w); break;
		case 2: INV ? butterfly_uniform_inv<2>(U, V, w) : butterfly_uniform<2>(U, V, w); break;
		case 3: INV ? butterfly_uniform_inv<3>(U, V, w) : butterfly_uniform<3>(U, V, w); break;
		case 4: INV ? butterfly_uniform_inv<4>(U, V, w) : butterfly_uniform<4>(U, V, w); break;
		default: INV ? butterfly_uniform_inv<5>(U, V, w) : butterfly_uniform<5>(U, V, w); break;
	}
}

template <bool INV>
B2_HD void in_line_dispatch(int cls, uint32_t* A, uint32_t* B, int p, const uint32_t* lt, uint32_t wa, uint32_t wb) {
	switch (cls) {
		case 0: butterfly_in_line<0, INV>(A, B, p, lt, wa, wb); break;
		case 2: butterfly_in_line<2, INV>(A, B, p, lt, wa, wb); break;
		case 3: butterfly_in_line<3, INV>(A, B, p, lt, wa, wb); break;
		case 4: butterfly_in_line<4, INV>(A, B, p, lt, wa, wb); break;
		default: butterfly_in_line<5, INV>(A, B, p, lt, wa, wb); break;
	}
}

// The layers inside a line for the two lines (xu, xv of tile tig) held by a thread.
// block index of lane pair y of line l at layer s: (l << (L0-1-s)) | (y >> (p+1)); the line part
// is uniform over the line (inl_hi ^ inl_lo), the lane part comes from lane_tw.
template <bool INV>
B2_HD void in_line_layers(const PassParams& P, const PassSmem& sm, Bundles& st, int tig, int xu, int xv) {
	const int L0 = P.n_inline, G = P.group_lines;
#pragma unroll 1
	for (int n = 0; n < L0; n++) {
		const int s = INV ? n : L0 - 1 - n;
		const int p = s + P.lane_shift;
		const uint32_t wh = sm.inl_hi[s * (G / 2) + tig];
		const uint32_t wa = wh ^ sm.inl_lo[s * G + xu];
		const uint32_t wb = wh ^ sm.inl_lo[s * G + xv];
		in_line_dispatch<INV>(P.cls[s], st.A, st.B, p, &sm.lane_tw[s * 32], wa, wb);
	}
}

// One line-stride layer (local bit i) for one thread.
//   first: the bundles were just copied in (transpose compact lines; inverse: in-line layers first)
//   last : forward: in-line layers last; transpose back for a compact output
template <bool INV>
B2_HD void pass_layer(const PassParams& P, PassSmem& sm, uint64_t group, int tid, int i, bool first, bool last) {
	const TileCoord c = tile_coord(P, group, tid);
	if (!c.active) return;
	const int xu = ((c.t >> i) << (i + 1)) | (c.t & ((1 << i) - 1));
	const int xv = xu | (1 << i);
	Bundles st;
	slot_load(sm, c.slot_base + xu, st.A);
	slot_load(sm, c.slot_base + xv, st.B);
	if (first) {
		if (!P.src_sliced) {
			transpose32(st.A);
			transpose32(st.B);
		}
		if (INV && P.in_line) in_line_layers<true>(P, sm, st, c.tile_in_group, xu, xv);
	}
	const uint32_t w = sm.tw_hi[c.tile_in_group * P.k + i] ^ sm.tw_lo[(1 << (P.k - 1 - i)) + (c.t >> i)];
	butterfly_lane_dispatch<INV>(P.cls[pass_stage(P, i)], st.A, st.B, w, sm.rows);  // two-pipe product, any class (round 1: uniform_dispatch)
	if (last) {
		if (!INV && P.in_line) in_line_layers<false>(P, sm, st, c.tile_in_group, xu, xv);
		if (!P.dst_sliced) {
			transpose32(st.A);
			transpose32(st.B);
		}
	}
	slot_store(sm, c.slot_base + xu, st.A);
	slot_store(sm, c.slot_base + xv, st.B);
}

// layer order of a pass: forward runs local bits k-1 .. 0, inverse 0 .. k-1
B2_HD int pass_layer_bit(const PassParams& P, int step) { return P.inverse ? step : P.k - 1 - step; }

}  // namespace b2
