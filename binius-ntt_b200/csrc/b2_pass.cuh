// b2_pass.cuh -- one fused pass of the additive NTT over bundles.
//
// The transform of ntt/additive_ntt.cuh:127-159 runs its layers from the highest stride down.
// A pass fuses `k` consecutive line-stride layers (and, in the last pass, the layers inside a
// line) so that the data set crosses HBM once per pass:
//
//   element index  e  = line * EPL + lane_elem          EPL = 32 / limbs elements per 128-B line
//   line index     l  = (hi << (q0+k)) | (x << q0) | lo   x = the k line bits this pass owns
//   a tile             = the 2^k lines of one (coset row, hi, lo); it lives in shared memory,
//                        bit-sliced, for the whole pass
//   thread t of a tile = one butterfly per layer: layer i (line bit q0+i) pairs the lines
//                        x_u = t with a 0 inserted at bit i, x_v = x_u | 1 << i
//
// The per-thread work is written as phase functions (prologue / layer / epilogue) that take the
// thread identity as arguments: the __global__ wrapper in b2_kernels.cu calls them with
// __syncthreads() between phases, and tests/emul drives the very same functions thread by thread
// on the CPU, which is how the indexing is checked in a container without a GPU.
#pragma once
#include "b2_bundle.cuh"

namespace b2 {

constexpr int kMaxPassBits = 9;        // up to 512 lines (64 KiB) per tile group
constexpr int kCtaThreads = 256;       // 2^(kMaxPassBits-1)
constexpr int kLinesPerCta = 512;      // tiles of 2^k lines are grouped to fill this
constexpr int kConstPitch = 32;        // words per layer in the constants table

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
	int sliced_io;   // 1: lines in memory are already bit-sliced words (no transposes on load / store)
	int nb;          // line bits per row = log_h + limbs_log - 5
	int q0;          // first line bit of this pass (>= batch_bits)
	int k;           // line bits fused in this pass (1..kMaxPassBits)
	int in_line;     // 1: also run the layers inside a line (the pass with q0 == batch_bits)
	int inverse;     // 0: forward, 1: inverse transform (layers in the opposite order)
	int coset0;      // coset index of row 0 (inverse transform of one coset)
	uint64_t n_tiles;  // rows << (nb - k)
	uint8_t cls[32];   // per layer: 0 = twiddle is always 0, else log2 of the twiddle subfield bits (2..5)
};

struct Bundles {
	uint32_t A[32];
	uint32_t B[32];
};

// shared-memory image: constants, then bundle slots
struct PassSmem {
	uint32_t consts[32 * kConstPitch];
	uint32_t lane_tw[5 * 32];
	uint32_t pad[96];
	uint32_t slots[kLinesPerCta * 32];
};
static_assert(sizeof(PassSmem) == 4096 + 640 + 384 + 65536, "layout");

// 16-byte groups of a slot are XOR-swizzled by the slot number so that a quarter warp that reads
// group g of eight different slots touches eight different bank groups.
B2_HD int slot_word(int slot, int word) {
	int g = (word >> 2) ^ ((slot ^ (slot >> 3) ^ (slot >> 6)) & 7);
	return slot * 32 + (g << 2) + (word & 3);
}

B2_HD void pass_prologue(const PassParams& P, PassSmem& sm, int tid, int nthreads) {
	for (int i = tid; i < 32 * kConstPitch; i += nthreads) sm.consts[i] = P.consts[i];
	for (int i = tid; i < 5 * 32; i += nthreads) sm.lane_tw[i] = P.lane_tw[i];
}

struct TileCoord {
	bool active;
	int slot_base;     // first slot of this thread's tile inside the CTA
	int t;             // thread index inside the tile
	uint32_t row;      // coset row
	uint64_t hi, lo;   // line index parts
};

B2_HD TileCoord tile_coord(const PassParams& P, uint32_t cta, int tid) {
	TileCoord c;
	const int tpt = 1 << (P.k - 1);  // threads per tile
	const int tile_in_cta = tid / tpt;
	c.t = tid % tpt;
	const uint64_t tile = (uint64_t)cta * (kCtaThreads / tpt) + tile_in_cta;
	c.active = tile < P.n_tiles;
	c.slot_base = tile_in_cta << P.k;
	const int tb = P.nb - P.k;  // tile bits per row
	c.row = (uint32_t)(tile >> tb);
	const uint64_t tau = tile & ((uint64_t(1) << tb) - 1);
	c.lo = tau & ((uint64_t(1) << P.q0) - 1);
	c.hi = tau >> P.q0;
	return c;
}

B2_HD uint64_t line_of(const PassParams& P, const TileCoord& c, int x) {
	return (c.hi << (P.q0 + P.k)) | ((uint64_t)x << P.q0) | c.lo;
}

B2_HD void load_line(const uint32_t* p, uint32_t* w) {
#if defined(__CUDA_ARCH__)
	const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
	for (int g = 0; g < 8; g++) {
		uint4 v = q[g];
		w[4 * g] = v.x, w[4 * g + 1] = v.y, w[4 * g + 2] = v.z, w[4 * g + 3] = v.w;
	}
#else
	for (int i = 0; i < 32; i++) w[i] = p[i];
#endif
}
B2_HD void store_line(uint32_t* p, const uint32_t* w) {
#if defined(__CUDA_ARCH__)
	uint4* q = reinterpret_cast<uint4*>(p);
#pragma unroll
	for (int g = 0; g < 8; g++) q[g] = make_uint4(w[4 * g], w[4 * g + 1], w[4 * g + 2], w[4 * g + 3]);
#else
	for (int i = 0; i < 32; i++) p[i] = w[i];
#endif
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

// twiddle of layer `stage` for the butterfly block `block` of coset row `row`
B2_HD uint32_t layer_twiddle(const PassParams& P, const PassSmem& sm, int stage, uint32_t row, uint64_t block) {
	const uint64_t ind = ((uint64_t)(P.coset0 + (row & P.coset_mask)) << (P.log_h - 1 - stage)) | block;
	return xor_select(&sm.consts[stage * kConstPitch], (uint32_t)ind);
}

// The layers inside a line for the two lines held by a thread (global line indices la, lb).
template <bool INV>
B2_HD void in_line_layers(const PassParams& P, const PassSmem& sm, Bundles& st, uint32_t row, uint64_t la, uint64_t lb) {
	const int L0 = P.n_inline;
	la >>= P.batch_bits;
	lb >>= P.batch_bits;
#pragma unroll 1
	for (int n = 0; n < L0; n++) {
		const int s = INV ? n : L0 - 1 - n;
		const int p = s + P.lane_shift;
		// block index of lane pair x of line l at layer s: (l << (L0-1-s)) | (x >> (p+1));
		// the line part is uniform over the line, the lane part comes from lane_tw.
		const uint32_t wa = layer_twiddle(P, sm, s, row, la << (L0 - 1 - s));
		const uint32_t wb = layer_twiddle(P, sm, s, row, lb << (L0 - 1 - s));
		in_line_dispatch<INV>(P.cls[s], st.A, st.B, p, &sm.lane_tw[s * 32], wa, wb);
	}
}

// One line-stride layer (local bit i) for one thread.
//   first: operands come from global memory (compact) instead of shared memory
//   last : results stay in registers (st) for the epilogue instead of going to shared memory
template <bool INV>
B2_HD void pass_layer(const PassParams& P, PassSmem& sm, uint32_t cta, int tid, int i, bool first, bool last, Bundles& st) {
	const TileCoord c = tile_coord(P, cta, tid);
	if (!c.active) return;
	const int xu = ((c.t >> i) << (i + 1)) | (c.t & ((1 << i) - 1));
	const int xv = xu | (1 << i);
	if (first) {
		const uint32_t* src = P.src + (uint64_t)(c.row >> P.src_row_shift) * P.src_row_words;
		load_line(src + line_of(P, c, xu) * 32, st.A);
		load_line(src + line_of(P, c, xv) * 32, st.B);
		if (!P.sliced_io) {
			transpose32(st.A);
			transpose32(st.B);
		}
		if (INV && P.in_line) in_line_layers<true>(P, sm, st, c.row, line_of(P, c, xu), line_of(P, c, xv));
	} else {
		slot_load(sm, c.slot_base + xu, st.A);
		slot_load(sm, c.slot_base + xv, st.B);
	}
	const int stage = P.n_inline + P.q0 - P.batch_bits + i;
	const uint64_t block = (c.hi << (P.k - 1 - i)) | (uint64_t)(c.t >> i);
	const uint32_t w = layer_twiddle(P, sm, stage, c.row, block);
	uniform_dispatch<INV>(P.cls[stage], st.A, st.B, w);
	if (!last) {
		slot_store(sm, c.slot_base + xu, st.A);
		slot_store(sm, c.slot_base + xv, st.B);
	} else {
		uint32_t* dst = P.dst + (uint64_t)c.row * P.dst_row_words;
		if (!INV && P.in_line) in_line_layers<false>(P, sm, st, c.row, line_of(P, c, xu), line_of(P, c, xv));
		if (!P.sliced_io) {
			transpose32(st.A);
			transpose32(st.B);
		}
		store_line(dst + line_of(P, c, xu) * 32, st.A);
		store_line(dst + line_of(P, c, xv) * 32, st.B);
	}
}

// layer order of a pass: forward runs local bits k-1 .. 0, inverse 0 .. k-1
B2_HD int pass_layer_bit(const PassParams& P, int step) { return P.inverse ? step : P.k - 1 - step; }

}  // namespace b2
