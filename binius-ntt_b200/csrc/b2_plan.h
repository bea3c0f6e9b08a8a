// b2_plan.h -- host-side plan of an additive NTT: domain constants, twiddle classes, per-lane
// twiddle words and the pass schedule.  Pure C++ (no CUDA calls) so the plan logic is unit-tested
// on the CPU; b2ntt.cu owns the device side.
//
// Reference counterparts: AdditiveNTTConf validation (ntt/nttconf.cuh:55-60),
// precompute_subspace_evals (ntt/additive_ntt.cuh:273-309), and the launch-shape derivation
// (ntt/nttconf.cuh:65-110, ntt/nttconf.cu) which this schedule replaces.
#pragma once
#include <stdint.h>
#include <string.h>

#include <vector>

#include "b2_pass.cuh"
#include "b2_bottom.cuh"
#include "b2_bsl.cuh"

namespace b2 {

struct PassDesc {
	int q0;       // first line bit
	int k;        // line bits fused
	int in_line;  // run the in-line layers in this pass
	int kind;     // 0: bundle pass through shared memory (b2_pass.cuh), 1: upper pass with warp-uniform twiddles (b2_upper.cuh),
	              // 2: lowest pass on 512-line tiles (b2_bottom.cuh), 4: block-sliced lowest pass on 32-block tiles (b2_bsl.cuh)
};

struct HostPlan {
	int field_bits = 0, layout = 0, log_h = 0, log_rate = 0, batch = 1;
	int limbs_log = 0;   // 0 or 2
	int n_inline = 0, lane_shift = 0, batch_bits = 0, sliced_io = 0;
	int nb = 0;          // line bits per row
	bool tiny = false;   // fewer than two lines of butterfly work per row: element-wise kernels
	int cta_threads = 128;  // T threads per CTA
	int group_lines = 512;  // G lines a CTA holds at a time (G / 2T butterflies per thread per layer); <= log2(G) line layers per pass
	uint32_t consts[32 * kConstPitch];
	uint32_t lane_tw[5 * 32];
	uint8_t cls[32];
	std::vector<PassDesc> passes;  // in forward execution order (highest line bits first)
};

inline bool plan_args_ok(int field_bits, int layout, int log_h, int log_rate, int batch) {
	if (field_bits != 32 && field_bits != 128) return false;
	if (layout != 0 && layout != 1) return false;
	if (log_h < 1 || log_rate < 0 || log_rate > 4 || log_h + log_rate > 32) return false;  // nttconf.cuh:55-60
	if (layout == 1 && log_h < 5) return false;  // a bit-sliced block is 32 elements
	if (batch < 1) return false;
	return true;
}

// normalised subspace-polynomial evaluations C[i][j] = W^_i(beta_{i+1+j}), beta_i = 1 << i
inline void plan_constants(int log_h, int log_rate, uint32_t* C /* [32][kConstPitch] */) {
	memset(C, 0, sizeof(uint32_t) * 32 * kConstPitch);
	uint32_t norm[32];
	const int width = log_h + log_rate;
	for (int i = 1; i < width; i++) C[i - 1] = uint32_t(1) << i;
	norm[0] = 1;
	for (int i = 1; i < log_h; i++) {
		const uint32_t* prev = C + (i - 1) * kConstPitch;
		uint32_t* row = C + i * kConstPitch;
		const uint32_t np = norm[i - 1];
		// subspace map e -> e^2 + norm * e (additive_ntt.cuh:16-19)
		norm[i] = gf32_square(prev[0]) ^ gf32_mul(np, prev[0]);
		for (int j = 1; j < width - i; j++) row[j - 1] = gf32_square(prev[j]) ^ gf32_mul(np, prev[j]);
	}
	for (int i = 0; i < log_h; i++) {
		const uint32_t inv = gf32_inverse(norm[i]);
		uint32_t* row = C + i * kConstPitch;
		for (int j = 0; j < width - i - 1; j++) row[j] = gf32_mul(inv, row[j]);
	}
}

inline int twiddle_class(uint32_t or_of_values) {
	if (or_of_values == 0) return 0;
	if (or_of_values < (1u << 4)) return 2;
	if (or_of_values < (1u << 8)) return 3;
	if (or_of_values < (1u << 16)) return 4;
	return 5;
}

inline int plan_build(HostPlan& hp, int field_bits, int layout, int log_h, int log_rate, int batch, int cta_threads = 0, int group_lines = 0,
                      int use_upper = 2) {
	if (!plan_args_ok(field_bits, layout, log_h, log_rate, batch)) return -1;
	hp.field_bits = field_bits, hp.layout = layout, hp.log_h = log_h, hp.log_rate = log_rate, hp.batch = batch;
	hp.limbs_log = field_bits == 128 ? 2 : 0;
	if (layout == 0) {
		hp.n_inline = 5 - hp.limbs_log, hp.lane_shift = hp.limbs_log, hp.batch_bits = 0, hp.sliced_io = 0;
	} else {
		hp.n_inline = 5, hp.lane_shift = 0, hp.batch_bits = hp.limbs_log, hp.sliced_io = 1;
	}
	hp.nb = log_h + hp.limbs_log - 5;
	if (cta_threads != 64 && cta_threads != 128 && cta_threads != 256) cta_threads = 128;
	hp.cta_threads = cta_threads;
	if (group_lines != 128 && group_lines != 256 && group_lines != 512) group_lines = 512;
	if (group_lines < 2 * cta_threads) group_lines = 2 * cta_threads;
	hp.group_lines = group_lines;
	int max_bits = 0;
	while ((1 << max_bits) < group_lines) max_bits++;
	plan_constants(log_h, log_rate, hp.consts);
	for (int s = 0; s < 32; s++) {
		uint32_t m = 0;
		for (int b = 0; b < kConstPitch; b++) m |= hp.consts[s * kConstPitch + b];
		hp.cls[s] = (uint8_t)twiddle_class(m);
	}
	memset(hp.lane_tw, 0, sizeof hp.lane_tw);
	for (int s = 0; s < hp.n_inline && s < log_h; s++)
		for (int x = 0; x < 32; x++) {
			const uint32_t y = (uint32_t)x >> (s + hp.lane_shift + 1);
			uint32_t val = 0;
			for (int b = 0; (y >> b) != 0; b++)
				if ((y >> b) & 1) val ^= hp.consts[s * kConstPitch + b];
			for (int i = 0; i < 32; i++) hp.lane_tw[s * 32 + i] |= ((val >> i) & 1u) << x;
		}
	const int nbf = hp.nb - hp.batch_bits;  // butterfly line bits
	hp.passes.clear();
	hp.tiny = nbf < 1;
	const bool use_upper6 = use_upper != 3;  // 3: as 2 but without the 6-bit passes (A/B runs)
	if (use_upper == 3) use_upper = 2;
	// 5: as 2 with the block-sliced lowest pass (compact images with at least 2^15 words per row); it owns 5 line bits
	const bool bsl = use_upper == 5 && !hp.sliced_io && hp.nb >= 10;
	if (use_upper == 5) use_upper = 2;
	if (!hp.tiny && use_upper == 2 && nbf >= 9 - hp.batch_bits) {
		const int kb = bsl ? 5 : 9 - hp.batch_bits;
		// lowest pass: 512-line tiles (kind 2) = 9 - batch_bits butterfly line bits; above it upper passes of 4
		// (or 2) line bits; an odd remainder goes to one bundle pass of 3 (or 1) bits at the top
		int left = nbf - kb, q = hp.batch_bits + nbf;
		if (left & 1) {
			const int kt = left >= 3 ? 3 : 1;
			q -= kt, left -= kt;
			hp.passes.push_back(PassDesc{q, kt, 0, 0});
		}
		// from the top: 6 line bits per pass while none of their stages has a full GF(2^32) twiddle (class 5 needs the
		// 108-row coefficient table, which only fits for 15 twiddles), then 4, with one pass of 2 when the count asks for it
		auto cheap6 = [&](int qtop) {  // stages of the line bits [qtop - 6, qtop)
			for (int bq = qtop - 6; bq < qtop; bq++)
				if (hp.cls[hp.n_inline + bq - hp.batch_bits] == 5) return false;
			return true;
		};
		while (use_upper6 && left >= 6 && (left - 6) % 2 == 0 && cheap6(q)) {
			q -= 6, left -= 6;
			hp.passes.push_back(PassDesc{q, 6, 0, 1});
		}
		if (left & 3) {
			q -= 2, left -= 2;
			hp.passes.push_back(PassDesc{q, 2, 0, 1});
		}
		while (left > 0) {
			q -= 4, left -= 4;
			hp.passes.push_back(PassDesc{q, 4, 0, 1});
		}
		hp.passes.push_back(PassDesc{hp.batch_bits, kb, 1, bsl ? 4 : 2});
	} else if (!hp.tiny && use_upper && nbf > max_bits && max_bits >= 6) {
		// lowest pass: the bundle pass with the in-line layers (8 or 9 line bits); everything above it
		// has strides >= 32 lines and runs as upper passes of 4 (or 2) line bits, highest first
		int kb = max_bits;
		if ((nbf - kb) & 1) kb--;
		int left = nbf - kb, q = hp.batch_bits + nbf;
		if (left & 3) {
			q -= 2, left -= 2;
			hp.passes.push_back(PassDesc{q, 2, 0, 1});
		}
		while (left > 0) {
			q -= 4, left -= 4;
			hp.passes.push_back(PassDesc{q, 4, 0, 1});
		}
		hp.passes.push_back(PassDesc{hp.batch_bits, kb, 1, 0});
	} else if (!hp.tiny) {
		const int np = (nbf + max_bits - 1) / max_bits;
		// the lowest pass also runs the in-line layers, so it gets the smallest share of line bits
		int left = nbf;
		std::vector<int> ks(np);
		for (int p = np - 1; p >= 0; p--) {  // p = np-1 is the highest pass
			int share = (left + p) / (p + 1);
			ks[p] = share;
			left -= share;
		}
		// ks[0] lowest ... ks[np-1] highest; emit highest first
		int q = hp.batch_bits + nbf;
		for (int p = np - 1; p >= 0; p--) {
			q -= ks[p];
			hp.passes.push_back(PassDesc{q, ks[p], p == 0 ? 1 : 0, 0});
		}
	}
	return 0;
}

// the block-sliced pass reads and writes COMPACT lines: the pass executed before it stores compact, the one after it loads compact
inline bool plan_next_is_bsl(const HostPlan& hp, int n, bool inverse) {
	const int np = (int)hp.passes.size();
	return n + 1 < np && (inverse ? hp.passes[np - 2 - n] : hp.passes[n + 1]).kind == 4;
}
inline bool plan_prev_is_bsl(const HostPlan& hp, int n, bool inverse) {
	const int np = (int)hp.passes.size();
	return n >= 1 && (inverse ? hp.passes[np - n] : hp.passes[n - 1]).kind == 4;
}

// Parameters of the n-th pass in EXECUTION order (forward: highest line bits first; inverse: the
// reverse).  consts / lane_tw point to wherever the kernels will read the tables (device memory on
// the GPU, hp's own arrays in the CPU emulation).  Intermediate passes leave the lines bit-sliced.
inline void plan_fill_pass(const HostPlan& hp, PassParams& P, int n, const uint32_t* src, uint32_t* dst, const uint32_t* consts,
                           const uint32_t* lane_tw, bool inverse, int coset, bool one_coset = false) {
	const int np = (int)hp.passes.size();
	const PassDesc& pd = inverse ? hp.passes[np - 1 - n] : hp.passes[n];
	memset(&P, 0, sizeof P);
	const uint64_t row_words = (uint64_t)32 << hp.nb;
	const bool single = inverse || one_coset;  // rows are columns of ONE coset
	const int rows = single ? hp.batch : hp.batch << hp.log_rate;
	const bool first_pass = n == 0, last_pass = n == np - 1;
	P.src = first_pass ? src : dst;
	P.dst = dst;
	P.src_row_words = row_words;
	P.dst_row_words = row_words;
	P.src_row_shift = (first_pass && !single) ? hp.log_rate : 0;
	P.coset_mask = single ? 0u : ((1u << hp.log_rate) - 1);
	P.consts = consts;
	P.lane_tw = lane_tw;
	P.log_h = hp.log_h, P.log_rate = hp.log_rate, P.limbs_log = hp.limbs_log;
	P.n_inline = hp.n_inline, P.lane_shift = hp.lane_shift, P.batch_bits = hp.batch_bits;
	P.src_sliced = (hp.sliced_io || (!first_pass && !plan_prev_is_bsl(hp, n, inverse))) ? 1 : 0;
	P.dst_sliced = (hp.sliced_io || (!last_pass && !plan_next_is_bsl(hp, n, inverse))) ? 1 : 0;
	P.nb = hp.nb, P.q0 = pd.q0, P.k = pd.k, P.in_line = pd.in_line;
	P.inverse = inverse ? 1 : 0;
	P.coset0 = single ? coset : 0;
	P.cta_threads = hp.cta_threads;
	P.group_lines = hp.group_lines;
	P.log_tpg = 0;
	while ((hp.group_lines >> pd.k) >> (P.log_tpg + 1)) P.log_tpg++;
	P.n_tiles = (uint64_t)rows << (hp.nb - pd.k);
	const uint64_t tpg = (uint64_t)hp.group_lines >> pd.k;
	P.n_groups = (P.n_tiles + tpg - 1) / tpg;
	memcpy(P.cls, hp.cls, 32);
}

// Parameters of the n-th pass in execution order when it is an upper pass (kind 1).
inline void plan_fill_upper(const HostPlan& hp, UpperParams& U, int n, const uint32_t* src, uint32_t* dst, const uint32_t* consts,
                            bool inverse, int coset, bool one_coset = false) {
	const int np = (int)hp.passes.size();
	const PassDesc& pd = inverse ? hp.passes[np - 1 - n] : hp.passes[n];
	memset(&U, 0, sizeof U);
	const uint64_t row_words = (uint64_t)32 << hp.nb;
	const bool single = inverse || one_coset;
	const int rows = single ? hp.batch : hp.batch << hp.log_rate;
	const bool first_pass = n == 0, last_pass = n == np - 1;
	U.src = first_pass ? src : dst;
	U.dst = dst;
	U.src_row_words = U.dst_row_words = row_words;
	U.src_row_shift = (first_pass && !single) ? hp.log_rate : 0;
	U.coset_mask = single ? 0u : ((1u << hp.log_rate) - 1);
	U.coset0 = single ? coset : 0;
	U.consts = consts;
	U.log_h = hp.log_h, U.nb = hp.nb, U.q0 = pd.q0, U.k = pd.k;
	U.ls = pd.k == 6 ? 3 : 5;
	U.stage0 = hp.n_inline + pd.q0 - hp.batch_bits;
	U.src_sliced = (hp.sliced_io || (!first_pass && !plan_prev_is_bsl(hp, n, inverse))) ? 1 : 0;
	U.dst_sliced = (hp.sliced_io || (!last_pass && !plan_next_is_bsl(hp, n, inverse))) ? 1 : 0;
	U.inverse = inverse ? 1 : 0;
	U.n_groups = (uint64_t)rows << (hp.nb - pd.k - U.ls);
	U.lpitch = kLeafRows4;
	for (int i = 0; i < pd.k; i++) {
		U.cls[i] = hp.cls[U.stage0 + i];
		if (U.cls[i] == 5) U.lpitch = kLeafRows5;
	}
}

// Parameters of the lowest pass (kind 2).
inline void plan_fill_bottom(const HostPlan& hp, BottomParams& B, int n, const uint32_t* src, uint32_t* dst, const uint32_t* consts,
                             const uint32_t* lane_tw, bool inverse, int coset, bool one_coset = false) {
	memset(&B, 0, sizeof B);
	plan_fill_upper(hp, B.up, n, src, dst, consts, inverse, coset, one_coset);
	UpperParams& U = B.up;
	U.q0 = 5, U.k = 4, U.ls = 5, U.lpitch = kLeafRows5;
	U.stage0 = hp.n_inline + 5 - hp.batch_bits;
	const bool single = inverse || one_coset;
	const int rows = single ? hp.batch : hp.batch << hp.log_rate;
	U.n_groups = (uint64_t)rows << (hp.nb - 9);
	for (int i = 0; i < 4; i++) U.cls[i] = hp.cls[U.stage0 + i];
	B.lane_tw = lane_tw;
	B.n_inline = hp.n_inline, B.lane_shift = hp.lane_shift, B.batch_bits = hp.batch_bits;
	memcpy(B.cls, hp.cls, 32);
	B.swap_inline = in_line_swap_ok(hp.n_inline, hp.lane_shift, hp.batch_bits, hp.cls, hp.consts, kConstPitch) ? 1 : 0;
}

// Parameters of the block-sliced lowest pass (kind 4).
inline void plan_fill_bsl(const HostPlan& hp, BslParams& L, int n, const uint32_t* src, uint32_t* dst, const uint32_t* consts, bool inverse,
                          int coset, bool one_coset = false) {
	memset(&L, 0, sizeof L);
	const int np = (int)hp.passes.size();
	const uint64_t row_words = (uint64_t)32 << hp.nb;
	const bool single = inverse || one_coset;
	const int rows = single ? hp.batch : hp.batch << hp.log_rate;
	const bool first_pass = n == 0;
	(void)np;
	L.src = first_pass ? src : dst;
	L.dst = dst;
	L.src_row_words = L.dst_row_words = row_words;
	L.src_row_shift = (first_pass && !single) ? hp.log_rate : 0;
	L.coset_mask = single ? 0u : ((1u << hp.log_rate) - 1);
	L.coset0 = single ? coset : 0;
	L.consts = consts;
	L.log_h = hp.log_h, L.limbs_log = hp.limbs_log;
	L.inverse = inverse ? 1 : 0;
	L.tiles_per_row_log = hp.nb - 10;
	L.n_tiles = (uint64_t)rows << L.tiles_per_row_log;
}

// Top phase of a transform split over 2^log_seg devices: the buffer holds
// 2^log_seg segments of 2^(log_h - 2*log_seg) elements each; segment h is the
// slice (h, m, .) of the global index (h: top log_seg bits, m: this device).  Runs the top log_seg
// stages (global stages log_h-1 .. log_h-log_seg) in place, element-wise across the segments: their
// twiddles depend only on the h bits above the butterfly bit, which the local index keeps on top.
// hp is the plan of the FULL transform (log_rate 0, compact or bit-sliced images).  Returns 0 or -1.
inline int plan_fill_top(const HostPlan& hp, PassParams& P, int log_seg, uint32_t* buf, const uint32_t* consts, const uint32_t* lane_tw,
                         void* const* scatter = nullptr, int scatter_rank = 0) {
	if (hp.log_rate != 0 || log_seg < 1 || log_seg > 4) return -1;
	const int nb_local = hp.nb - log_seg;                     // line bits of the local buffer
	const int q0 = nb_local - log_seg;                        // the segment-index bits are its top bits
	if (q0 < hp.batch_bits) return -1;                        // a segment is at least one line
	memset(&P, 0, sizeof P);
	const uint64_t row_words = (uint64_t)32 << nb_local;
	P.src = buf, P.dst = buf;
	P.src_row_words = P.dst_row_words = row_words;
	P.consts = consts, P.lane_tw = lane_tw;
	P.log_h = hp.log_h, P.log_rate = 0, P.limbs_log = hp.limbs_log;
	P.n_inline = hp.n_inline, P.lane_shift = hp.lane_shift, P.batch_bits = hp.batch_bits;
	P.src_sliced = hp.sliced_io, P.dst_sliced = hp.sliced_io;
	P.nb = nb_local, P.q0 = q0, P.k = log_seg, P.in_line = 0;
	P.stage_shift = log_seg;
	P.cta_threads = hp.cta_threads;
	P.group_lines = hp.group_lines;
	while ((hp.group_lines >> log_seg) >> (P.log_tpg + 1)) P.log_tpg++;
	P.n_tiles = (uint64_t)hp.batch << (nb_local - log_seg);
	const uint64_t tpg = (uint64_t)hp.group_lines >> log_seg;
	P.n_groups = (P.n_tiles + tpg - 1) / tpg;
	memcpy(P.cls, hp.cls, 32);
	if (scatter) {  // results leave for the devices' block buffers instead of going back to `buf`
		P.scatter_n = 1 << log_seg, P.scatter_rank = scatter_rank;
		for (int x = 0; x < P.scatter_n; x++) P.scatter[x] = (uint32_t*)scatter[x];
	}
	return 0;
}

}  // namespace b2
