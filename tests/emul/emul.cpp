// emul.cpp -- CPU emulation of the pass kernels: drives the SAME phase functions the CUDA kernels
// call (binius-ntt_b200/csrc/b2_pass.cuh), thread by thread, phase by phase, over host memory.
// TEST INFRASTRUCTURE ONLY (built into tests/emul/libb2emul.so, loaded by the "not gpu" tests to
// check indexing, schedules and circuits where no GPU exists).  Never linked into libb2ntt.so.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../binius-ntt_b200/csrc/b2_plan.h"

using namespace b2;

template <bool INV>
static void run_pass(const PassParams& P) {
	const int T = P.cta_threads;
	std::vector<uint32_t> mem(pass_smem_words(P.group_lines, 1));
	PassSmem sm = pass_smem_view(mem.data(), P.group_lines);
	for (int t = 0; t < T; t++) pass_prologue(P, sm, t);
	for (uint64_t grp = 0; grp < P.n_groups; grp++) {
		for (int t = 0; t < T; t++) group_prologue(P, sm, grp, t);
		for (int t = 0; t < T; t++) group_copy<false>(P, sm, grp, t);
		for (int step = 0; step < P.k; step++) {
			const int i = pass_layer_bit(P, step);
			for (int t = 0; t < P.group_lines / 2; t++) pass_layer<INV>(P, sm, grp, t, i, step == 0, step == P.k - 1);
		}
		for (int t = 0; t < T; t++) group_copy<true>(P, sm, grp, t);
	}
}

// upper passes (b2_upper.cuh): same phase functions as upper_kernel, CTA of 4 warps, one group at a time
template <bool INV>
static void run_upper(const UpperParams& P) {
	std::vector<uint32_t> mem(upper_smem_words(4) + 8);
	UpperSmem sm = upper_smem_view(mem.data());
	const int T = kUpperWarps * 32;
	for (int t = 0; t < T; t++) upper_rows16_fill(sm, t, T);
	for (uint64_t grp = 0; grp < P.n_groups; grp++) {
		const UGroup c = upper_group(P, grp);
		for (int t = 0; t < T; t++) upper_twiddles(P, sm, c, t, P.consts);
		for (int t = 0; t < T; t++) upper_leaves(P, sm, t, T);
		if (P.k >= 4) {
			const int k = P.k, ls = P.ls, upw = 1 << (k - 4), nph = k >> 1;
			for (int lane = 0; lane < 32; lane++) upper_copy<false>(P, sm, c, 0, 0, 1, 1 << k, 0, 1, lane);
			for (int e = 0; e < nph; e++) {
				const int j = INV ? nph - 1 - e : e, b1 = k - 1 - 2 * j, b0 = b1 - 1;
				for (int t = 0; t < T; t++) {
					const int warp = t >> 5, lane = t & 31, u = warp * upw + (lane >> ls);
					const int xbase = ((u >> b0) << (b1 + 1)) | (u & ((1 << b0) - 1));
					upper_pair<INV>(P, sm, lane, xbase, b0, 0, b1, e == 0, e == nph - 1);
				}
			}
			for (int lane = 0; lane < 32; lane++) upper_copy<true>(P, sm, c, 0, 0, 1, 1 << k, 0, 1, lane);
		} else {
			const int w = (int)(grp & 3);  // any warp slot
			for (int t = 0; t < 32; t++) upper_copy<false>(P, sm, c, w * 4, 0, 1, 4, 0, 1, t);
			for (int t = 0; t < 32; t++) upper_pair<INV>(P, sm, t, 0, 0, w * 4, 1, true, true);
			for (int t = 0; t < 32; t++) upper_copy<true>(P, sm, c, w * 4, 0, 1, 4, 0, 1, t);
		}
	}
}

// lowest pass on 512-line tiles (b2_bottom.cuh): same phase functions as bottom_kernel
template <bool INV>
static void run_bottom(const BottomParams& B) {
	std::vector<uint32_t> mem(bottom_smem_words() + 8);
	BottomSmem bs = bottom_smem_view(mem.data());
	const UpperParams& P = B.up;
	const int T = kUpperWarps * 32, nline = 5 - B.batch_bits;
	for (int t = 0; t < T; t++) bottom_prologue(B, bs, t, T, P.consts);
	for (uint64_t grp = 0; grp < P.n_groups; grp++) {
		const UGroup c = upper_group(P, grp);
		for (int t = 0; t < T; t++) upper_twiddles(P, bs.us, c, t, P.consts), bottom_tile_tables(B, bs, c, t - 16, P.consts);
		for (int lane = 0; lane < 32; lane++) upper_copy<false>(P, bs.us, c, 0, 0, 1, 16, 0, 1, lane);
		// as bottom_kernel: no coefficient table when the four phase-A layers are all of class 5
		const bool nt = P.cls[0] == 5 && P.cls[1] == 5 && P.cls[2] == 5 && P.cls[3] == 5;
		if (!nt)
			for (int t = 0; t < T; t++) upper_leaves(P, bs.us, t, T);
		if (!INV) {
			for (int t = 0; t < T; t++) {
				if (nt) upper_pair<false, true>(P, bs.us, t & 31, t >> 5, 2, 0, 3, true, false);
				else upper_pair<false, false>(P, bs.us, t & 31, t >> 5, 2, 0, 3, true, false);
			}
			for (int t = 0; t < T; t++) {
				if (nt) upper_pair<false, true>(P, bs.us, t & 31, (t >> 5) << 2, 0, 0, 1, false, false);
				else upper_pair<false, false>(P, bs.us, t & 31, (t >> 5) << 2, 0, 0, 1, false, false);
			}
		}
		for (int iter = 0; iter < 2; iter++)
			for (int n = 0; n < nline; n++)
				for (int t = 0; t < T; t++) {
					// as launch_bottom: the exchanged in-line form when the plan allows it and the in-line side is compact
					if (B.swap_inline && (INV ? !P.src_sliced : !P.dst_sliced))
						bottom_phase_b<INV, true>(B, bs, t >> 5, t & 31, iter, n);
					else
						bottom_phase_b<INV, false>(B, bs, t >> 5, t & 31, iter, n);
				}
		if (INV) {
			for (int t = 0; t < T; t++) {
				if (nt) upper_pair<true, true>(P, bs.us, t & 31, (t >> 5) << 2, 0, 0, 1, false, false);
				else upper_pair<true, false>(P, bs.us, t & 31, (t >> 5) << 2, 0, 0, 1, false, false);
			}
			for (int t = 0; t < T; t++) {
				if (nt) upper_pair<true, true>(P, bs.us, t & 31, t >> 5, 2, 0, 3, false, true);
				else upper_pair<true, false>(P, bs.us, t & 31, t >> 5, 2, 0, 3, false, true);
			}
		}
		for (int lane = 0; lane < 32; lane++) upper_copy<true>(P, bs.us, c, 0, 0, 1, 16, 0, 1, lane);
	}
}

// block-sliced lowest pass (b2_bsl.cuh): the same phase functions as bsl_kernel, one tile at a time
template <bool INV>
static void run_bsl(const BslParams& P) {
	std::vector<uint32_t> mem(bsl_smem_words() + 8);
	BslSmem sm = bsl_smem_view(mem.data());
	const int T = kBslThreads, S = bsl_stages(P);
	for (int i = 0; i < 1024; i++) sm.consts[i] = P.consts[i];
	for (int t = 0; t < T; t++) bsl_prologue(P, sm, t, T);
	for (uint64_t tile = 0; tile < P.n_tiles; tile++) {
		const BslTile bt = bsl_tile_of(P, tile);
		for (int t = 0; t < T; t++) bsl_copy<false>(P, sm, bt, t), bsl_tile_twiddles(P, sm, bt, t);
		for (int t = 0; t < T; t++) bsl_transpose(sm, t);
		for (int step = 0; step < S; step++) {
			const int s = INV ? step : S - 1 - step;
			if (bsl_has_table(P, s))
				for (int t = 0; t < T; t++) bsl_build_tables(P, sm, s, t);
			for (int t = 0; t < T; t++) bsl_stage<INV>(P, sm, s, t);
		}
		for (int t = 0; t < T; t++) bsl_transpose(sm, t);
		for (int t = 0; t < T; t++) bsl_copy<true>(P, sm, bt, t);
	}
}

extern "C" int emul_antt_sched(int field_bits, int layout, int log_h, int log_rate, int batch, const uint32_t* in,
                               uint32_t* out, int inverse, int coset, int cta_threads, int group_lines, int use_upper) {
	HostPlan hp;
	if (plan_build(hp, field_bits, layout, log_h, log_rate, batch, cta_threads, group_lines, use_upper)) return -1;
	if (hp.tiny) return -4;
	PassParams P;
	UpperParams U;
	BottomParams B;
	const int np = (int)hp.passes.size();
	for (int n = 0; n < np; n++) {
		const int kind = (inverse ? hp.passes[np - 1 - n] : hp.passes[n]).kind;
		if (kind == 2) {
			plan_fill_bottom(hp, B, n, in, out, hp.consts, hp.lane_tw, inverse != 0, coset);
			if (inverse) run_bottom<true>(B); else run_bottom<false>(B);
			continue;
		}
		if (kind == 1) {
			plan_fill_upper(hp, U, n, in, out, hp.consts, inverse != 0, coset);
			if (inverse) run_upper<true>(U); else run_upper<false>(U);
			continue;
		}
		if (kind == 4) {
			BslParams L;
			plan_fill_bsl(hp, L, n, in, out, hp.consts, inverse != 0, coset);
			if (inverse) run_bsl<true>(L); else run_bsl<false>(L);
			continue;
		}
		plan_fill_pass(hp, P, n, in, out, hp.consts, hp.lane_tw, inverse != 0, coset);
		if (inverse) run_pass<true>(P); else run_pass<false>(P);
	}
	return 0;
}
extern "C" int emul_antt_cta(int field_bits, int layout, int log_h, int log_rate, int batch, const uint32_t* in,
                             uint32_t* out, int inverse, int coset, int cta_threads, int group_lines) {
	return emul_antt_sched(field_bits, layout, log_h, log_rate, batch, in, out, inverse, coset, cta_threads, group_lines, 2);
}
extern "C" int emul_antt(int field_bits, int layout, int log_h, int log_rate, int batch, const uint32_t* in,
                         uint32_t* out, int inverse, int coset) {
	return emul_antt_cta(field_bits, layout, log_h, log_rate, batch, in, out, inverse, coset, 0, 0);
}

extern "C" int emul_plan_passes(int field_bits, int layout, int log_h, int log_rate, int* q0, int* k, int* in_line,
                                uint8_t* cls) {
	HostPlan hp;
	if (plan_build(hp, field_bits, layout, log_h, log_rate, 1)) return -1;
	for (size_t i = 0; i < hp.passes.size(); i++) q0[i] = hp.passes[i].q0, k[i] = hp.passes[i].k, in_line[i] = hp.passes[i].in_line;
	memcpy(cls, hp.cls, 32);
	return (int)hp.passes.size();
}

extern "C" void emul_constants(int log_h, int log_rate, uint32_t* out) { plan_constants(log_h, log_rate, out); }

extern "C" uint32_t emul_gf32_mul(uint32_t a, uint32_t b) { return gf32_mul(a, b); }
extern "C" void emul_gf128_mul(const uint32_t* a, const uint32_t* b, uint32_t* o) {
	gf128 x, y;
	memcpy(x.w, a, 16), memcpy(y.w, b, 16);
	gf128 z = gf128_mul(x, y);
	memcpy(o, z.w, 16);
}
extern "C" void emul_sliced_mul(const uint32_t* a, const uint32_t* b, uint32_t* o, int h) {
	switch (h) {
		case 2: Sliced<2>::mul(a, b, o); break;
		case 3: Sliced<3>::mul(a, b, o); break;
		case 4: Sliced<4>::mul(a, b, o); break;
		case 5: Sliced<5>::mul(a, b, o); break;
		case 6: Sliced<6>::mul(a, b, o); break;
		default: Sliced<7>::mul(a, b, o); break;
	}
}
extern "C" void emul_transpose32(uint32_t* a) { transpose32(a); }
