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

static void fill_params(const HostPlan& hp, PassParams& P, const PassDesc& pd, const uint32_t* src, uint32_t* dst,
                        bool first_pass, bool inverse, int coset) {
	memset(&P, 0, sizeof P);
	const uint64_t row_words = (uint64_t)32 << hp.nb;
	const int rows = inverse ? hp.batch : hp.batch << hp.log_rate;
	P.src = first_pass ? src : dst;
	P.dst = dst;
	P.src_row_words = row_words;
	P.dst_row_words = row_words;
	P.src_row_shift = (first_pass && !inverse) ? hp.log_rate : 0;
	P.coset_mask = inverse ? 0u : ((1u << hp.log_rate) - 1);
	P.consts = hp.consts;
	P.lane_tw = hp.lane_tw;
	P.log_h = hp.log_h, P.log_rate = hp.log_rate, P.limbs_log = hp.limbs_log;
	P.n_inline = hp.n_inline, P.lane_shift = hp.lane_shift, P.batch_bits = hp.batch_bits, P.sliced_io = hp.sliced_io;
	P.nb = hp.nb, P.q0 = pd.q0, P.k = pd.k, P.in_line = pd.in_line;
	P.inverse = inverse ? 1 : 0;
	P.coset0 = inverse ? coset : 0;
	P.n_tiles = (uint64_t)rows << (hp.nb - pd.k);
	memcpy(P.cls, hp.cls, 32);
}

template <bool INV>
static void run_pass(const PassParams& P) {
	const uint64_t tiles_per_cta = kCtaThreads >> (P.k - 1);
	const uint64_t ctas = (P.n_tiles + tiles_per_cta - 1) / tiles_per_cta;
	PassSmem* sm = new PassSmem;
	std::vector<Bundles> st(kCtaThreads);
	for (uint64_t cta = 0; cta < ctas; cta++) {
		for (int t = 0; t < kCtaThreads; t++) pass_prologue(P, *sm, t, kCtaThreads);
		for (int step = 0; step < P.k; step++) {
			const int i = pass_layer_bit(P, step);
			for (int t = 0; t < kCtaThreads; t++)
				pass_layer<INV>(P, *sm, (uint32_t)cta, t, i, step == 0, step == P.k - 1, st[t]);
		}
	}
	delete sm;
}

extern "C" int emul_antt(int field_bits, int layout, int log_h, int log_rate, int batch, const uint32_t* in,
                         uint32_t* out, int inverse, int coset) {
	HostPlan hp;
	if (plan_build(hp, field_bits, layout, log_h, log_rate, batch)) return -1;
	if (hp.tiny) return -4;
	PassParams P;
	const int np = (int)hp.passes.size();
	for (int n = 0; n < np; n++) {
		const PassDesc& pd = inverse ? hp.passes[np - 1 - n] : hp.passes[n];
		fill_params(hp, P, pd, in, out, n == 0, inverse != 0, coset);
		if (inverse) run_pass<true>(P); else run_pass<false>(P);
	}
	return 0;
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
