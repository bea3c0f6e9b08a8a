// tools/ubench/pipes.cu -- pipe-rate and butterfly-circuit microbenchmarks for the B200 SM (sm_100a).
// Measures, in registers only: LOP3 (ALU pipe), IMAD (FMA pipe), both interleaved, and the thread-level
// constant-multiplication butterfly in its two forms (ConstMul: AND masks on the ALU pipe; MatMul: 0/1
// coefficients through IMAD).  Prints one JSON line per measurement.  Build: see tools/ubench/Makefile.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../binius-ntt_b200/csrc/b2_constmat.cuh"
#include "../../binius-ntt_b200/csrc/b2_upper.cuh"

using namespace b2;

#define CK(x)                                                                       \
	do {                                                                            \
		cudaError_t e_ = (x);                                                       \
		if (e_ != cudaSuccess) {                                                    \
			fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
			exit(1);                                                                \
		}                                                                           \
	} while (0)

__device__ __forceinline__ uint32_t lop3_mix(uint32_t a, uint32_t b, uint32_t c) {
	uint32_t r;
	asm("lop3.b32 %0, %1, %2, %3, 0x6a;" : "=r"(r) : "r"(a), "r"(b), "r"(c));  // a ^ (b & c)
	return r;
}
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
	uint32_t r;
	asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
	return r;
}

// MODE 0: LOP3 only; 1: IMAD only; 2: LOP3 and IMAD 1:1; 3: IMAD with a uniform (kernel parameter) multiplier; 4: 2 LOP3 : 1 IMAD;
// 5: two-register XOR + two-register multiply 1:1 (fewer register-file reads); 6: LOP3 (3 registers) + multiply (2 registers) 1:1;
// 7: LOP3 + IMAD with the uniform multiplier 1:1
template <int MODE>
__global__ void __launch_bounds__(256) k_pipes(uint32_t* out, int iters, uint32_t um) {
	uint32_t x[8], y[8];
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
	for (int i = 0; i < 8; i++) x[i] = t * 2654435761u + i, y[i] = t ^ (i * 40503u);
	const uint32_t m = out[t] & 1u, c = t >> 3;  // unknown at compile time
#pragma unroll 1
	for (int it = 0; it < iters; it++) {
#pragma unroll
		for (int r = 0; r < 4; r++) {
#pragma unroll
			for (int i = 0; i < 8; i++) {
				if (MODE == 0) {
					x[i] = lop3_mix(x[i], y[i], c);
					y[i] = lop3_mix(y[i], x[i], m);
				} else if (MODE == 1) {
					x[i] = imad(x[i], m, c);
					y[i] = imad(y[i], m, x[i]);
				} else if (MODE == 2) {
					x[i] = lop3_mix(x[i], y[i], c);
					y[i] = imad(y[i], m, c);
				} else if (MODE == 3) {
					x[i] = imad(x[i], um, c);
					y[i] = imad(y[i], um, x[i]);
				} else if (MODE == 4) {
					x[i] = lop3_mix(x[i], y[i], c);
					y[i] = imad(y[i], m, c);
					x[i] = lop3_mix(x[i], y[i], m);
				} else if (MODE == 5) {
					x[i] = x[i] ^ y[(i + 1) & 7];
					y[i] = b2::sel01(y[i], m);
				} else if (MODE == 6) {
					x[i] = lop3_mix(x[i], y[(i + 1) & 7], c);
					y[i] = b2::sel01(y[i], m);
				} else {
					x[i] = lop3_mix(x[i], y[(i + 1) & 7], c);
					y[i] = imad(y[i], um, c);
				}
			}
		}
	}
	uint32_t s = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) s ^= x[i] ^ y[i];
	out[t] = s;
}

// thread-level butterflies with a warp-uniform twiddle from shared-memory tables; FORM 0: ConstMul<5> masks (GF(4) leaves),
// FORM 1 + NI: MatMul<5, NI> (GF(16) matrix leaves, NI of 4 terms per output bit through IMAD)
template <int FORM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_bfly(uint32_t* io, const uint32_t* tws, int ntw, int iters) {
	extern __shared__ __align__(16) uint32_t smem[];
	Mask4* leaves = reinterpret_cast<Mask4*>(smem);
	Coef4* coefs = reinterpret_cast<Coef4*>(smem);
	if (FORM == 0) {
		for (int item = threadIdx.x; item < ntw * 81; item += THREADS) leaves[item] = leaf_group(tws[item / 81], 5, item % 81);
	} else {
		for (int item = threadIdx.x; item < ntw * 108; item += THREADS) {
			const int tw = item / 108, r = item % 108;
			coefs[item] = coef_row<FORM - 1>(coef_leaf_value(tws[tw], 5, r >> 2), r & 3);
		}
	}
	__syncthreads();
	uint32_t U[32], V[32];
	uint32_t* p = io + ((size_t)blockIdx.x * THREADS + threadIdx.x) * 64;
#pragma unroll
	for (int i = 0; i < 32; i++) U[i] = p[i], V[i] = p[32 + i];
#pragma unroll 1
	for (int it = 0; it < iters; it++) {
		const int tw = it % ntw;
		if (FORM == 0)
			butterfly_const<5, false>(U, V, leaves + tw * 81);
		else
			butterfly_mat<5, false, FORM - 1>(U, V, coefs + tw * 108);
	}
#pragma unroll
	for (int i = 0; i < 32; i++) p[i] = U[i], p[32 + i] = V[i];
}

static float time_ms(void (*launch)(void*), void* ctx, int reps) {
	cudaEvent_t a, b;
	CK(cudaEventCreate(&a));
	CK(cudaEventCreate(&b));
	launch(ctx);
	CK(cudaDeviceSynchronize());
	float best = 1e30f;
	for (int r = 0; r < reps; r++) {
		CK(cudaEventRecord(a));
		launch(ctx);
		CK(cudaEventRecord(b));
		CK(cudaEventSynchronize(b));
		float ms;
		CK(cudaEventElapsedTime(&ms, a, b));
		if (ms < best) best = ms;
	}
	return best;
}

struct PipeCtx {
	int mode, ctas, iters;
	uint32_t* out;
};
static void launch_pipes(void* v) {
	PipeCtx* c = (PipeCtx*)v;
	switch (c->mode) {
		case 0: k_pipes<0><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 1: k_pipes<1><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 2: k_pipes<2><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 3: k_pipes<3><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 4: k_pipes<4><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 5: k_pipes<5><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		case 6: k_pipes<6><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
		default: k_pipes<7><<<c->ctas, 256>>>(c->out, c->iters, 1); break;
	}
}
struct BflyCtx {
	int form, threads, minb, ctas, iters, ntw;
	uint32_t *io, *tws;
};
template <int FORM, int THREADS, int MINB>
static void launch_bfly_t(BflyCtx* c) {
	const size_t smem = (size_t)c->ntw * (FORM == 0 ? 81 : 108) * 16;
	CK(cudaFuncSetAttribute(k_bfly<FORM, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	k_bfly<FORM, THREADS, MINB><<<c->ctas, THREADS, smem>>>(c->io, c->tws, c->ntw, c->iters);
}
template <int FORM>
static void launch_bfly_f(BflyCtx* c) {
	if (c->threads == 128 && c->minb == 2) launch_bfly_t<FORM, 128, 2>(c);
	else if (c->threads == 128 && c->minb == 3) launch_bfly_t<FORM, 128, 3>(c);
	else if (c->threads == 64 && c->minb == 6) launch_bfly_t<FORM, 64, 6>(c);
	else if (c->threads == 128 && c->minb == 1) launch_bfly_t<FORM, 128, 1>(c);
	else launch_bfly_t<FORM, 256, 1>(c);
}
static void launch_bfly(void* v) {
	BflyCtx* c = (BflyCtx*)v;
	switch (c->form) {
		case 0: launch_bfly_f<0>(c); break;
		case 1: launch_bfly_f<1>(c); break;
		case 2: launch_bfly_f<2>(c); break;
		case 3: launch_bfly_f<3>(c); break;
		case 4: launch_bfly_f<4>(c); break;
		default: launch_bfly_f<5>(c); break;
	}
}

int main() {
	cudaDeviceProp pr;
	CK(cudaGetDeviceProperties(&pr, 0));
	const int sms = pr.multiProcessorCount;
	const double ghz = pr.clockRate * 1e-6;
	uint32_t* out;
	CK(cudaMalloc(&out, (size_t)sms * 8 * 256 * 4));
	const char* names[8] = {"lop3", "imad", "lop3+imad 1:1", "imad uniform multiplier", "lop3:imad 2:1", "xor(2 reg)+mul(2 reg) 1:1", "lop3(3 reg)+mul(2 reg) 1:1",
	                        "lop3+imad(uniform multiplier) 1:1"};
	for (int mode = 0; mode < 0; mode++) {
		for (int cps = 4; cps <= 4; cps *= 2) {
			PipeCtx c{mode, sms * cps, 2000, out};
			const float ms = time_ms(launch_pipes, &c, 5);
			const double per_iter = mode == 4 ? 96.0 : 64.0;
			const double ops = (double)c.ctas * 256 * c.iters * per_iter;
			printf("{\"bench\": \"pipes\", \"mode\": \"%s\", \"ctas_per_sm\": %d, \"ms\": %.4f, \"lane_ops_per_clk_per_sm\": %.2f, \"clock_ghz\": %.3f}\n", names[mode], cps, ms,
			       ops / (ms * 1e-3) / (ghz * 1e9) / sms, ghz);
		}
	}
	// butterflies
	const int ntw = 15;
	std::vector<uint32_t> tws(ntw);
	for (int i = 0; i < ntw; i++) tws[i] = 0x9e3779b9u * (i + 1) ^ 0x7f4a7c15u;
	uint32_t* d_tws;
	CK(cudaMalloc(&d_tws, ntw * 4));
	CK(cudaMemcpy(d_tws, tws.data(), ntw * 4, cudaMemcpyHostToDevice));
	const size_t max_threads = (size_t)sms * 4 * 128;
	std::vector<uint32_t> h0(max_threads * 64), h1(max_threads * 64), href(max_threads * 64);
	for (size_t i = 0; i < h0.size(); i++) h0[i] = (uint32_t)(i * 2654435761u) ^ (uint32_t)(i >> 7) * 40503u;
	uint32_t* io;
	CK(cudaMalloc(&io, h0.size() * 4));
	// parity of the two forms against each other and against the scalar tower product (one thread's lanes)
	{
		BflyCtx c{0, 128, 2, sms * 2, 7, ntw, io, d_tws};
		CK(cudaMemcpy(io, h0.data(), h0.size() * 4, cudaMemcpyHostToDevice));
		launch_bfly(&c);
		CK(cudaMemcpy(h1.data(), io, (size_t)c.ctas * 128 * 64 * 4, cudaMemcpyDeviceToHost));
		bool same = true;
		for (int form = 1; form <= 5; form++) {
			c.form = form;
			CK(cudaMemcpy(io, h0.data(), h0.size() * 4, cudaMemcpyHostToDevice));
			launch_bfly(&c);
			CK(cudaMemcpy(href.data(), io, (size_t)c.ctas * 128 * 64 * 4, cudaMemcpyDeviceToHost));
			same = same && memcmp(h1.data(), href.data(), (size_t)c.ctas * 128 * 64 * 4) == 0;
		}
		// scalar check of thread 0, bit position 5: element = bit 5 of the 32 words
		uint32_t u = 0, v = 0;
		for (int i = 0; i < 32; i++) u |= ((h0[i] >> 5) & 1u) << i, v |= ((h0[32 + i] >> 5) & 1u) << i;
		for (int it = 0; it < 7; it++) {
			u ^= gf32_mul(tws[it % ntw], v);
			v ^= u;
		}
		uint32_t gu = 0, gv = 0;
		for (int i = 0; i < 32; i++) gu |= ((href[i] >> 5) & 1u) << i, gv |= ((href[32 + i] >> 5) & 1u) << i;
		printf("{\"bench\": \"bfly_parity\", \"forms_agree\": %s, \"scalar_agrees\": %s}\n", same ? "true" : "false", (gu == u && gv == v) ? "true" : "false");
	}
	const int cfgs[4][2] = {{128, 2}, {256, 1}, {128, 3}, {128, 1}};
	const char* fnames[6] = {"ConstMul GF(4) leaves, LOP3 only", "MatMul 0 IMAD terms (LOP3 only)", "MatMul 1 IMAD term", "MatMul 2 IMAD terms", "MatMul 3 IMAD terms", "MatMul 4 IMAD terms"};
	for (int form = 0; form < 6; form += 5) {
		for (int k = 0; k < 4; k++) {
			BflyCtx c{form, cfgs[k][0], cfgs[k][1], sms * cfgs[k][1], 400, ntw, io, d_tws};
			const float ms = time_ms(launch_bfly, &c, 5);
			const double bf = (double)c.ctas * c.threads * c.iters;
			printf("{\"bench\": \"bfly\", \"form\": \"%s\", \"threads\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"thread_butterflies_per_s\": %.4e, \"clk_per_warp_butterfly_per_smsp\": %.1f}\n",
			       fnames[form], c.threads, cfgs[k][1], ms, bf / (ms * 1e-3), (ms * 1e-3) * ghz * 1e9 / ((double)c.iters * c.threads * cfgs[k][1] / 128.0));
		}
	}
	return 0;
}
