// tools/ubench/uconst.cu -- does the class-5 butterfly run faster when the 0/1 coefficient rows come from CONSTANT memory through
// the uniform datapath?  The twiddle is made provably warp-uniform (shuffle from lane 0), its 27 GF(16) leaf values are computed
// on uniform registers, the rows of the static 16-matrix table are fetched with LDCU and IMAD takes the coefficient as a uniform
// register operand (one vector register read per multiply, no shared-memory traffic for coefficients).
// Compared with butterfly_lane (same static table in shared memory).  Prints JSON lines.
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

__constant__ Coef4 c_rows[64];  // [i * 16 + lambda]
__constant__ uint32_t c_tw[64];

struct UniTwiddle {
	uint32_t y[9];
	__device__ explicit UniTwiddle(uint32_t w) {
		const uint32_t x0 = w & 0x0f0f0f0fu, x1 = (w >> 4) & 0x0f0f0f0fu, x[3] = {x0, x1, x0 ^ x1};
#pragma unroll
		for (int d3 = 0; d3 < 3; d3++) {
			const uint32_t lo = x[d3] & 0x00ff00ffu, hi = (x[d3] >> 8) & 0x00ff00ffu;
			y[3 * d3] = lo, y[3 * d3 + 1] = hi, y[3 * d3 + 2] = lo ^ hi;
		}
	}
	template <int G>
	__device__ uint32_t leaf() const {
		constexpr int d5 = G / 9, d4 = (G / 3) % 3, d3 = G % 3;
		const uint32_t v = y[3 * d3 + d4];
		return d5 == 0 ? (v & 15u) : d5 == 1 ? (v >> 16) : ((v ^ (v >> 16)) & 15u);
	}
};
template <int H, int G0>
struct UniMul {
	static constexpr int N = 1 << H, M = N / 2;
	static constexpr int L = MatMul<H - 1>::L;
	static __device__ void mul(const uint32_t* a, const UniTwiddle& tw, uint32_t* out) {
		uint32_t sa[M], z0[M], z2[M], zz[M], za[M];
#pragma unroll
		for (int i = 0; i < M; i++) sa[i] = a[i] ^ a[M + i];
		UniMul<H - 1, G0>::mul(a, tw, z0);
		UniMul<H - 1, G0 + L>::mul(a + M, tw, z2);
		UniMul<H - 1, G0 + 2 * L>::mul(sa, tw, zz);
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
struct UniMul<2, G0> {
	static __device__ void mul(const uint32_t* a, const UniTwiddle& tw, uint32_t* out) {
		const uint32_t lam = tw.template leaf<G0>();
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const Coef4 q = c_rows[16 * i + lam];
			out[i] = (a[0] * q.x) ^ (a[1] * q.y) ^ (a[2] * q.z) ^ (a[3] * q.w);
		}
	}
};
__device__ __forceinline__ void butterfly_uni(uint32_t* U, uint32_t* V, uint32_t w) {
	const UniTwiddle tw(__shfl_sync(0xffffffffu, w, 0));
	uint32_t p[32];
	UniMul<5, 0>::mul(V, tw, p);
#pragma unroll
	for (int i = 0; i < 32; i++) U[i] ^= p[i];
#pragma unroll
	for (int i = 0; i < 32; i++) V[i] ^= U[i];
}

__device__ __forceinline__ void butterfly_uni2(uint32_t* U, uint32_t* V, uint32_t w) {
	const UniTwiddle tw(w);
	uint32_t p[32];
	UniMul<5, 0>::mul(V, tw, p);
#pragma unroll
	for (int i = 0; i < 32; i++) U[i] ^= p[i];
#pragma unroll
	for (int i = 0; i < 32; i++) V[i] ^= U[i];
}
// FORM 0: butterfly_lane (static rows in shared memory); 1: butterfly_uni (rows in constant memory, uniform datapath)
template <int FORM, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_bfly(uint32_t* io, const uint32_t* tws, int ntw, int iters) {
	__shared__ __align__(16) Coef4 rows[64];
	__shared__ uint32_t stw[64];
	for (int i = threadIdx.x; i < 64; i += THREADS) rows[i] = lane_row_entry(i), stw[i] = tws[i % ntw];
	__syncthreads();
	uint32_t U[32], V[32];
	uint32_t* p = io + ((size_t)blockIdx.x * THREADS + threadIdx.x) * 64;
#pragma unroll
	for (int i = 0; i < 32; i++) U[i] = p[i], V[i] = p[32 + i];
#pragma unroll 1
	for (int it = 0; it < iters; it++) {
		const uint32_t w = stw[(it + (threadIdx.x >> 5)) & 63];  // warp-uniform, from shared memory like the kernels' sm.tw
		if (FORM == 0)
			butterfly_lane<false>(U, V, w, rows);
		else if (FORM == 1)
			butterfly_uni(U, V, w);
		else
			butterfly_uni2(U, V, c_tw[(it + blockIdx.x) & 63]);  // provably uniform: constant table, index from blockIdx and the loop
	}
#pragma unroll
	for (int i = 0; i < 32; i++) p[i] = U[i], p[32 + i] = V[i];
}

template <int FORM, int THREADS, int MINB>
static float run(uint32_t* io, const uint32_t* tws, int ntw, int ctas, int iters, int reps) {
	cudaEvent_t a, b;
	CK(cudaEventCreate(&a));
	CK(cudaEventCreate(&b));
	k_bfly<FORM, THREADS, MINB><<<ctas, THREADS>>>(io, tws, ntw, iters);
	CK(cudaDeviceSynchronize());
	float best = 1e30f;
	for (int r = 0; r < reps; r++) {
		CK(cudaEventRecord(a));
		k_bfly<FORM, THREADS, MINB><<<ctas, THREADS>>>(io, tws, ntw, iters);
		CK(cudaEventRecord(b));
		CK(cudaEventSynchronize(b));
		float ms;
		CK(cudaEventElapsedTime(&ms, a, b));
		if (ms < best) best = ms;
	}
	return best;
}

int main() {
	int sms = 0, khz = 0;
	CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
	CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
	const double ghz = khz * 1e-6;
	Coef4 hrows[64];
	for (int i = 0; i < 64; i++) hrows[i] = lane_row_entry(i);
	CK(cudaMemcpyToSymbol(c_rows, hrows, sizeof hrows));
	{
		uint32_t t[64];
		for (int i = 0; i < 64; i++) t[i] = 0x9e3779b9u * (i + 1) | 0x80000000u;
		CK(cudaMemcpyToSymbol(c_tw, t, sizeof t));
	}
	const int ntw = 64;
	std::vector<uint32_t> tws(ntw);
	for (int i = 0; i < ntw; i++) tws[i] = 0x9e3779b9u * (i + 1) | 0x80000000u;
	uint32_t *d_tws, *io;
	const size_t words = (size_t)sms * 2 * 256 * 64;
	CK(cudaMalloc(&d_tws, ntw * 4));
	CK(cudaMemcpy(d_tws, tws.data(), ntw * 4, cudaMemcpyHostToDevice));
	CK(cudaMalloc(&io, words * 4));
	std::vector<uint32_t> h0(words), h1(words), h2(words);
	for (size_t i = 0; i < words; i++) h0[i] = (uint32_t)(i * 2654435761u) ^ (uint32_t)(i >> 7);
	// parity of the two forms
	CK(cudaMemcpy(io, h0.data(), words * 4, cudaMemcpyHostToDevice));
	k_bfly<0, 128, 2><<<sms * 2, 128>>>(io, d_tws, ntw, 7);
	CK(cudaMemcpy(h1.data(), io, words * 4, cudaMemcpyDeviceToHost));
	CK(cudaMemcpy(io, h0.data(), words * 4, cudaMemcpyHostToDevice));
	k_bfly<1, 128, 2><<<sms * 2, 128>>>(io, d_tws, ntw, 7);
	CK(cudaMemcpy(h2.data(), io, words * 4, cudaMemcpyDeviceToHost));
	printf("{\"bench\": \"uconst_parity\", \"forms_agree\": %s}\n", memcmp(h1.data(), h2.data(), (size_t)sms * 2 * 128 * 64 * 4) == 0 ? "true" : "false");
	const int iters = 400;
	const char* names[3] = {"butterfly_lane (rows in shared memory)", "rows in constant memory, twiddle from a shuffle (compiler emits vector LDC)", "rows in constant memory, provably uniform twiddle (LDCU + uniform-register IMAD operands)"};
	for (int form = 0; form < 3; form++) {
		const float a = form == 0 ? run<0, 128, 2>(io, d_tws, ntw, sms * 2, iters, 5) : form == 1 ? run<1, 128, 2>(io, d_tws, ntw, sms * 2, iters, 5) : run<2, 128, 2>(io, d_tws, ntw, sms * 2, iters, 5);
		const float b = form == 0 ? run<0, 256, 1>(io, d_tws, ntw, sms, iters, 5) : form == 1 ? run<1, 256, 1>(io, d_tws, ntw, sms, iters, 5) : run<2, 256, 1>(io, d_tws, ntw, sms, iters, 5);
		printf("{\"bench\": \"uconst\", \"form\": \"%s\", \"clk_per_warp_butterfly_per_smsp_128x2\": %.1f, \"clk_per_warp_butterfly_per_smsp_256x1\": %.1f}\n", names[form],
		       (a * 1e-3) * ghz * 1e9 / ((double)iters * 2), (b * 1e-3) * ghz * 1e9 / ((double)iters * 2));
	}
	return 0;
}
