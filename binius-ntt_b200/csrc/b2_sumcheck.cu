// b2_sumcheck.cu -- bit-sliced GF(2^128) sumcheck prover for products of C multilinears.  sm_100a only.
//
// Replaces Sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED> (sumcheck/sumcheck.cuh:10-300) with its kernels
// compute_compositions (sumcheck/core/kernels.cuh:5-102) and fold_large_list_halves (sumcheck/core/kernels.cu:9-30) and
// the batch helpers of sumcheck/core/core.cu:9-96.
//
// Data: C columns of 2^num_vars evaluations, bit-sliced in batches of 32 (128 words per batch: word 32k + b = bit b of
// limb k); column stride stays 2^num_vars elements for the whole protocol, round i uses the first 2^(num_vars-i) of each.
//   round kernel: a CTA takes 32 batch rows x of the lower half at a time; for every interpolation point t = 0..C it stages
//     f_c = lo_c ^ t * (lo_c ^ hi_c) (t = 0: lo, t = 1: hi; t >= 2: a GF(16) constant product on each 4-word group),
//     multiplies the C columns with the CTA-level product of b2_mul128.cuh and XORs the 32 product rows into a per-CTA
//     accumulator; one atomicXor per word per CTA at the end.  The claimed sum is points[0] ^ points[1] (the same XOR
//     of all 2^(num_vars-i) products the reference forms separately).  From the second round on p(1) is not computed
//     on the device: the folded multilinears sum to q(r), q the previous round polynomial (degree C, known at C + 1
//     points) and r the challenge just bound, so p(1) = q(r) ^ p(0) exactly -- (C - 1) of the (C + 1)(C - 1) + C
//     products per pair less.
//   fold kernel: a_c[x] <- lo ^ r * (lo ^ hi), r broadcast to all 32 positions.
// With 32 or fewer evaluations left the reference moves to the host (sumcheck.cuh:159-191, fold_small core.cu:57-79); so
// does this file, on the compact values of the last batch (a few hundred field products).
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/b2ntt.h"
#include "b2_field.cuh"
#include "b2_launch.h"
#include "b2_mul128.cuh"

namespace b2 {

constexpr int kScMaxC = 4;

// lo ^ t * (lo ^ hi) on one GF(16) group (4 consecutive bit-sliced words)
__device__ __forceinline__ uint4 sc_point(const uint4 lo, const uint4 hi, int t) {
	if (t == 0) return lo;
	if (t == 1) return hi;
	uint32_t d[4] = {lo.x ^ hi.x, lo.y ^ hi.y, lo.z ^ hi.z, lo.w ^ hi.w}, m[4], p[4];
#pragma unroll
	for (int i = 0; i < 4; i++) m[i] = (t >> i) & 1 ? ~0u : 0u;
	Sliced<2>::mul(d, m, p);
	return make_uint4(lo.x ^ p[0], lo.y ^ p[1], lo.z ^ p[2], lo.w ^ p[3]);
}

// out[(C + 1) * 128] ^= per-point XOR over the batch rows [0, half_rows) of PROD_c f_c
__global__ void __launch_bounds__(kMulThreads, 2) sc_round_kernel(const uint32_t* __restrict__ evals, uint64_t col_words, uint64_t half_rows, int C, int skip1, uint32_t* out) {
	extern __shared__ __align__(16) uint32_t msm[];
	const MulSmem m = mul_smem_view(msm);
	uint32_t* acc = msm + kMulSmemWords;  // [(C + 1)][128]
	for (int i = threadIdx.x; i < (C + 1) * 128; i += kMulThreads) acc[i] = 0;
	for (uint64_t g0 = (uint64_t)blockIdx.x * 32; g0 < half_rows; g0 += (uint64_t)gridDim.x * 32) {
		const int nact = half_rows - g0 < 32 ? (int)(half_rows - g0) : 32;
#pragma unroll 1
		for (int t = 0; t <= C; t++) {
			if (t == 1 && skip1) continue;  // the host derives p(1) from the verifier's claim
#pragma unroll 1
			for (int c = 1; c < C; c++) {
				// stage column c into sb and, for the first product, column 0 into sa
				for (int ch = threadIdx.x; ch < nact * 32; ch += kMulThreads) {
					const int row = ch >> 5, c4 = (ch & 31) * 4;
					const uint32_t* p = evals + (g0 + row) * 128 + c4;
					const uint4 lo = *reinterpret_cast<const uint4*>(p + c * col_words), hi = *reinterpret_cast<const uint4*>(p + c * col_words + half_rows * 128);
					*reinterpret_cast<uint4*>(m.sb + row * kMulPitch + c4) = sc_point(lo, hi, t);
					if (c == 1) {
						const uint4 lo0 = *reinterpret_cast<const uint4*>(p), hi0 = *reinterpret_cast<const uint4*>(p + half_rows * 128);
						*reinterpret_cast<uint4*>(m.sa + row * kMulPitch + c4) = sc_point(lo0, hi0, t);
					}
				}
				__syncthreads();
				cta_mul128(m.sa, m.sb, m.sp, m.sz, nact);
			}
			if (threadIdx.x < 128) {
				uint32_t x = 0;
				for (int row = 0; row < nact; row++) x ^= m.sa[row * kMulPitch + threadIdx.x];
				acc[t * 128 + threadIdx.x] ^= x;
			}
			__syncthreads();
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < (C + 1) * 128; i += kMulThreads)
		if (acc[i]) atomicXor(out + i, acc[i]);
}

// a_c[x] <- lo ^ coeff * (lo ^ hi) for the batch rows x < half_rows of every column; coeff: one bit-sliced batch
__global__ void __launch_bounds__(kMulThreads, 2) sc_fold_kernel(uint32_t* evals, uint64_t col_words, uint64_t half_rows, int C, const uint32_t* __restrict__ coeff) {
	extern __shared__ __align__(16) uint32_t msm[];
	const MulSmem m = mul_smem_view(msm);
	for (int ch = threadIdx.x; ch < 32 * 32; ch += kMulThreads) {
		const int row = ch >> 5, c4 = (ch & 31) * 4;
		*reinterpret_cast<uint4*>(m.sb + row * kMulPitch + c4) = *reinterpret_cast<const uint4*>(coeff + c4);
	}
	const uint64_t groups = (half_rows + 31) / 32;
	for (uint64_t it = blockIdx.x; it < groups * C; it += gridDim.x) {
		const uint64_t c = it / groups, g0 = (it - c * groups) * 32;
		const int nact = half_rows - g0 < 32 ? (int)(half_rows - g0) : 32;
		uint32_t* col = evals + c * col_words;
		for (int ch = threadIdx.x; ch < nact * 32; ch += kMulThreads) {
			const int row = ch >> 5, c4 = (ch & 31) * 4;
			const uint4 lo = *reinterpret_cast<const uint4*>(col + (g0 + row) * 128 + c4), hi = *reinterpret_cast<const uint4*>(col + (g0 + row + half_rows) * 128 + c4);
			*reinterpret_cast<uint4*>(m.sa + row * kMulPitch + c4) = make_uint4(lo.x ^ hi.x, lo.y ^ hi.y, lo.z ^ hi.z, lo.w ^ hi.w);
		}
		__syncthreads();
		cta_mul128(m.sa, m.sb, m.sp, m.sz, nact);
		for (int ch = threadIdx.x; ch < nact * 32; ch += kMulThreads) {
			const int row = ch >> 5, c4 = (ch & 31) * 4;
			uint4* g = reinterpret_cast<uint4*>(col + (g0 + row) * 128 + c4);
			const uint4 lo = *g, p = *reinterpret_cast<const uint4*>(m.sa + row * kMulPitch + c4);
			*g = make_uint4(lo.x ^ p.x, lo.y ^ p.y, lo.z ^ p.z, lo.w ^ p.w);
		}
		__syncthreads();
	}
}

// ---- host side -----------------------------------------------------------------------------------------------------
static void host_unbitslice128(const uint32_t* batch, gf128* out /* 32 */) {
	for (int k = 0; k < 4; k++) {
		uint32_t r[32];
		for (int b = 0; b < 32; b++) r[b] = batch[32 * k + b];
		transpose32(r);
		for (int e = 0; e < 32; e++) out[e].w[k] = r[e];
	}
}
static gf128 gx(const gf128& a, const gf128& b) {
	gf128 r;
	for (int i = 0; i < 4; i++) r.w[i] = a.w[i] ^ b.w[i];
	return r;
}

}  // namespace b2

using namespace b2;

struct b2ntt_sumcheck {
	int num_vars = 0, C = 0, round = 0, device = 0, sms = 0;
	uint64_t n0 = 0;
	uint32_t* d_evals = nullptr;  // C columns, bit-sliced
	bool owns_evals = false;
	uint32_t *d_out = nullptr, *d_coeff = nullptr;
	std::vector<gf128> tail;  // [C][32] compact values once 32 or fewer evaluations are left
	uint32_t last_points[(kScMaxC + 1) * 4] = {0};
	int points_round = -1;    // round whose points are in last_points
	gf128 claim{};            // q(r) of the round before the current one
	bool have_claim = false;
	bool recompute_p1 = false;  // b2ntt_sumcheck_set_recompute
	bool loaded = false;
	uint64_t launches = 0;
};

#define SCU(call)                                                  \
	do {                                                           \
		cudaError_t e_ = (call);                                   \
		if (e_ != cudaSuccess) return report_cuda_error(e_, #call); \
	} while (0)

// Lagrange evaluation at r of the polynomial through (t, points[t]), t = 0..np-1 (what the verifier computes,
// evaluate_univariate_given_points, sumcheck/test/verifier.cu:9-31); the nodes lie in GF(16)
static gf128 sc_interpolate(const uint32_t* points, int np, const gf128& r) {
	gf128 e{};
	for (int i = 0; i < np; i++) {
		gf128 p;
		for (int k = 0; k < 4; k++) p.w[k] = points[4 * i + k];
		for (int j = 0; j < np; j++)
			if (j != i) {
				gf128 f = r, inv{};
				f.w[0] ^= (uint32_t)j;
				inv.w[0] = (uint32_t)Tower<2>::inverse((uint64_t)(i ^ j));
				p = gf128_mul(gf128_mul(p, f), inv);
			}
		e = gx(e, p);
	}
	return e;
}

static int sc_fetch_tail(b2ntt_sumcheck* s) {
	std::vector<uint32_t> batch(128);
	s->tail.assign((size_t)s->C * 32, gf128{});
	for (int c = 0; c < s->C; c++) {
		SCU(cudaMemcpy(batch.data(), s->d_evals + (uint64_t)c * s->n0 * 4, 128 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
		host_unbitslice128(batch.data(), &s->tail[(size_t)c * 32]);
	}
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_create(b2ntt_sumcheck** out, int num_vars, int composition_size, int device) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	if (num_vars < 5 || num_vars > 30 || composition_size < 2 || composition_size > kScMaxC) return B2NTT_EINVAL;
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return report_cuda_error(cudaErrorNoDevice, "b2ntt_sumcheck_create");
	b2ntt_sumcheck* s = new (std::nothrow) b2ntt_sumcheck;
	if (!s) return B2NTT_ENOMEM;
	s->num_vars = num_vars, s->C = composition_size, s->device = device, s->n0 = uint64_t(1) << num_vars;
	DeviceGuard guard(device);
	cudaDeviceGetAttribute(&s->sms, cudaDevAttrMultiProcessorCount, device);
	if (cudaMalloc(&s->d_out, (kScMaxC + 1) * 128 * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&s->d_coeff, 128 * sizeof(uint32_t)) != cudaSuccess) {
		cudaFree(s->d_out);
		delete s;
		return B2NTT_ENOMEM;
	}
	const int smem = (int)((kMulSmemWords + (kScMaxC + 1) * 128) * sizeof(uint32_t));
	cudaFuncSetAttribute(sc_round_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
	cudaFuncSetAttribute(sc_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMulSmemBytes);
	*out = s;
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_destroy(b2ntt_sumcheck* s) {
	if (!s) return B2NTT_OK;
	DeviceGuard guard(s->device);
	if (s->owns_evals) cudaFree(s->d_evals);
	cudaFree(s->d_out), cudaFree(s->d_coeff);
	delete s;
	return B2NTT_OK;
}

static int sc_adopt(b2ntt_sumcheck* s, int transposed) {
	if (!transposed) {
		if (launch_bitslice(s->d_evals, s->n0 / 32 * s->C, 128, 0, nullptr)) return B2NTT_ECUDA;
		s->launches++;
	}
	s->round = 0, s->loaded = true, s->points_round = -1, s->have_claim = false;
	s->tail.clear();
	if (s->n0 == 32) {
		SCU(cudaDeviceSynchronize());
		return sc_fetch_tail(s);
	}
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_load_host(b2ntt_sumcheck* s, const void* h_evals, int transposed) {
	if (!s || !h_evals) return B2NTT_EINVAL;
	DeviceGuard guard(s->device);
	SCU(guard.err);
	const size_t bytes = (size_t)s->C * s->n0 * 16;
	if (!s->owns_evals) {
		s->d_evals = nullptr;
		if (cudaMalloc(&s->d_evals, bytes) != cudaSuccess) return cudaGetLastError(), B2NTT_ENOMEM;
		s->owns_evals = true;
	}
	SCU(cudaMemcpy(s->d_evals, h_evals, bytes, cudaMemcpyHostToDevice));
	return sc_adopt(s, transposed);
}

extern "C" int b2ntt_sumcheck_load_device(b2ntt_sumcheck* s, void* d_evals, int transposed) {
	if (!s || !d_evals) return B2NTT_EINVAL;
	if (pointer_device(d_evals) != s->device) return B2NTT_EINVAL;
	DeviceGuard guard(s->device);
	SCU(guard.err);
	if (s->owns_evals) cudaFree(s->d_evals);
	s->owns_evals = false, s->d_evals = (uint32_t*)d_evals;
	return sc_adopt(s, transposed);
}

extern "C" int b2ntt_sumcheck_round_messages(b2ntt_sumcheck* s, uint32_t* sum, uint32_t* points) {
	if (!s || !sum || !points || !s->loaded || s->round > s->num_vars) return B2NTT_EINVAL;
	const uint64_t n = s->n0 >> s->round;
	const int C = s->C;
	if (n <= 32) {
		const int half = (int)(n / 2);
		gf128 acc{};
		for (int x = 0; x < (int)n; x++) {
			gf128 p = s->tail[x];
			for (int c = 1; c < C; c++) p = gf128_mul(p, s->tail[(size_t)c * 32 + x]);
			acc = gx(acc, p);
		}
		for (int i = 0; i < 4; i++) sum[i] = acc.w[i];
		for (int t = 0; t <= C; t++) {
			gf128 tt{}, pt{};
			tt.w[0] = (uint32_t)t;
			for (int x = 0; x < half; x++) {
				gf128 p{};
				for (int c = 0; c < C; c++) {
					const gf128 lo = s->tail[(size_t)c * 32 + x], hi = s->tail[(size_t)c * 32 + x + half];
					const gf128 f = gx(lo, gf128_mul(gx(lo, hi), tt));
					p = c ? gf128_mul(p, f) : f;
				}
				pt = gx(pt, p);
			}
			for (int i = 0; i < 4; i++) points[4 * t + i] = pt.w[i];
		}
		memcpy(s->last_points, points, (C + 1) * 16), s->points_round = s->round;
		return B2NTT_OK;
	}
	DeviceGuard guard(s->device);
	SCU(guard.err);
	const uint64_t half_rows = n / 64;
	SCU(cudaMemsetAsync(s->d_out, 0, (C + 1) * 128 * sizeof(uint32_t), nullptr));
	uint64_t ctas = (half_rows + 31) / 32;
	if (ctas > (uint64_t)s->sms * 2) ctas = (uint64_t)s->sms * 2;
	const size_t smem = (kMulSmemWords + (C + 1) * 128) * sizeof(uint32_t);
	const int skip1 = (s->have_claim && !s->recompute_p1) ? 1 : 0;
	sc_round_kernel<<<(unsigned)ctas, kMulThreads, smem>>>(s->d_evals, s->n0 * 4, half_rows, C, skip1, s->d_out);
	SCU(cudaGetLastError());
	s->launches++;
	std::vector<uint32_t> h((C + 1) * 128);
	SCU(cudaMemcpy(h.data(), s->d_out, h.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
	// compute_sum (core.cu:81-95): un-bit-slice the batch of partial sums and XOR its 32 elements
	for (int t = 0; t <= C; t++) {
		gf128 e[32], acc{};
		host_unbitslice128(h.data() + 128 * t, e);
		for (int x = 0; x < 32; x++) acc = gx(acc, e[x]);
		for (int i = 0; i < 4; i++) points[4 * t + i] = acc.w[i];
	}
	if (skip1)
		for (int i = 0; i < 4; i++) points[4 + i] = s->claim.w[i] ^ points[i];
	else if (s->have_claim)  // p(1) came from the data: it must agree with the verifier's claim
		for (int i = 0; i < 4; i++)
			if (points[4 + i] != (s->claim.w[i] ^ points[i])) return B2NTT_ECHECK;
	for (int i = 0; i < 4; i++) sum[i] = points[i] ^ points[4 + i];
	memcpy(s->last_points, points, (C + 1) * 16), s->points_round = s->round;
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_set_recompute(b2ntt_sumcheck* s, int on) {
	if (!s) return B2NTT_EINVAL;
	s->recompute_p1 = on != 0;
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_next_round(b2ntt_sumcheck* s, const uint32_t* challenge) {
	if (!s || !challenge || !s->loaded || s->round >= s->num_vars) return B2NTT_EINVAL;
	const uint64_t n = s->n0 >> s->round;
	{
		gf128 r;
		for (int i = 0; i < 4; i++) r.w[i] = challenge[i];
		s->have_claim = s->points_round == s->round;
		if (s->have_claim) s->claim = sc_interpolate(s->last_points, s->C + 1, r);
	}
	if (n <= 32) {
		const int half = (int)(n / 2);
		gf128 r;
		for (int i = 0; i < 4; i++) r.w[i] = challenge[i];
		for (int c = 0; c < s->C; c++)
			for (int x = 0; x < half; x++) {
				gf128& lo = s->tail[(size_t)c * 32 + x];
				lo = gx(lo, gf128_mul(gx(lo, s->tail[(size_t)c * 32 + x + half]), r));
			}
		s->round++;
		return B2NTT_OK;
	}
	DeviceGuard guard(s->device);
	SCU(guard.err);
	// repeat_value_bitsliced (utils/bitslicing.cuh:66-78): word 32k + b = all-ones iff bit b of limb k of the challenge
	uint32_t coeff[128];
	for (int k = 0; k < 4; k++)
		for (int b = 0; b < 32; b++) coeff[32 * k + b] = (challenge[k] >> b) & 1 ? ~0u : 0u;
	SCU(cudaMemcpy(s->d_coeff, coeff, sizeof coeff, cudaMemcpyHostToDevice));
	const uint64_t half_rows = n / 64;
	uint64_t ctas = (half_rows + 31) / 32 * s->C;
	if (ctas > (uint64_t)s->sms * 2) ctas = (uint64_t)s->sms * 2;
	sc_fold_kernel<<<(unsigned)ctas, kMulThreads, kMulSmemBytes>>>(s->d_evals, s->n0 * 4, half_rows, s->C, s->d_coeff);
	SCU(cudaGetLastError());
	s->launches++;
	s->round++;
	if (n / 2 == 32) {
		SCU(cudaDeviceSynchronize());
		return sc_fetch_tail(s);
	}
	return B2NTT_OK;
}

extern "C" int b2ntt_sumcheck_round(const b2ntt_sumcheck* s) { return s ? s->round : B2NTT_EINVAL; }
extern "C" long long b2ntt_sumcheck_launches(const b2ntt_sumcheck* s) { return s ? (long long)s->launches : B2NTT_EINVAL; }

// ================================================================================================================================
// b2ntt_sumcheck_round_static: the round kernel for callers that cannot allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// The only scratch the round needs is the accumulator that sc_round_kernel XORs its per-CTA partial sums into:
// (C + 1) x 128 words.  Here it lives in a file-scope __device__ array sized for the largest composition (C = 4, 2560 bytes;
// every GPU that loads the library holds it until exit).  In HOST code the name `g_round_scratch` is only a stand-in for the
// symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device with that device current, and
// is what the kernel receives.  Device code (sc_points_kernel) names the array directly.
// All calls on one device share the array: it is cleared at the start of every call, and calls that overlap on the GPU
// (different streams, or graph branches) race on it.
namespace b2 {

constexpr int kStaticRoundWords = (kScMaxC + 1) * 128;
static __device__ __align__(256) uint32_t g_round_scratch[kStaticRoundWords];

// compute_sum (sumcheck/core/core.cu:81-95) on the device: limb k of the XOR of the 32 elements of batch t has bit b =
// parity of word 128 t + 32 k + b.  One warp per (t, k); thread 0 of the block then forms sum = p(0) ^ p(1).
__global__ void __launch_bounds__(32 * 4 * (kScMaxC + 1)) sc_points_kernel(int C, uint32_t* __restrict__ points, uint32_t* __restrict__ sum) {
	const int w = threadIdx.x >> 5, b = threadIdx.x & 31;  // w = 4 t + k
	if (w < 4 * (C + 1)) {
		const uint32_t limb = __ballot_sync(0xffffffffu, __popc(g_round_scratch[32 * w + b]) & 1);
		if (b == 0) points[w] = limb;
	}
	__syncthreads();
	if (threadIdx.x < 4) sum[threadIdx.x] = points[threadIdx.x] ^ points[4 + threadIdx.x];
}

namespace {
struct StaticRoundDevice {
	uint32_t* scratch = nullptr;  // device address of g_round_scratch on this device
	int sms = 0;
};
StaticRoundDevice g_static_round[64];
std::mutex g_static_round_mutex;

// once per device, with that device current
int static_round_setup(int device, StaticRoundDevice** out) {
	if (device < 0 || device >= 64) return B2NTT_EINVAL;
	std::lock_guard<std::mutex> lock(g_static_round_mutex);
	StaticRoundDevice& d = g_static_round[device];
	if (!d.scratch) {
		uint32_t* ptr = nullptr;
		int sms = 0;
		SCU(cudaGetSymbolAddress((void**)&ptr, g_round_scratch));
		SCU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
		SCU(cudaFuncSetAttribute(sc_round_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((kMulSmemWords + kStaticRoundWords) * sizeof(uint32_t))));
		d.sms = sms, d.scratch = ptr;
	}
	*out = &d;
	return B2NTT_OK;
}
}  // namespace

}  // namespace b2

extern "C" int b2ntt_sumcheck_round_static(const void* d_evals, int num_vars, int composition_size, void* d_points, void* d_sum, void* stream) {
	if (!d_evals || !d_points || !d_sum) return B2NTT_EINVAL;
	if (composition_size > B2NTT_STATIC_ROUND_MAX_COMPOSITION || num_vars > B2NTT_STATIC_ROUND_MAX_VARS)
		return report_limit_error("b2ntt_sumcheck_round_static: composition_size > 4 (the static scratch holds (4 + 1) x 128 words) or num_vars > 30");
	if (composition_size < 2 || num_vars < 6) return B2NTT_EINVAL;  // 32 or fewer evaluations are host work in every entry point
	const int device = pointer_device(d_evals);
	if (device < 0 || pointer_device(d_points) != device || pointer_device(d_sum) != device) return B2NTT_EINVAL;
	DeviceGuard guard(device);
	SCU(guard.err);
	StaticRoundDevice* dev = nullptr;
	if (int rc = static_round_setup(device, &dev)) return rc;
	const int C = composition_size;
	const uint64_t n0 = uint64_t(1) << num_vars, half_rows = n0 / 64;
	cudaStream_t st = (cudaStream_t)stream;
	SCU(cudaMemsetAsync(dev->scratch, 0, (C + 1) * 128 * sizeof(uint32_t), st));  // the last call's sums are still there
	uint64_t ctas = (half_rows + 31) / 32;
	if (ctas > (uint64_t)dev->sms * 2) ctas = (uint64_t)dev->sms * 2;
	const size_t smem = (kMulSmemWords + (C + 1) * 128) * sizeof(uint32_t);
	sc_round_kernel<<<(unsigned)ctas, kMulThreads, smem, st>>>((const uint32_t*)d_evals, n0 * 4, half_rows, C, 0, dev->scratch);
	SCU(cudaGetLastError());
	sc_points_kernel<<<1, 32 * 4 * (kScMaxC + 1), 0, st>>>(C, (uint32_t*)d_points, (uint32_t*)d_sum);
	SCU(cudaGetLastError());
	return B2NTT_OK;
}
