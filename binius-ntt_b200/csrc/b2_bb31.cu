// b2_bb31.cu -- BabyBear (p = 15 * 2^27 + 1) radix-2 NTT, the prime-field transform that shares `ntt_tests` with the additive NTT
// (SURVEY 8f.4).  Same function as the reference's NTT<BB31> (ntt/gpuntt.cuh:126-209: bit-reverse the input, then stages
// s = 0 .. L-1 of  U = u + v ; V = (u - v) * tw_br[t >> s]  with u = (t mod 2^s) | (t >> s) << (s+1)), different shape:
//   * words are risc0::Fp raw values (Montgomery form, R = 2^32: finite_fields/risc0_baby_bear.h:163-177), so host arrays of
//     BB31 pass through unconverted; add and subtract are the same in both forms, the product is a Montgomery product;
//   * the bit-reversed twiddle table is built on the device from the L-1 squarings of the root (no host loop, no upload);
//   * stages run in fused passes on shared-memory tiles: the first pass takes 11 stages on 2048 consecutive elements, the
//     following ones 6 stages on 64 rows x 32 consecutive elements (128-byte runs: every global access is a full line; the
//     reference's kernels 1 and 2 read one 4-byte element per 8 KB / 16 MB stride, gpuntt.cuh:79-97).
// sm_100a only.  HBM-bound integer work: no tensor cores.
#include <cuda_runtime.h>
#include <stdint.h>

#include <new>

#include "../../include/b2ntt.h"
#include "b2_launch.h"

namespace b2 {

constexpr uint32_t kBbP = 15u * (1u << 27) + 1;
constexpr uint32_t kBbNinv = 0x77FFFFFFu;  // -p^-1 mod 2^32
constexpr uint32_t kBbR2 = 1172168163u;    // 2^64 mod p
constexpr int kBbThreads = 256;
constexpr int kBbFirstBits = 11;           // stages of the first pass (2048-element tiles)
constexpr int kBbRowBits = 6, kBbRunBits = 5;  // later passes: 2^6 rows x 2^5 consecutive elements

__host__ __device__ __forceinline__ uint32_t bb_add(uint32_t a, uint32_t b) {
	const uint32_t r = a + b;
	return r >= kBbP ? r - kBbP : r;
}
__host__ __device__ __forceinline__ uint32_t bb_sub(uint32_t a, uint32_t b) { return a >= b ? a - b : a + kBbP - b; }
// Montgomery product a * b / 2^32 mod p, operands and result in [0, p)
__host__ __device__ __forceinline__ uint32_t bb_mul(uint32_t a, uint32_t b) {
	const uint64_t t = (uint64_t)a * b;
	const uint32_t q = (uint32_t)t * kBbNinv;
	const uint32_t r = (uint32_t)((t + (uint64_t)q * kBbP) >> 32);
	return r >= kBbP ? r - kBbP : r;
}

struct BbRoots {
	uint32_t sq[27];  // sq[b] = w^(2^b), Montgomery form
};

// tw_br[i] = w^(bit reversal of i over `bits` bits): a product over the set bits of the exponent
__global__ void __launch_bounds__(kBbThreads) bb31_twiddle_kernel(uint32_t* tw, uint32_t count, int bits, const __grid_constant__ BbRoots R, uint32_t one) {
	for (uint32_t i = blockIdx.x * kBbThreads + threadIdx.x; i < count; i += gridDim.x * kBbThreads) {
		const uint32_t e = bits ? __brev(i) >> (32 - bits) : 0u;
		uint32_t v = one;
		for (int b = 0; b < bits; b++)
			if ((e >> b) & 1u) v = bb_mul(v, R.sq[b]);
		tw[i] = v;
	}
}

// out[bit reversal of i] = in[i], reduced into [0, p) on the way (raw words above p - 1 never come from BB31, but cost nothing to fold)
__global__ void __launch_bounds__(kBbThreads) bb31_bitrev_kernel(const uint32_t* in, uint32_t* out, uint32_t n, int bits, int reverse) {
	for (uint32_t i = blockIdx.x * kBbThreads + threadIdx.x; i < n; i += gridDim.x * kBbThreads) {
		uint32_t v = in[i];
		v = v >= kBbP ? v - kBbP : v;
		out[reverse ? __brev(i) >> (32 - bits) : i] = v;
	}
}

// stages s0 .. s0+k-1 on tiles of 2^(k+c) elements: element (x, lo) of a tile is base + (x << s0) + lo, lo < 2^c consecutive
__global__ void __launch_bounds__(kBbThreads) bb31_pass_kernel(uint32_t* data, const uint32_t* __restrict__ tw, int log_n, int s0, int k, int c) {
	__shared__ uint32_t tile[1 << kBbFirstBits];
	const int tbits = k + c;
	const uint32_t tsize = 1u << tbits, ntiles = 1u << (log_n - tbits);
	for (uint32_t tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
		// tile number = (hi, lo_hi): lo_hi counts the 2^c-element runs below bit s0
		const uint32_t lo_hi = tl & ((1u << (s0 - c)) - 1), hi = tl >> (s0 - c);
		const uint32_t base = (hi << (s0 + k)) | (lo_hi << c);
		for (uint32_t m = threadIdx.x; m < tsize; m += kBbThreads) tile[m] = data[base + ((m >> c) << s0) + (m & ((1u << c) - 1))];
		__syncthreads();
		for (int j = 0; j < k; j++) {
			for (uint32_t b = threadIdx.x; b < tsize / 2; b += kBbThreads) {
				const uint32_t lo = b & ((1u << c) - 1), r = b >> c;
				const uint32_t xu = (r & ((1u << j) - 1)) | ((r >> j) << (j + 1));
				const uint32_t mu = (xu << c) | lo, mv = mu | (1u << (j + c));
				const uint32_t eu = base + (xu << s0) + lo;
				const uint32_t w = tw[eu >> (s0 + j + 1)];
				const uint32_t u = tile[mu], v = tile[mv];
				tile[mu] = bb_add(u, v);
				tile[mv] = bb_mul(bb_sub(u, v), w);
			}
			__syncthreads();
		}
		for (uint32_t m = threadIdx.x; m < tsize; m += kBbThreads) data[base + ((m >> c) << s0) + (m & ((1u << c) - 1))] = tile[m];
		__syncthreads();
	}
}

}  // namespace b2

using namespace b2;

struct b2ntt_bb31 {
	int device = 0, log_n = 0, sms = 0;
	uint32_t* d_tw = nullptr;    // [2^(log_n-1)] bit-reversed twiddles
	uint32_t* d_work = nullptr;  // exec_host staging: input image
	uint32_t* d_data = nullptr;  // exec_host staging: transform buffer
	cudaStream_t stream = nullptr;
};

static uint32_t bb_pow(uint32_t x_mont, uint64_t n, uint32_t one) {
	uint32_t r = one;
	while (n) {
		if (n & 1) r = bb_mul(r, x_mont);
		x_mont = bb_mul(x_mont, x_mont);
		n >>= 1;
	}
	return r;
}

extern "C" int b2ntt_bb31_create(b2ntt_bb31** out, uint32_t generator_raw, int log_group_order, int log_n, int device) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	// NTTConfRad2 (ntt/nttconf.cuh:32-39): 1 <= log_n <= 27, log_group_order >= log_n
	if (log_n < 1 || log_n > 27 || log_group_order < log_n || log_group_order > 31 || generator_raw >= kBbP) return B2NTT_EINVAL;
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return report_cuda_error(cudaErrorNoDevice, "b2ntt_bb31_create");
	if (device < 0 || device >= ndev) return B2NTT_EINVAL;
	b2ntt_bb31* h = new (std::nothrow) b2ntt_bb31;
	if (!h) return B2NTT_ENOMEM;
	h->device = device, h->log_n = log_n;
	DeviceGuard guard(device);
	cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, device);
	const uint32_t half = 1u << (log_n - 1), one = bb_mul(1, kBbR2);
	if (cudaMalloc(&h->d_tw, sizeof(uint32_t) * half) != cudaSuccess) {
		cudaGetLastError();
		delete h;
		return B2NTT_ENOMEM;
	}
	BbRoots R;
	R.sq[0] = bb_pow(generator_raw, (uint64_t)1 << (log_group_order - log_n), one);  // gpuntt.cuh:195-196
	for (int b = 1; b < 27; b++) R.sq[b] = bb_mul(R.sq[b - 1], R.sq[b - 1]);
	const unsigned grid = (unsigned)((half + kBbThreads - 1) / kBbThreads < (uint32_t)h->sms * 8 ? (half + kBbThreads - 1) / kBbThreads : h->sms * 8);
	bb31_twiddle_kernel<<<grid, kBbThreads>>>(h->d_tw, half, log_n - 1, R, one);
	cudaError_t e = cudaDeviceSynchronize();
	if (e != cudaSuccess) {
		cudaFree(h->d_tw);
		delete h;
		return report_cuda_error(e, "b2ntt_bb31_create");
	}
	*out = h;
	return B2NTT_OK;
}

extern "C" int b2ntt_bb31_destroy(b2ntt_bb31* h) {
	if (!h) return B2NTT_OK;
	DeviceGuard guard(h->device);
	cudaFree(h->d_tw);
	if (h->d_work) cudaFree(h->d_work);
	if (h->d_data) cudaFree(h->d_data);
	if (h->stream) cudaStreamDestroy(h->stream);
	delete h;
	return B2NTT_OK;
}

extern "C" int b2ntt_bb31_launches(const b2ntt_bb31* h) {
	if (!h) return B2NTT_EINVAL;
	int passes = 1, s = h->log_n < kBbFirstBits ? h->log_n : kBbFirstBits;
	for (; s < h->log_n; s += kBbRowBits) passes++;
	return 1 + passes;
}

extern "C" int b2ntt_bb31_exec(b2ntt_bb31* h, const void* d_in, void* d_out, int input_bit_reversed, void* stream) {
	if (!h || !d_in || !d_out) return B2NTT_EINVAL;
	if (d_in == d_out && !input_bit_reversed) return B2NTT_EINVAL;  // the permutation is out of place
	if (pointer_device(d_in) != h->device || pointer_device(d_out) != h->device) return B2NTT_EINVAL;
	DeviceGuard guard(h->device);
	if (guard.err != cudaSuccess) return report_cuda_error(guard.err, "b2ntt_bb31_exec");
	cudaStream_t st = (cudaStream_t)stream;
	const int L = h->log_n;
	const uint32_t n = 1u << L;
	const unsigned cap = (unsigned)h->sms * 8;
	unsigned grid = (n + kBbThreads - 1) / kBbThreads;
	if (d_in != d_out || !input_bit_reversed) bb31_bitrev_kernel<<<grid < cap ? grid : cap, kBbThreads, 0, st>>>((const uint32_t*)d_in, (uint32_t*)d_out, n, L, input_bit_reversed ? 0 : 1);
	int s0 = 0;
	while (s0 < L) {
		const int k = s0 == 0 ? (L < kBbFirstBits ? L : kBbFirstBits) : (L - s0 < kBbRowBits ? L - s0 : kBbRowBits);
		const int c = s0 == 0 ? 0 : kBbRunBits;
		const uint32_t ntiles = 1u << (L - k - c);
		bb31_pass_kernel<<<ntiles < cap ? ntiles : cap, kBbThreads, 0, st>>>((uint32_t*)d_out, h->d_tw, L, s0, k, c);
		s0 += k;
	}
	cudaError_t e = cudaGetLastError();
	return e == cudaSuccess ? B2NTT_OK : report_cuda_error(e, "b2ntt_bb31_exec");
}

// NTT<E>::apply (ntt/gpuntt.cuh:158-183): host in, host out, synchronous
extern "C" int b2ntt_bb31_exec_host(b2ntt_bb31* h, const void* h_in, size_t n, int input_bit_reversed, void* h_out) {
	if (!h || !h_in || !h_out || n != ((size_t)1 << h->log_n)) return B2NTT_EINVAL;
	DeviceGuard guard(h->device);
	if (guard.err != cudaSuccess) return report_cuda_error(guard.err, "b2ntt_bb31_exec_host");
	const size_t bytes = n * sizeof(uint32_t);
	cudaError_t e = cudaSuccess;
	if (!h->stream) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
	if (e == cudaSuccess && !h->d_work) e = cudaMalloc(&h->d_work, bytes);
	if (e == cudaSuccess && !h->d_data) e = cudaMalloc(&h->d_data, bytes);
	if (e != cudaSuccess) return e == cudaErrorMemoryAllocation ? (cudaGetLastError(), B2NTT_ENOMEM) : report_cuda_error(e, "b2ntt_bb31_exec_host");
	e = cudaMemcpyAsync(h->d_work, h_in, bytes, cudaMemcpyHostToDevice, h->stream);
	if (e != cudaSuccess) return report_cuda_error(e, "b2ntt_bb31_exec_host H2D");
	const int rc = b2ntt_bb31_exec(h, h->d_work, h->d_data, input_bit_reversed, h->stream);
	if (rc) return rc;
	e = cudaMemcpyAsync(h_out, h->d_data, bytes, cudaMemcpyDeviceToHost, h->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
	return e == cudaSuccess ? B2NTT_OK : report_cuda_error(e, "b2ntt_bb31_exec_host D2H");
}
