// b2_general.cu -- additive NTT over GF(2^128) on a caller-supplied basis.  sm_100a only.
//
// The reference's transform (ntt/additive_ntt.cuh:91-160, :273-309) fixes the basis beta_i = 1 << i, so every twiddle
// lies in GF(2^32) and log_h + log_rate <= 32.  Here beta_0 .. beta_{log_h-1} are arbitrary GF(2^128) elements, linearly
// independent over GF(2): the domain constants C[i][j] = W^_i(beta_{i+1+j}) are full-width (same recurrence as
// precompute_subspace_evals with norm_0 = beta_0 instead of 1) and the butterfly u ^= w * v; v ^= u needs a true
// GF(2^128) product (b2::gf128_mul, the tower of sumcheck/test/utils/tower_7_mul.cu:4-20).
//
// From log_h = 11 up: one launch per two layers, a CTA takes 1024 butterflies per layer and iteration: it stages the V operands and the
// 1024 twiddles (compact), bit-slices both in shared memory, multiplies with the CTA-level product of b2_mul128.cuh (the
// one behind b2ntt_mul and the sumcheck kernels), transposes the products back and applies the XORs on the way out.
// Below that size: one butterfly per thread with the scalar table-free product.  The default basis reproduces b2ntt_exec
// bit for bit (tests), which is the faster route whenever the basis is 1 << i.
#include <cuda_runtime.h>

#include <new>
#include <vector>

#include "../../include/b2ntt.h"
#include "b2_field.cuh"
#include "b2_launch.h"
#include "b2_mul128.cuh"

using namespace b2;

namespace {

gf128 g_xor(const gf128& a, const gf128& b) {
	gf128 r;
	for (int l = 0; l < 4; l++) r.w[l] = a.w[l] ^ b.w[l];
	return r;
}
bool g_zero(const gf128& a) { return !(a.w[0] | a.w[1] | a.w[2] | a.w[3]); }
// inverse through the norm to GF(2^64) (finite_fields/binary_tower.cuh:70-80 one level up); inverse(0) = 0
gf128 g_inverse(const gf128& a) {
	const uint64_t a0 = a.w[0] | (uint64_t)a.w[1] << 32, a1 = a.w[2] | (uint64_t)a.w[3] << 32;
	const uint64_t inter = a0 ^ Tower<6>::mul_alpha(a1);
	const uint64_t delta = gf64_mul(a0, inter) ^ Tower<6>::square(a1);
	const uint64_t dinv = Tower<6>::inverse(delta);
	const uint64_t lo = gf64_mul(dinv, inter), hi = gf64_mul(dinv, a1);
	gf128 r;
	r.w[0] = (uint32_t)lo, r.w[1] = (uint32_t)(lo >> 32), r.w[2] = (uint32_t)hi, r.w[3] = (uint32_t)(hi >> 32);
	return r;
}

constexpr int kMaxLog = 34;  // 2^34 elements = 256 GiB: beyond one B200, the bound only sizes the shared table

// one layer: butterflies t = (block j, offset k), u = j << (layer + 1) | k, v = u | 1 << layer (additive_ntt.cuh:21-33);
// twiddle = XOR of row[b] over the set bits b of j (calculate_twiddle, :59-77, with full-width constants)
template <bool INV>
__global__ void __launch_bounds__(256) general_layer_kernel(uint4* __restrict__ data, const uint4* __restrict__ row, int layer, int nbits, uint64_t n_half) {
	__shared__ uint4 sc[kMaxLog];
	if ((int)threadIdx.x < nbits) sc[threadIdx.x] = row[threadIdx.x];
	__syncthreads();
	const uint64_t step = (uint64_t)gridDim.x * blockDim.x, kmask = (uint64_t(1) << layer) - 1;
	for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n_half; t += step) {
		const uint64_t j = t >> layer, ui = (j << (layer + 1)) | (t & kmask), vi = ui | (uint64_t(1) << layer);
		gf128 w = {{0, 0, 0, 0}};
		for (int b = 0; b < nbits; b++)
			if ((j >> b) & 1) {
				const uint4 c = sc[b];
				w.w[0] ^= c.x, w.w[1] ^= c.y, w.w[2] ^= c.z, w.w[3] ^= c.w;
			}
		uint4 u = data[ui], v = data[vi];
		if (INV) v.x ^= u.x, v.y ^= u.y, v.z ^= u.z, v.w ^= u.w;
		const gf128 vv = {{v.x, v.y, v.z, v.w}};
		const gf128 p = gf128_mul(w, vv);
		u.x ^= p.w[0], u.y ^= p.w[1], u.z ^= p.w[2], u.w ^= p.w[3];
		if (!INV) v.x ^= u.x, v.y ^= u.y, v.z ^= u.z, v.w ^= u.w;
		data[ui] = u, data[vi] = v;
	}
}

// 1024 butterflies per CTA iteration and layer (n_half is a multiple of 1024).  Butterfly ch of the iteration is staged as
// 16-byte element (ch & 31) of block row ch >> 5, the image mul_sliced128_kernel<true> (b2_misc.cu) works on.
// nl = 1: one layer `hi`, butterflies t = 1024 * it + ch.  nl = 2: layers hi and hi - 1 on 512 quads of elements that differ
// in index bits hi and hi - 1 (quad q = 512 * it + (ch & 511), pair ch >> 9 of the quad): the second layer reads what the
// same CTA wrote a moment earlier (L2/L1, made visible by the CTA barrier), so two layers cost one HBM round trip.
template <bool INV>
__global__ void __launch_bounds__(kMulThreads, 2) general_layer_cta_kernel(uint4* __restrict__ data, const uint4* __restrict__ rows, int log_h, int hi, int nl, uint64_t n_half) {
	extern __shared__ __align__(16) uint32_t gsm[];
	const MulSmem m = mul_smem_view(gsm);
	uint32_t *sa = m.sa, *sb = m.sb;
	uint4* sc = reinterpret_cast<uint4*>(gsm + kMulSmemWords);  // [2][kMaxLog]: constants of layer hi, of layer hi - 1
	const int j = threadIdx.x >> 5, lb = threadIdx.x & 31;
	if ((int)threadIdx.x < log_h) {
		sc[threadIdx.x] = rows[(size_t)hi * log_h + threadIdx.x];
		if (nl == 2) sc[kMaxLog + threadIdx.x] = rows[(size_t)(hi - 1) * log_h + threadIdx.x];
	}
	__syncthreads();
	const uint64_t n_iter = n_half / 1024;  // nl = 2: n / 4 quads, 512 per iteration
	for (uint64_t it = blockIdx.x; it < n_iter; it += gridDim.x) {
#pragma unroll 1
		for (int ph = 0; ph < nl; ph++) {
			// layer of this phase: forward runs hi first, the inverse hi - 1 first
			const int layer = nl == 1 ? hi : ((ph == 0) != INV ? hi : hi - 1);
			const int nbits = log_h - 1 - layer;
			const uint4* tab = sc + (layer == hi ? 0 : kMaxLog);
			const uint64_t kmask = (uint64_t(1) << layer) - 1, lowmask = nl == 2 ? (uint64_t(1) << (hi - 1)) - 1 : 0;
			auto pair = [&](int ch, uint64_t& ui, uint64_t& vi) {
				if (nl == 1) {
					const uint64_t t = it * 1024 + ch;
					ui = ((t >> layer) << (layer + 1)) | (t & kmask);
				} else {
					const uint64_t q = it * 512 + (ch & 511), e0 = ((q >> (hi - 1)) << (hi + 1)) | (q & lowmask);
					ui = e0 | ((uint64_t)(ch >> 9) << (layer == hi ? hi - 1 : hi));
				}
				vi = ui | (uint64_t(1) << layer);
			};
			for (int ch = threadIdx.x; ch < 1024; ch += kMulThreads) {
				uint64_t ui, vi;
				pair(ch, ui, vi);
				const uint64_t blk = ui >> (layer + 1);
				uint4 v = data[vi];
				if (INV) {
					const uint4 u = data[ui];
					v.x ^= u.x, v.y ^= u.y, v.z ^= u.z, v.w ^= u.w;
				}
				uint4 w = make_uint4(0, 0, 0, 0);
				for (int b = 0; b < nbits; b++)
					if ((blk >> b) & 1) {
						const uint4 c = tab[b];
						w.x ^= c.x, w.y ^= c.y, w.z ^= c.z, w.w ^= c.w;
					}
				const int off = (ch >> 5) * kMulPitch + (ch & 31) * 4;
				*reinterpret_cast<uint4*>(sa + off) = v;
				*reinterpret_cast<uint4*>(sb + off) = w;
			}
			__syncthreads();
			if (j < 8) {  // compact -> bit-sliced, both operands (lane = 4 * block + limb)
				uint32_t r[32];
				uint32_t* rw = (j < 4 ? sa : sb) + (8 * (j & 3) + (lb >> 2)) * kMulPitch;
				const int k = lb & 3;
#pragma unroll
				for (int e = 0; e < 32; e++) r[e] = rw[4 * e + k];
				transpose32(r);
				__syncwarp();
#pragma unroll
				for (int i = 0; i < 32; i++) rw[32 * k + i] = r[i];
			}
			__syncthreads();
			cta_mul128(sa, sb, m.sp, m.sz, 32);
			if (j < 4) {  // products back to compact
				uint32_t r[32];
				uint32_t* rw = sa + (8 * j + (lb >> 2)) * kMulPitch;
				const int k = lb & 3;
#pragma unroll
				for (int i = 0; i < 32; i++) r[i] = rw[32 * k + i];
				transpose32(r);
				__syncwarp();
#pragma unroll
				for (int e = 0; e < 32; e++) rw[4 * e + k] = r[e];
			}
			__syncthreads();
			for (int ch = threadIdx.x; ch < 1024; ch += kMulThreads) {
				uint64_t ui, vi;
				pair(ch, ui, vi);
				const uint4 p = *reinterpret_cast<const uint4*>(sa + (ch >> 5) * kMulPitch + (ch & 31) * 4);
				uint4 u = data[ui], v = data[vi];  // second read of the pair: L2-resident, this CTA fetched it a few microseconds ago
				if (INV) v.x ^= u.x, v.y ^= u.y, v.z ^= u.z, v.w ^= u.w;
				u.x ^= p.x, u.y ^= p.y, u.z ^= p.z, u.w ^= p.w;
				if (!INV) v.x ^= u.x, v.y ^= u.y, v.z ^= u.z, v.w ^= u.w;
				data[ui] = u, data[vi] = v;
			}
			__syncthreads();  // also publishes this phase's stores to the CTA's next phase
		}
	}
}

constexpr size_t kGeneralSmemBytes = kMulSmemBytes + 2 * kMaxLog * sizeof(uint4);

}  // namespace

struct b2ntt_general {
	int log_h = 0, device = 0, sms = 0;
	std::vector<gf128> C;  // [log_h][log_h], row i holds log_h - 1 - i constants
	uint4* d_C = nullptr;
};

extern "C" int b2ntt_general_create(b2ntt_general** out, int log_h, const uint32_t* h_basis, int device) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	if (log_h < 1 || log_h > kMaxLog) return B2NTT_EINVAL;
	if (!h_basis && log_h > 32) return B2NTT_EINVAL;  // the default basis 1 << i has 32 independent elements in limb 0
	// unnormalised E_i(beta_k), k >= i: E_0(x) = x, E_i(x) = E_{i-1}(x)^2 + E_{i-1}(beta_{i-1}) * E_{i-1}(x)
	std::vector<gf128> cur(log_h), C((size_t)log_h * log_h, gf128{{0, 0, 0, 0}});
	for (int k = 0; k < log_h; k++) {
		if (h_basis)
			for (int l = 0; l < 4; l++) cur[k].w[l] = h_basis[4 * k + l];
		else
			cur[k] = gf128{{uint32_t(1) << k, 0, 0, 0}};
	}
	for (int i = 0; i < log_h; i++) {
		if (i > 0) {
			const gf128 np = cur[i - 1];
			for (int k = i; k < log_h; k++) cur[k] = g_xor(gf128_mul(cur[k], cur[k]), gf128_mul(np, cur[k]));
		}
		if (g_zero(cur[i])) return B2NTT_EINVAL;  // beta_i lies in the span of beta_0..beta_{i-1}
		const gf128 inv = g_inverse(cur[i]);
		for (int k = i + 1; k < log_h; k++) C[(size_t)i * log_h + (k - i - 1)] = gf128_mul(inv, cur[k]);
	}
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return report_cuda_error(cudaErrorNoDevice, "b2ntt_general_create");
	b2ntt_general* g = new (std::nothrow) b2ntt_general;
	if (!g) return B2NTT_ENOMEM;
	g->log_h = log_h, g->device = device, g->C.swap(C);
	DeviceGuard guard(device);
	cudaDeviceGetAttribute(&g->sms, cudaDevAttrMultiProcessorCount, device);
	cudaFuncSetAttribute(general_layer_cta_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGeneralSmemBytes);
	cudaFuncSetAttribute(general_layer_cta_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGeneralSmemBytes);
	const size_t bytes = g->C.size() * sizeof(gf128);
	if (cudaMalloc(&g->d_C, bytes) != cudaSuccess) {
		delete g;
		return B2NTT_ENOMEM;
	}
	cudaError_t e = cudaMemcpy(g->d_C, g->C.data(), bytes, cudaMemcpyHostToDevice);
	if (e != cudaSuccess) {
		cudaFree(g->d_C);
		delete g;
		return report_cuda_error(e, "b2ntt_general_create");
	}
	*out = g;
	return B2NTT_OK;
}

extern "C" int b2ntt_general_destroy(b2ntt_general* g) {
	if (!g) return B2NTT_OK;
	DeviceGuard guard(g->device);
	cudaFree(g->d_C);
	delete g;
	return B2NTT_OK;
}

extern "C" int b2ntt_general_constants(const b2ntt_general* g, uint32_t* h_out) {
	if (!g || !h_out) return B2NTT_EINVAL;
	for (size_t i = 0; i < g->C.size(); i++)
		for (int l = 0; l < 4; l++) h_out[4 * i + l] = g->C[i].w[l];
	return B2NTT_OK;
}

extern "C" int b2ntt_general_exec(b2ntt_general* g, const void* d_in, void* d_out, int inverse, void* stream) {
	if (!g || !d_in || !d_out) return B2NTT_EINVAL;
	cudaStream_t st = (cudaStream_t)stream;
	if (pointer_device(d_in) != g->device || pointer_device(d_out) != g->device) return B2NTT_EINVAL;
	DeviceGuard guard(g->device);
	const uint64_t n = uint64_t(1) << g->log_h, n_half = n >> 1;
	if (d_in != d_out) {
		cudaError_t e = cudaMemcpyAsync(d_out, d_in, n * sizeof(gf128), cudaMemcpyDeviceToDevice, st);
		if (e != cudaSuccess) return report_cuda_error(e, "b2ntt_general_exec copy");
	}
	if (g->log_h >= 11) {
		// layers in pairs (hi, hi - 1), one launch each; an odd layer count leaves one single-layer launch
		const uint64_t cap = (uint64_t)g->sms * 2;
		int done = 0;
		while (done < g->log_h) {
			const int nl = g->log_h - done >= 2 ? 2 : 1;
			// forward walks the strides down (additive_ntt.cuh:233-247), the inverse up
			const int hi = inverse ? done + nl - 1 : g->log_h - 1 - done;
			const uint64_t want = n_half / 1024;
			const unsigned grid = (unsigned)(want < cap ? want : cap);
			if (inverse)
				general_layer_cta_kernel<true><<<grid, kMulThreads, kGeneralSmemBytes, st>>>((uint4*)d_out, g->d_C, g->log_h, hi, nl, n_half);
			else
				general_layer_cta_kernel<false><<<grid, kMulThreads, kGeneralSmemBytes, st>>>((uint4*)d_out, g->d_C, g->log_h, hi, nl, n_half);
			done += nl;
		}
	} else {
		const uint64_t want = (n_half + 255) / 256, cap = (uint64_t)g->sms * 8 * 4;
		const unsigned grid = (unsigned)(want < cap ? want : cap);
		for (int s = 0; s < g->log_h; s++) {
			const int layer = inverse ? s : g->log_h - 1 - s;
			const uint4* row = g->d_C + (size_t)layer * g->log_h;
			const int nbits = g->log_h - 1 - layer;
			if (inverse)
				general_layer_kernel<true><<<grid, 256, 0, st>>>((uint4*)d_out, row, layer, nbits, n_half);
			else
				general_layer_kernel<false><<<grid, 256, 0, st>>>((uint4*)d_out, row, layer, nbits, n_half);
		}
	}
	cudaError_t e = cudaGetLastError();
	return e == cudaSuccess ? B2NTT_OK : report_cuda_error(e, "b2ntt_general_exec");
}
