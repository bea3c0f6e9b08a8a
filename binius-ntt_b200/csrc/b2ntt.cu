// b2ntt.cu -- implementation of the C ABI declared in include/b2ntt.h (libb2ntt.so).
// Host logic only: plan building (b2_plan.h), device buffers, launches.  No CPU fallback: every
// compute entry point needs a CUDA device and returns B2NTT_ECUDA without one.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>

#include "../../include/b2ntt.h"
#include "b2_launch.h"
#include "b2_plan.h"

using namespace b2;

static thread_local char g_cuda_err[256] = "";

static int cuda_fail(cudaError_t e, const char* where) {
	snprintf(g_cuda_err, sizeof g_cuda_err, "%s: %s", where, cudaGetErrorString(e));
	cudaGetLastError();
	return B2NTT_ECUDA;
}
namespace b2 {
int report_cuda_error(cudaError_t e, const char* where) { return cuda_fail(e, where); }
int report_limit_error(const char* text) {
	snprintf(g_cuda_err, sizeof g_cuda_err, "%s", text);
	return B2NTT_ELIMIT;
}
}  // namespace b2
#define CU(call)                                              \
	do {                                                      \
		cudaError_t e_ = (call);                              \
		if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
	} while (0)

struct b2ntt_plan {
	HostPlan hp;
	int device = 0;
	uint32_t* d_consts = nullptr;   // [32][kConstPitch]
	uint32_t* d_lane_tw = nullptr;  // [5][32]
	// exec_host staging
	cudaStream_t stream = nullptr;
	uint32_t* d_in = nullptr;
	uint32_t* d_out = nullptr;
	// exec_host_batch pipeline: 3 buffer sets, 3 streams (H2D, transform, D2H)
	static constexpr int kPipe = 3;
	cudaStream_t ps[3] = {nullptr, nullptr, nullptr};
	uint32_t* pd_in[kPipe] = {nullptr, nullptr, nullptr};
	uint32_t* pd_out[kPipe] = {nullptr, nullptr, nullptr};
	cudaEvent_t ev_in[kPipe] = {}, ev_cmp[kPipe] = {}, ev_out[kPipe] = {};
	bool pipe_ready = false;
};

static size_t elem_words(const HostPlan& hp) { return (size_t)1 << hp.limbs_log; }
static size_t in_words(const HostPlan& hp) { return ((size_t)hp.batch << hp.log_h) * elem_words(hp); }
static size_t out_words(const HostPlan& hp) { return in_words(hp) << hp.log_rate; }

extern "C" int b2ntt_plan_create(b2ntt_plan** out, int field_bits, int layout, int log_h, int log_rate, int batch, int device) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	if (!plan_args_ok(field_bits, layout, log_h, log_rate, batch)) return B2NTT_EINVAL;
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return cuda_fail(cudaErrorNoDevice, "cudaGetDeviceCount");
	if (device < 0 || device >= ndev) return B2NTT_EINVAL;
	b2ntt_plan* p = new (std::nothrow) b2ntt_plan;
	if (!p) return B2NTT_ENOMEM;
	int cta = 0;
	if (const char* e = getenv("B2NTT_CTA_THREADS")) cta = atoi(e);  // tuning knob: 64, 128 or 256
	int glines = 0;
	if (const char* e = getenv("B2NTT_GROUP_LINES")) glines = atoi(e);  // tuning knob: 128, 256 or 512
	// schedule: the upper + tile passes win from 2^21 lines (256 MiB) of total work; below, the bundle passes (smaller
	// groups, fewer launches) fill the SMs better (profiles/r1w_*: 1.5x at log_h = 20 GF(2^32); round-2 kernels, profiles/
	// r2z_schedule_sweep.log: 2^21 lines 1.62 vs 1.72 ms (GF128) and 1.95 vs 1.94 ms (GF32), 2^20 lines 0.91 vs 0.84 and 1.11 vs 0.95)
	const int line_bits = log_h + log_rate + (field_bits == 128 ? 2 : 0) - 5;
	const uint64_t total_lines = (uint64_t)(batch > 0 ? batch : 1) << (line_bits > 0 ? line_bits : 0);
	int upper = total_lines >= (uint64_t(1) << 21) ? 2 : 0;
	if (const char* e = getenv("B2NTT_UPPER")) upper = atoi(e);  // tuning knob: 0 = bundle passes only, 1 = upper passes over a bundle pass, 2 = upper passes over the tile pass, 5 = upper passes over the block-sliced pass
	if (plan_build(p->hp, field_bits, layout, log_h, log_rate, batch, cta, glines, upper)) {
		delete p;
		return B2NTT_EINVAL;
	}
	p->device = device;
	DeviceGuard guard(device);
	cudaError_t e = guard.err;
	if (e == cudaSuccess) e = cudaMalloc(&p->d_consts, sizeof p->hp.consts);
	if (e == cudaSuccess) e = cudaMalloc(&p->d_lane_tw, sizeof p->hp.lane_tw);
	if (e == cudaSuccess) e = cudaMemcpy(p->d_consts, p->hp.consts, sizeof p->hp.consts, cudaMemcpyHostToDevice);
	if (e == cudaSuccess) e = cudaMemcpy(p->d_lane_tw, p->hp.lane_tw, sizeof p->hp.lane_tw, cudaMemcpyHostToDevice);
	if (e != cudaSuccess) {
		const int rc = e == cudaErrorMemoryAllocation ? B2NTT_ENOMEM : cuda_fail(e, "plan_create");
		b2ntt_plan_destroy(p);
		return rc;
	}
	*out = p;
	return B2NTT_OK;
}

extern "C" int b2ntt_plan_destroy(b2ntt_plan* p) {
	if (!p) return B2NTT_OK;
	DeviceGuard guard(p->device);
	if (p->d_consts) cudaFree(p->d_consts);
	if (p->d_lane_tw) cudaFree(p->d_lane_tw);
	if (p->d_in) cudaFree(p->d_in);
	if (p->d_out) cudaFree(p->d_out);
	if (p->stream) cudaStreamDestroy(p->stream);
	for (int i = 0; i < b2ntt_plan::kPipe; i++) {
		if (p->pd_in[i]) cudaFree(p->pd_in[i]);
		if (p->pd_out[i]) cudaFree(p->pd_out[i]);
		if (p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
		if (p->ev_cmp[i]) cudaEventDestroy(p->ev_cmp[i]);
		if (p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
	}
	for (int i = 0; i < 3; i++)
		if (p->ps[i]) cudaStreamDestroy(p->ps[i]);
	delete p;
	return B2NTT_OK;
}

static int run(b2ntt_plan* p, const void* d_in, void* d_out, bool inverse, int coset, cudaStream_t st, bool one_coset = false,
               cudaEvent_t* ev = nullptr /* np + 1 events recorded around the pass launches */) {
	const HostPlan& hp = p->hp;
	if (!d_in || !d_out) return B2NTT_EINVAL;
	const bool single = inverse || one_coset;
	if (d_in == d_out && hp.log_rate != 0 && !single) return B2NTT_EINVAL;
	// the buffers must live on the plan's device (a pointer of another GPU would be dereferenced by this one's kernels)
	if (pointer_device(d_in) != p->device || pointer_device(d_out) != p->device) return B2NTT_EINVAL;
	DeviceGuard guard(p->device);
	CU(guard.err);
	const uint32_t* src = (const uint32_t*)d_in;
	uint32_t* dst = (uint32_t*)d_out;
	if (hp.tiny) {
		TinyParams T;
		memset(&T, 0, sizeof T);
		T.src = src, T.dst = dst, T.consts = p->d_consts;
		T.log_h = hp.log_h, T.log_rate = hp.log_rate, T.limbs_log = hp.limbs_log;
		T.rows = single ? hp.batch : hp.batch << hp.log_rate;
		T.src_row_shift = single ? 0 : hp.log_rate;
		T.coset_mask = single ? 0u : ((1u << hp.log_rate) - 1);
		T.coset0 = single ? (uint32_t)coset : 0u;
		T.inverse = inverse ? 1 : 0;
		if (launch_tiny_fanout(T, st)) return cuda_fail(cudaGetLastError(), "tiny_fanout");
		const uint64_t words = (uint64_t)T.rows << (hp.log_h + hp.limbs_log);
		const int width = 32 << hp.limbs_log;
		if (hp.layout == B2NTT_LAYOUT_BITSLICED && launch_bitslice(dst, words / width, width, 1, st))
			return cuda_fail(cudaGetLastError(), "bitslice");
		if (launch_tiny_layers(T, st)) return cuda_fail(cudaGetLastError(), "tiny_layers");
		if (hp.layout == B2NTT_LAYOUT_BITSLICED && launch_bitslice(dst, words / width, width, 0, st))
			return cuda_fail(cudaGetLastError(), "bitslice");
		return B2NTT_OK;
	}
	PassParams P;
	UpperParams U;
	BottomParams B;
	const int np = (int)hp.passes.size();
	if (ev) CU(cudaEventRecord(ev[0], st));
	for (int n = 0; n < np; n++) {
		const int kind = (inverse ? hp.passes[np - 1 - n] : hp.passes[n]).kind;
		if (kind == 1) {
			plan_fill_upper(hp, U, n, src, dst, p->d_consts, inverse, coset, one_coset);
			// B2NTT_UPPER_PIPE=1: the producer / consumer form of the k = 4, 6 passes (b2_upper_pipe.cu); measured slower than
			// upper_kernel (4.36 / 5.13 / 4.62 ms against 3.80 / 4.32 / 3.92 on the headline transform), so it is opt-in
			const char* pe = getenv("B2NTT_UPPER_PIPE");
			if (pe && atoi(pe) != 0 && U.k >= 4 && U.n_groups >= 512) {
				if (launch_upper_pipe(U, st)) return cuda_fail(cudaGetLastError(), "launch_upper_pipe");
			} else if (launch_upper(U, st))
				return cuda_fail(cudaGetLastError(), "launch_upper");
		} else if (kind == 2) {
			plan_fill_bottom(hp, B, n, src, dst, p->d_consts, p->d_lane_tw, inverse, coset, one_coset);
			if (launch_bottom(B, st)) return cuda_fail(cudaGetLastError(), "launch_bottom");
		} else if (kind == 4) {
			BslParams L;
			plan_fill_bsl(hp, L, n, src, dst, p->d_consts, inverse, coset, one_coset);
			if (launch_bsl(L, st)) return cuda_fail(cudaGetLastError(), "launch_bsl");
		} else {
			plan_fill_pass(hp, P, n, src, dst, p->d_consts, p->d_lane_tw, inverse, coset, one_coset);
			if (launch_pass(P, st)) return cuda_fail(cudaGetLastError(), "launch_pass");
		}
		if (ev) CU(cudaEventRecord(ev[n + 1], st));
	}
	return B2NTT_OK;
}

extern "C" int b2ntt_exec(b2ntt_plan* p, const void* d_in, void* d_out, void* stream) {
	if (!p) return B2NTT_EINVAL;
	return run(p, d_in, d_out, false, 0, (cudaStream_t)stream);
}

extern "C" int b2ntt_exec_profile(b2ntt_plan* p, const void* d_in, void* d_out, void* stream, float* pass_ms, int* pass_kind, int max_passes) {
	if (!p || !pass_ms || max_passes < 1) return B2NTT_EINVAL;
	if (p->hp.tiny) return B2NTT_EUNSUPPORTED;
	const int np = (int)p->hp.passes.size();
	if (np > max_passes || np > 15) return B2NTT_EINVAL;
	DeviceGuard guard(p->device);
	CU(guard.err);
	cudaEvent_t ev[16] = {};
	for (int i = 0; i <= np; i++) {
		const cudaError_t e = cudaEventCreate(&ev[i]);
		if (e != cudaSuccess) {
			for (int k = 0; k < i; k++) cudaEventDestroy(ev[k]);
			return cuda_fail(e, "cudaEventCreate");
		}
	}
	int rc = run(p, d_in, d_out, false, 0, (cudaStream_t)stream, false, ev);
	if (rc == B2NTT_OK && cudaEventSynchronize(ev[np]) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventSynchronize");
	for (int i = 0; i < np && rc == B2NTT_OK; i++) {
		if (cudaEventElapsedTime(&pass_ms[i], ev[i], ev[i + 1]) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventElapsedTime");
		if (pass_kind) pass_kind[i] = p->hp.passes[i].kind;
	}
	for (int i = 0; i <= np; i++) cudaEventDestroy(ev[i]);
	return rc == B2NTT_OK ? np : rc;
}

extern "C" int b2ntt_exec_inverse(b2ntt_plan* p, const void* d_in, void* d_out, int coset, void* stream) {
	if (!p) return B2NTT_EINVAL;
	if (coset < 0 || coset >= (1 << p->hp.log_rate)) return B2NTT_EINVAL;
	return run(p, d_in, d_out, true, coset, (cudaStream_t)stream);
}

extern "C" int b2ntt_exec_coset(b2ntt_plan* p, const void* d_in, void* d_out, int coset, void* stream) {
	if (!p) return B2NTT_EINVAL;
	if (coset < 0 || coset >= (1 << p->hp.log_rate)) return B2NTT_EINVAL;
	return run(p, d_in, d_out, false, coset, (cudaStream_t)stream, true);
}

extern "C" int b2ntt_exec_top(b2ntt_plan* p, void* d_buf, int log_segments, void* stream) {
	if (!p || !d_buf) return B2NTT_EINVAL;
	PassParams P;
	if (p->hp.tiny || plan_fill_top(p->hp, P, log_segments, (uint32_t*)d_buf, p->d_consts, p->d_lane_tw)) return B2NTT_EINVAL;
	if (pointer_device(d_buf) != p->device) return B2NTT_EINVAL;
	DeviceGuard guard(p->device);
	CU(guard.err);
	if (launch_pass(P, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError(), "launch_pass");
	return B2NTT_OK;
}

extern "C" int b2ntt_exec_host(b2ntt_plan* p, const void* h_in, size_t n_in, void* h_out) {
	if (!p || !h_in || !h_out) return B2NTT_EINVAL;
	const HostPlan& hp = p->hp;
	if (n_in != ((size_t)1 << hp.log_h)) return B2NTT_EINVAL;  // additive_ntt.cuh:206-208
	DeviceGuard guard(p->device);
	CU(guard.err);
	if (!p->stream) CU(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
	const size_t ib = in_words(hp) * 4, ob = out_words(hp) * 4;
	if (!p->d_out) {
		cudaError_t e = cudaMalloc(&p->d_out, ob);
		if (e != cudaSuccess) return e == cudaErrorMemoryAllocation ? (cudaGetLastError(), B2NTT_ENOMEM) : cuda_fail(e, "cudaMalloc");
	}
	if (hp.log_rate != 0 && !p->d_in) {
		cudaError_t e = cudaMalloc(&p->d_in, ib);
		if (e != cudaSuccess) return e == cudaErrorMemoryAllocation ? (cudaGetLastError(), B2NTT_ENOMEM) : cuda_fail(e, "cudaMalloc");
	}
	uint32_t* din = hp.log_rate ? p->d_in : p->d_out;
	CU(cudaMemcpyAsync(din, h_in, ib, cudaMemcpyHostToDevice, p->stream));
	const int rc = run(p, din, p->d_out, false, 0, p->stream);
	if (rc) return rc;
	CU(cudaMemcpyAsync(h_out, p->d_out, ob, cudaMemcpyDeviceToHost, p->stream));
	CU(cudaStreamSynchronize(p->stream));
	return B2NTT_OK;
}

// Pipelined host path: transform i is copied in on one stream while transform i-1 runs on a second and transform
// i-2 is copied out on a third (PCIe is full duplex), three device buffer sets in rotation.
extern "C" int b2ntt_exec_host_batch(b2ntt_plan* p, const void* const* h_in, size_t n_in, void* const* h_out, int count) {
	if (!p || !h_in || !h_out || count < 0) return B2NTT_EINVAL;
	const HostPlan& hp = p->hp;
	if (n_in != ((size_t)1 << hp.log_h)) return B2NTT_EINVAL;
	for (int i = 0; i < count; i++)
		if (!h_in[i] || !h_out[i]) return B2NTT_EINVAL;
	if (count == 0) return B2NTT_OK;
	DeviceGuard guard(p->device);
	CU(guard.err);
	const size_t ib = in_words(hp) * 4, ob = out_words(hp) * 4;
	constexpr int K = b2ntt_plan::kPipe;
	if (!p->pipe_ready) {
		// every resource is created only if its handle is still null: a call that failed half-way (ENOMEM) is retried
		// without leaking what the first attempt made; b2ntt_plan_destroy frees whatever exists
		for (int i = 0; i < 3; i++)
			if (!p->ps[i]) CU(cudaStreamCreateWithFlags(&p->ps[i], cudaStreamNonBlocking));
		for (int i = 0; i < K; i++) {
			cudaError_t e = p->pd_out[i] ? cudaSuccess : cudaMalloc(&p->pd_out[i], ob);
			if (e == cudaSuccess && hp.log_rate && !p->pd_in[i]) e = cudaMalloc(&p->pd_in[i], ib);
			if (e != cudaSuccess) return e == cudaErrorMemoryAllocation ? (cudaGetLastError(), B2NTT_ENOMEM) : cuda_fail(e, "cudaMalloc");
			if (!p->ev_in[i]) CU(cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming));
			if (!p->ev_cmp[i]) CU(cudaEventCreateWithFlags(&p->ev_cmp[i], cudaEventDisableTiming));
			if (!p->ev_out[i]) CU(cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming));
		}
		p->pipe_ready = true;
	}
	int rc = B2NTT_OK;
	cudaError_t ce = cudaSuccess;
#define PIPE(call)                                      \
	if (rc == B2NTT_OK && (ce = (call)) != cudaSuccess) rc = cuda_fail(ce, #call)
	for (int i = 0; i < count && rc == B2NTT_OK; i++) {
		const int b = i % K;
		uint32_t* din = hp.log_rate ? p->pd_in[b] : p->pd_out[b];
		if (i >= K) PIPE(cudaStreamWaitEvent(p->ps[0], p->ev_out[b], 0));  // the buffer set is free again
		PIPE(cudaMemcpyAsync(din, h_in[i], ib, cudaMemcpyHostToDevice, p->ps[0]));
		PIPE(cudaEventRecord(p->ev_in[b], p->ps[0]));
		PIPE(cudaStreamWaitEvent(p->ps[1], p->ev_in[b], 0));
		if (rc == B2NTT_OK) rc = run(p, din, p->pd_out[b], false, 0, p->ps[1]);
		PIPE(cudaEventRecord(p->ev_cmp[b], p->ps[1]));
		PIPE(cudaStreamWaitEvent(p->ps[2], p->ev_cmp[b], 0));
		PIPE(cudaMemcpyAsync(h_out[i], p->pd_out[b], ob, cudaMemcpyDeviceToHost, p->ps[2]));
		PIPE(cudaEventRecord(p->ev_out[b], p->ps[2]));
	}
#undef PIPE
	// drain all three streams whatever happened: no copy on the caller's buffers is in flight after the call returns
	for (int i = 2; i >= 0; i--) {
		const cudaError_t e = cudaStreamSynchronize(p->ps[i]);
		if (e != cudaSuccess && rc == B2NTT_OK) rc = cuda_fail(e, "cudaStreamSynchronize");
	}
	return rc;
}

extern "C" int b2ntt_plan_launches(const b2ntt_plan* p) {
	if (!p) return B2NTT_EINVAL;
	const HostPlan& hp = p->hp;
	if (hp.tiny) return hp.log_h + (hp.log_rate ? 1 : 0) + (hp.layout == B2NTT_LAYOUT_BITSLICED ? 2 : 0);
	return (int)hp.passes.size();
}
extern "C" int b2ntt_plan_passes(const b2ntt_plan* p) {
	if (!p) return B2NTT_EINVAL;
	return p->hp.tiny ? p->hp.log_h : (int)p->hp.passes.size();
}
extern "C" int b2ntt_plan_constants(const b2ntt_plan* p, uint32_t* h_out) {
	if (!p || !h_out) return B2NTT_EINVAL;
	memcpy(h_out, p->hp.consts, sizeof p->hp.consts);
	return B2NTT_OK;
}

extern "C" int b2ntt_bitslice(void* d_buf, size_t n_blocks, int width_bits, int dir, void* stream) {
	if (width_bits != 32 && width_bits != 64 && width_bits != 128) return B2NTT_EINVAL;
	if (dir != 0 && dir != 1) return B2NTT_EINVAL;
	if (n_blocks == 0) return B2NTT_OK;
	if (!d_buf) return B2NTT_EINVAL;
	const int dev = pointer_device(d_buf);
	if (dev < 0) return B2NTT_EINVAL;
	DeviceGuard guard(dev);
	CU(guard.err);
	if (launch_bitslice((uint32_t*)d_buf, n_blocks, width_bits, dir, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError(), "bitslice");
	return B2NTT_OK;
}

extern "C" int b2ntt_mul(int field_bits, int layout, const void* a, const void* b, void* o, size_t n, void* stream) {
	if (layout == B2NTT_LAYOUT_COMPACT) {
		if (field_bits != 8 && field_bits != 16 && field_bits != 32 && field_bits != 128) return B2NTT_EINVAL;
	} else if (layout == B2NTT_LAYOUT_BITSLICED) {
		if ((field_bits != 32 && field_bits != 128) || n % 32) return B2NTT_EINVAL;
	} else
		return B2NTT_EINVAL;
	if (n == 0) return B2NTT_OK;
	if (!a || !b || !o) return B2NTT_EINVAL;
	const int dev = pointer_device(o);
	if (dev < 0 || pointer_device(a) != dev || pointer_device(b) != dev) return B2NTT_EINVAL;
	DeviceGuard guard(dev);
	CU(guard.err);
	if (launch_mul(field_bits, layout, (const uint32_t*)a, (const uint32_t*)b, (uint32_t*)o, n, (cudaStream_t)stream))
		return cuda_fail(cudaGetLastError(), "mul");
	return B2NTT_OK;
}

extern "C" int b2ntt_mul_repeat(int field_bits, void* d_io, size_t n_threads, int iters, void* stream) {
	if (field_bits != 32 && field_bits != 128) return B2NTT_EINVAL;
	if (!d_io || iters < 0) return B2NTT_EINVAL;
	const int dev = pointer_device(d_io);
	if (dev < 0) return B2NTT_EINVAL;
	DeviceGuard guard(dev);
	CU(guard.err);
	if (launch_mul_repeat(field_bits, (uint32_t*)d_io, n_threads, iters, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError(), "mul_repeat");
	return B2NTT_OK;
}

// utils/common.cu:6-43: >= 1024 threads per block, > 32 KB shared memory, grid >= (2^20, 2^12, 2^15)
extern "C" int b2ntt_check_gpu_capabilities(int device) {
	cudaDeviceProp pr;
	cudaError_t e = cudaGetDeviceProperties(&pr, device);
	if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
	const bool ok = pr.maxThreadsPerBlock >= 1024 && pr.sharedMemPerBlock > (1u << 15) && pr.maxGridSize[0] >= (1 << 20) &&
	                pr.maxGridSize[1] >= (1 << 12) && pr.maxGridSize[2] >= (1 << 15) && pr.major >= 10;
	return ok ? 1 : 0;
}

extern "C" const char* b2ntt_strerror(int code) {
	switch (code) {
		case B2NTT_OK: return "ok";
		case B2NTT_EINVAL: return "invalid argument";
		case B2NTT_ENOMEM: return "out of memory";
		case B2NTT_ECUDA: return "CUDA error or no usable device";
		case B2NTT_EUNSUPPORTED: return "not supported by this build";
		case B2NTT_ECHECK: return "self-check failed";
		case B2NTT_ELIMIT: return "size beyond the fixed scratch of a no-allocation entry point";
		default: return "unknown error";
	}
}
extern "C" const char* b2ntt_last_cuda_error(void) { return g_cuda_err; }

// ================================================================================================================================
// Multi-GPU: one transform split over 2^g devices with the exchange fused into the first phase's stores, and
// the single-process drivers b2ntt_mg_* for a C++ caller that owns all the GPUs of the box.
// ================================================================================================================================
struct b2ntt_fourstep {
	b2ntt_plan* full = nullptr;   // plan of the whole transform: constants and classes of the top g stages
	b2ntt_plan* local = nullptr;  // (log_h = log_n - g, log_rate = g): phase 2 is its coset `rank`
	int g = 0, rank = 0, device = 0, log_n = 0, field_bits = 0;
	void* peers[16] = {};         // block buffer of every rank, mapped in this process
};

extern "C" int b2ntt_fourstep_create(b2ntt_fourstep** out, int field_bits, int log_n, int log_ranks, int rank, int device) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	if (log_ranks < 0 || log_ranks > 4 || rank < 0 || rank >= (1 << log_ranks) || log_n - 2 * log_ranks < 1) return B2NTT_EINVAL;
	const int limbs_log = field_bits == 128 ? 2 : 0;
	if (log_ranks && log_n - 2 * log_ranks + limbs_log < 5) return B2NTT_EINVAL;  // a segment is at least one 128-byte line
	b2ntt_fourstep* fs = new (std::nothrow) b2ntt_fourstep;
	if (!fs) return B2NTT_ENOMEM;
	fs->g = log_ranks, fs->rank = rank, fs->device = device, fs->log_n = log_n, fs->field_bits = field_bits;
	int rc = log_ranks ? b2ntt_plan_create(&fs->full, field_bits, B2NTT_LAYOUT_COMPACT, log_n, 0, 1, device) : B2NTT_OK;
	if (rc == B2NTT_OK) rc = b2ntt_plan_create(&fs->local, field_bits, B2NTT_LAYOUT_COMPACT, log_n - log_ranks, log_ranks, 1, device);
	if (rc != B2NTT_OK) {
		b2ntt_plan_destroy(fs->full), b2ntt_plan_destroy(fs->local);
		delete fs;
		return rc;
	}
	*out = fs;
	return B2NTT_OK;
}
extern "C" int b2ntt_fourstep_destroy(b2ntt_fourstep* fs) {
	if (!fs) return B2NTT_OK;
	b2ntt_plan_destroy(fs->full), b2ntt_plan_destroy(fs->local);
	delete fs;
	return B2NTT_OK;
}
extern "C" size_t b2ntt_fourstep_block_bytes(const b2ntt_fourstep* fs) {
	return fs ? ((size_t)(fs->field_bits / 8) << (fs->log_n - fs->g)) : 0;
}
extern "C" int b2ntt_fourstep_launches(const b2ntt_fourstep* fs) {
	return fs ? (fs->g ? 1 : 0) + b2ntt_plan_launches(fs->local) : B2NTT_EINVAL;
}
extern "C" int b2ntt_fourstep_set_peer(b2ntt_fourstep* fs, int peer_rank, void* d_block) {
	if (!fs || peer_rank < 0 || peer_rank >= (1 << fs->g) || !d_block) return B2NTT_EINVAL;
	fs->peers[peer_rank] = d_block;
	return B2NTT_OK;
}
// Phase 1 + exchange: the top g stages across this rank's 2^g segments (h, m = rank, .), segment h stored straight into rank
// h's block buffer at slot `rank`.  d_shard is read only.  Every rank's block must be free of readers (a barrier before).
extern "C" int b2ntt_fourstep_scatter(b2ntt_fourstep* fs, const void* d_shard, void* stream) {
	if (!fs || !d_shard) return B2NTT_EINVAL;
	for (int r = 0; r < (1 << fs->g); r++)
		if (!fs->peers[r]) return B2NTT_EINVAL;
	if (pointer_device(d_shard) != fs->device) return B2NTT_EINVAL;
	DeviceGuard guard(fs->device);
	CU(guard.err);
	if (fs->g == 0) {
		if (fs->peers[0] != d_shard) CU(cudaMemcpyAsync(fs->peers[0], d_shard, b2ntt_fourstep_block_bytes(fs), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
		return B2NTT_OK;
	}
	PassParams P;
	b2ntt_plan* p = fs->full;
	if (p->hp.tiny || plan_fill_top(p->hp, P, fs->g, (uint32_t*)d_shard, p->d_consts, p->d_lane_tw, fs->peers, fs->rank)) return B2NTT_EINVAL;
	if (launch_pass(P, (cudaStream_t)stream)) return cuda_fail(cudaGetLastError(), "launch_pass (scatter)");
	return B2NTT_OK;
}
// Phase 2: the local transform of the block this rank owns, in place (needs every rank's scatter to have completed).
extern "C" int b2ntt_fourstep_local(b2ntt_fourstep* fs, void* d_block, void* stream) {
	if (!fs || !d_block) return B2NTT_EINVAL;
	return run(fs->local, d_block, d_block, false, fs->rank, (cudaStream_t)stream, true);
}

// ---- CUDA IPC helpers for one-process-per-GPU callers (torchrun): export a device buffer, map a peer's --------------------
extern "C" int b2ntt_ipc_export(const void* d_ptr, void* handle64, size_t* offset) {
	if (!d_ptr || !handle64 || !offset) return B2NTT_EINVAL;
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
	const int dev = pointer_device(d_ptr);
	if (dev < 0) return B2NTT_EINVAL;
	DeviceGuard guard(dev);
	CU(guard.err);
	// base of the allocation that holds d_ptr: cuMemGetAddressRange of the driver API, looked up at run time (the library
	// does not link libcuda, so it still loads on a machine without a driver)
	typedef int (*range_fn)(unsigned long long*, size_t*, unsigned long long);
	void* fn = nullptr;
	cudaDriverEntryPointQueryResult qres;
	CU(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qres));
	if (!fn || qres != cudaDriverEntryPointSuccess) return B2NTT_EUNSUPPORTED;
	unsigned long long base_addr = 0;
	size_t size = 0;
	if (((range_fn)fn)(&base_addr, &size, (unsigned long long)(uintptr_t)d_ptr) != 0) return cuda_fail(cudaErrorInvalidValue, "cuMemGetAddressRange");
	void* base = (void*)(uintptr_t)base_addr;
	CU(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)handle64, base));
	*offset = (size_t)((const char*)d_ptr - (const char*)base);
	return B2NTT_OK;
}
extern "C" int b2ntt_ipc_open(const void* handle64, size_t offset, int device, void** d_ptr) {
	if (!handle64 || !d_ptr) return B2NTT_EINVAL;
	DeviceGuard guard(device);
	CU(guard.err);
	void* base = nullptr;
	cudaIpcMemHandle_t h;
	memcpy(&h, handle64, sizeof h);
	CU(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
	*d_ptr = (char*)base + offset;
	return B2NTT_OK;
}
extern "C" int b2ntt_ipc_close(void* d_ptr, size_t offset, int device) {
	if (!d_ptr) return B2NTT_OK;
	DeviceGuard guard(device);
	CU(guard.err);
	CU(cudaIpcCloseMemHandle((char*)d_ptr - offset));
	return B2NTT_OK;
}

// ---- single-process drivers: one host thread, one stream per device -------------------------------------------------------------
struct b2ntt_mg {
	int mode = 0, n = 0, field_bits = 0, log_h = 0, log_rate = 0, batch = 0;
	int devices[16] = {};
	cudaStream_t streams[16] = {};
	cudaEvent_t ev_scatter[16] = {};
	b2ntt_plan* plans[16] = {};      // COLUMN_SHARD: device i transforms columns [first[i], first[i] + count[i])
	int first[16] = {}, count[16] = {};
	b2ntt_fourstep* fs[16] = {};     // FOUR_STEP
};

extern "C" int b2ntt_mg_destroy(b2ntt_mg* mg) {
	if (!mg) return B2NTT_OK;
	for (int i = 0; i < mg->n; i++) {
		DeviceGuard guard(mg->devices[i]);
		if (mg->streams[i]) cudaStreamDestroy(mg->streams[i]);
		if (mg->ev_scatter[i]) cudaEventDestroy(mg->ev_scatter[i]);
		b2ntt_plan_destroy(mg->plans[i]);
		b2ntt_fourstep_destroy(mg->fs[i]);
	}
	delete mg;
	return B2NTT_OK;
}

extern "C" int b2ntt_mg_create(b2ntt_mg** out, int mode, int n_gpus, const int* devices, int field_bits, int log_h, int log_rate, int batch) {
	if (!out) return B2NTT_EINVAL;
	*out = nullptr;
	if (!devices || n_gpus < 1 || n_gpus > 16 || (mode != B2NTT_MG_COLUMN_SHARD && mode != B2NTT_MG_FOUR_STEP)) return B2NTT_EINVAL;
	int g = 0;
	while ((1 << g) < n_gpus) g++;
	if (mode == B2NTT_MG_FOUR_STEP && ((1 << g) != n_gpus || log_rate != 0 || batch != 1)) return B2NTT_EINVAL;
	if (mode == B2NTT_MG_COLUMN_SHARD && batch < 1) return B2NTT_EINVAL;
	for (int i = 0; i < n_gpus; i++)
		for (int j = 0; j < i; j++)
			if (devices[i] == devices[j]) return B2NTT_EINVAL;
	b2ntt_mg* mg = new (std::nothrow) b2ntt_mg;
	if (!mg) return B2NTT_ENOMEM;
	mg->mode = mode, mg->n = n_gpus, mg->field_bits = field_bits, mg->log_h = log_h, mg->log_rate = log_rate, mg->batch = batch;
	int rc = B2NTT_OK;
	for (int i = 0; i < n_gpus && rc == B2NTT_OK; i++) {
		mg->devices[i] = devices[i];
		DeviceGuard guard(devices[i]);
		if (guard.err != cudaSuccess || cudaStreamCreateWithFlags(&mg->streams[i], cudaStreamNonBlocking) != cudaSuccess ||
		    cudaEventCreateWithFlags(&mg->ev_scatter[i], cudaEventDisableTiming) != cudaSuccess) {
			rc = cuda_fail(cudaGetLastError(), "b2ntt_mg_create");
			break;
		}
		if (mode == B2NTT_MG_COLUMN_SHARD) {
			// contiguous slabs of columns, sizes differing by at most one
			const int base = batch / n_gpus, extra = batch % n_gpus;
			mg->first[i] = i * base + (i < extra ? i : extra), mg->count[i] = base + (i < extra ? 1 : 0);
			if (mg->count[i]) rc = b2ntt_plan_create(&mg->plans[i], field_bits, B2NTT_LAYOUT_COMPACT, log_h, log_rate, mg->count[i], devices[i]);
		} else {
			rc = b2ntt_fourstep_create(&mg->fs[i], field_bits, log_h, g, i, devices[i]);
			for (int j = 0; j < n_gpus && rc == B2NTT_OK; j++) {  // the scatter stores into every other device's block
				if (j == i) continue;
				int can = 0;
				cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
				if (!can) {
					rc = B2NTT_EUNSUPPORTED;
					break;
				}
				const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
				if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = cuda_fail(e, "cudaDeviceEnablePeerAccess");
				cudaGetLastError();
			}
		}
	}
	if (rc != B2NTT_OK) {
		b2ntt_mg_destroy(mg);
		return rc;
	}
	*out = mg;
	return B2NTT_OK;
}

extern "C" int b2ntt_mg_columns(const b2ntt_mg* mg, int i, int* first, int* count) {
	if (!mg || i < 0 || i >= mg->n || mg->mode != B2NTT_MG_COLUMN_SHARD) return B2NTT_EINVAL;
	if (first) *first = mg->first[i];
	if (count) *count = mg->count[i];
	return B2NTT_OK;
}

// Launches on every device, then waits for all of them.  COLUMN_SHARD: d_in[i] / d_out[i] = device i's columns (no data moves
// between devices).  FOUR_STEP: d_in[i] = rank i's m-shard (2^g segments (h, m = i, .), read only), d_out[i] = block i of the
// transform in natural order; the only cross-device traffic is the scatter's peer stores.
extern "C" int b2ntt_mg_exec(b2ntt_mg* mg, void* const* d_in, void* const* d_out) {
	if (!mg || !d_in || !d_out) return B2NTT_EINVAL;
	int rc = B2NTT_OK;
	if (mg->mode == B2NTT_MG_COLUMN_SHARD) {
		for (int i = 0; i < mg->n && rc == B2NTT_OK; i++)
			if (mg->plans[i]) rc = b2ntt_exec(mg->plans[i], d_in[i], d_out[i], mg->streams[i]);
	} else {
		for (int i = 0; i < mg->n && rc == B2NTT_OK; i++)
			for (int j = 0; j < mg->n && rc == B2NTT_OK; j++) rc = b2ntt_fourstep_set_peer(mg->fs[i], j, d_out[j]);
		for (int i = 0; i < mg->n && rc == B2NTT_OK; i++) {
			rc = b2ntt_fourstep_scatter(mg->fs[i], d_in[i], mg->streams[i]);
			DeviceGuard guard(mg->devices[i]);
			if (rc == B2NTT_OK && cudaEventRecord(mg->ev_scatter[i], mg->streams[i]) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventRecord");
		}
		for (int i = 0; i < mg->n && rc == B2NTT_OK; i++) {
			DeviceGuard guard(mg->devices[i]);
			for (int j = 0; j < mg->n; j++)  // block i is complete when every device's scatter has finished
				if (cudaStreamWaitEvent(mg->streams[i], mg->ev_scatter[j], 0) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaStreamWaitEvent");
			if (rc == B2NTT_OK) rc = b2ntt_fourstep_local(mg->fs[i], d_out[i], mg->streams[i]);
		}
	}
	for (int i = 0; i < mg->n; i++) {  // drain every stream even after an error: no launch outlives the call
		DeviceGuard guard(mg->devices[i]);
		if (cudaStreamSynchronize(mg->streams[i]) != cudaSuccess && rc == B2NTT_OK) rc = cuda_fail(cudaGetLastError(), "cudaStreamSynchronize");
	}
	return rc;
}
