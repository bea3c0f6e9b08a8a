// ulvt/ntt/additive_ntt.cuh -- drop-in for AdditiveNTT<T,P> (ntt/additive_ntt.cuh:176-270) on top of
// the C ABI of libb2ntt.so.  Same constructor, `bool apply(const NTTData<T>&, NTTData<T>&)` and
// destructor semantics: apply returns false iff the input size or order is wrong (:206-208), never
// checks output.size, writes 2^(log_h+log_rate) elements coset-major (:250-259) and is synchronous.
// Supported (T, P): (uint32_t, FanPaarTowerField<5>) as in the reference, and the new
// (b128, FanPaarTowerField<7>) whose twiddles stay in GF(2^32) (log_h + log_rate <= 32).
#pragma once
#include <type_traits>

#include "../finite_fields/binary_tower.cuh"
#include "nttconf.cuh"

template <typename T, typename P>
class AdditiveNTT {
	static_assert((std::is_same<T, uint32_t>::value && P::N_BITS() == 32) || (sizeof(T) == 16 && P::N_BITS() == 128),
	              "AdditiveNTT: supported instantiations are (uint32_t, FanPaarTowerField<5>) and (b128, FanPaarTowerField<7>)");

public:
	explicit AdditiveNTT(const AdditiveNTTConf<T, P>& conf) : ntt_conf(conf) {
		// a GF(2^128) transform still draws its domain from GF(2^32): log_h + log_rate <= 32
		// the plan lives on the caller's CURRENT device, as the reference's cudaMalloc'd members do (additive_ntt.cuh:184-198)
		int device = 0;
		cudaGetDevice(&device);
		const int rc = b2ntt_plan_create(&plan, (int)(sizeof(T) * 8), B2NTT_LAYOUT_COMPACT, conf.log_h, conf.log_rate, 1, device);
		if (rc != B2NTT_OK) {
			std::cerr << "AdditiveNTT: " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
			plan = nullptr;
		}
	}
	AdditiveNTT(const AdditiveNTT&) = delete;
	AdditiveNTT& operator=(const AdditiveNTT&) = delete;
	~AdditiveNTT() { b2ntt_plan_destroy(plan); }

	bool apply(const NTTData<T>& input, NTTData<T>& output) {
		if (!plan) return false;
		if (input.size != ((size_t)1 << ntt_conf.log_h) || input.order != DataOrder::IN_ORDER) return false;
		const int rc = b2ntt_exec_host(plan, input.data.get(), input.size, output.data.get());
		if (rc != B2NTT_OK) {
			std::cerr << "AdditiveNTT::apply: " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
			return false;
		}
		output.order = DataOrder::IN_ORDER;
		return true;
	}

	// device-resident variant (no reference counterpart): data already in HBM, asynchronous on `stream`
	bool apply_device(const T* d_in, T* d_out, cudaStream_t stream = 0) {
		return plan && b2ntt_exec(plan, d_in, d_out, (void*)stream) == B2NTT_OK;
	}

private:
	AdditiveNTTConf<T, P> ntt_conf;
	b2ntt_plan* plan = nullptr;
};

// GF(2^128) transform on a caller-supplied basis (no reference counterpart: additive_ntt.cuh:281-283 fixes
// beta_i = 1 << i).  Same apply() contract as AdditiveNTT with log_rate = 0, plus the inverse; basis = nullptr is the
// reference basis and gives AdditiveNTT<b128, FanPaarTowerField<7>>'s output word for word.
class GeneralAdditiveNTT {
public:
	GeneralAdditiveNTT(int log_h, const b128* basis) : log_h(log_h) {
		cudaGetDevice(&device);
		const int rc = b2ntt_general_create(&g, log_h, reinterpret_cast<const uint32_t*>(basis), device);
		if (rc != B2NTT_OK) {
			std::cerr << "GeneralAdditiveNTT: " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
			g = nullptr;
		}
	}
	GeneralAdditiveNTT(const GeneralAdditiveNTT&) = delete;
	GeneralAdditiveNTT& operator=(const GeneralAdditiveNTT&) = delete;
	~GeneralAdditiveNTT() {
		b2ntt_general_destroy(g);
		if (d_stage) {
			int cur = 0;
			cudaGetDevice(&cur);
			cudaSetDevice(device);
			cudaFree(d_stage);
			cudaSetDevice(cur);
		}
	}
	bool valid() const { return g != nullptr; }

	bool apply(const NTTData<b128>& input, NTTData<b128>& output, bool inverse = false) {
		if (!g || input.size != ((size_t)1 << log_h) || input.order != DataOrder::IN_ORDER) return false;
		const size_t bytes = input.size * sizeof(b128);
		int cur = 0;
		cudaGetDevice(&cur);
		cudaSetDevice(device);  // the staging buffer (kept across calls) lives on the object's device
		bool ok = d_stage || cudaMalloc(&d_stage, bytes) == cudaSuccess;
		ok = ok && cudaMemcpy(d_stage, input.data.get(), bytes, cudaMemcpyHostToDevice) == cudaSuccess &&
		     b2ntt_general_exec(g, d_stage, d_stage, inverse, nullptr) == B2NTT_OK &&
		     cudaMemcpy(output.data.get(), d_stage, bytes, cudaMemcpyDeviceToHost) == cudaSuccess;
		cudaSetDevice(cur);
		if (ok) output.order = DataOrder::IN_ORDER;
		return ok;
	}
	bool apply_device(const b128* d_in, b128* d_out, bool inverse = false, cudaStream_t stream = 0) {
		return g && b2ntt_general_exec(g, d_in, d_out, inverse, (void*)stream) == B2NTT_OK;
	}

private:
	int log_h;
	int device = 0;
	void* d_stage = nullptr;
	b2ntt_general* g = nullptr;
};
