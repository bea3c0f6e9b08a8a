// ulvt/ntt/nttconf.cuh -- drop-in for ntt/nttconf.cuh:9-60: DataOrder, NTTData<E> and
// AdditiveNTTConf<T,P>.  The launch-shape tables of the reference (KERNEL_LAUNCH_CONFIGS,
// get_kernel_launch_confs, nttconf.cuh:65-110 / nttconf.cu) have no role here: the pass schedule is
// built inside libb2ntt.so by b2ntt_plan_create.
#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>

#include "../utils/common.cuh"

// The reference exports its hand-tuned launch table as a data symbol of ulvt_gpu (nttconf.cu:1-190).  The symbol is
// kept so that code naming it still links; nothing here consults it (the pass schedule is built by the library).
#if defined(__CUDACC__)
inline dim3 KERNEL_LAUNCH_CONFIGS[31][3][2] = {};
#endif

enum class DataOrder : char { INVALID = -1, IN_ORDER, BIT_REVERSED };

// host buffer + order tag (nttconf.cuh:11-21); owns its memory
template <typename E>
class NTTData {
public:
	DataOrder order;
	std::unique_ptr<E[]> data;
	size_t size;

	NTTData(DataOrder o, size_t n) : order(o), data(new E[n]()), size(n) {}
	explicit NTTData(size_t n) : order(DataOrder::INVALID), data(new E[n]()), size(n) {}
	size_t byte_len() const { return size * sizeof(E); }
};

// nttconf.cuh:49-60: 1 <= log_h, log_h + log_rate <= P::N_BITS(), 0 <= log_rate <= 4
template <typename T, typename P>
class AdditiveNTTConf {
public:
	int log_h;
	int log_rate;

	AdditiveNTTConf(int lh, int lr) : log_h(lh), log_rate(lr) {
		ASSERT(log_h >= 1);
		ASSERT(log_h + log_rate <= (int)P::N_BITS());
		ASSERT(log_rate >= 0 && log_rate <= 4);
	}
};
