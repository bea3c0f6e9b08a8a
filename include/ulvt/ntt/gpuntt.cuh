// ulvt/ntt/gpuntt.cuh -- drop-in for NTT<E> (ntt/gpuntt.cuh:126-209), the prime-field radix-2 transform that shares ntt_tests
// with the additive NTT.  Same constructor (NTTConfRad2<E>), `void apply(const NTTData<E>&, NTTData<E>&)` and
// destructor; E must be a 4-byte field whose memory image is the BabyBear Montgomery word (BB31 = risc0::Fp).  The work is
// done by b2ntt_bb31_* (include/b2ntt.h): bit reversal, device-built twiddles, fused shared-memory passes.
#pragma once
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <vector>

#include "../finite_fields/baby_bear.cuh"
#include "nttconf.cuh"

// index with its low `size` bits reversed (gpuntt.cuh:12-19)
static inline B2_BB_HD int reverse_bits(uint32_t n, const int size) {
	uint32_t r = 0;
	for (int i = 0; i < size; i++) r |= ((n >> i) & 1u) << (size - 1 - i);
	return (int)r;
}

template <typename E>
class NTT {
	static_assert(sizeof(E) == 4, "NTT<E>: 4-byte field elements (BB31)");

public:
	explicit NTT(const NTTConfRad2<E>& nttconf) : ntt_conf(nttconf) {
		int device = 0;
		cudaGetDevice(&device);  // the caller's current device, like the reference's cudaMalloc'd members (gpuntt.cuh:134-148)
		const int rc = b2ntt_bb31_create(&h, ntt_conf.generator.asRaw(), (int)ntt_conf.log_group_order, ntt_conf.log_inp_size, device);
		if (rc != B2NTT_OK) {
			std::cerr << "NTT: " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
			h = nullptr;
		}
	}
	NTT(const NTT&) = delete;
	NTT& operator=(const NTT&) = delete;
	~NTT() { b2ntt_bb31_destroy(h); }

	// gpuntt.cuh:158-183: the input size must match (ASSERT), a BIT_REVERSED input skips the permutation, the output is IN_ORDER
	void apply(const NTTData<E>& input, NTTData<E>& output) {
		const size_t n = (size_t)1 << ntt_conf.log_inp_size;
		ASSERT(n == input.size);
		ASSERT(output.size == n);
		if (!h) return;
		const int rc = b2ntt_bb31_exec_host(h, input.data.get(), input.size, input.order == DataOrder::BIT_REVERSED ? 1 : 0, output.data.get());
		if (rc != B2NTT_OK) std::cerr << "NTT::apply: " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
		output.order = DataOrder::IN_ORDER;
	}

private:
	NTTConfRad2<E> ntt_conf;
	b2ntt_bb31* h = nullptr;
};
