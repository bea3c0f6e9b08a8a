// ulvt/sumcheck/sumcheck.cuh -- drop-in for Sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED>
// (sumcheck/sumcheck.cuh:10-300) on top of the C ABI of libb2ntt.so (b2ntt_sumcheck_*).  Same constructor (host vector
// of COMPOSITION_SIZE columns of 2^NUM_VARS GF(2^128) evaluations, compact or bit-sliced; the `benchmarking` flag keeps
// the three time stamps), same this_round_messages / move_to_next_round.  NUM_VARS 5..30 (the reference: 20, 24, 28).
#pragma once
#include <cuda_runtime.h>

#include <array>
#include <chrono>
#include <cstdint>
#include <iostream>
#include <vector>

#include "../../b2ntt.h"
#include "utils/constants.hpp"

template <uint32_t NUM_VARS, uint32_t COMPOSITION_SIZE, bool DATA_IS_TRANSPOSED>
class Sumcheck {
	static_assert(NUM_VARS >= 5 && NUM_VARS <= 30, "NUM_VARS must be 5..30");
	static_assert(COMPOSITION_SIZE >= 2 && COMPOSITION_SIZE <= 4, "COMPOSITION_SIZE must be 2, 3, or 4");
	static constexpr uint32_t INTERPOLATION_POINTS = COMPOSITION_SIZE + 1;
	b2ntt_sumcheck* sc = nullptr;

	static void report(int rc, const char* where) {
		if (rc != B2NTT_OK) std::cerr << where << ": " << b2ntt_strerror(rc) << " " << b2ntt_last_cuda_error() << std::endl;
	}

public:
	std::chrono::time_point<std::chrono::high_resolution_clock> start_before_memcpy, start_before_transpose, start_raw;

	Sumcheck(const std::vector<uint32_t>& evals_span, const bool benchmarking) {
		int device = 0;
		cudaGetDevice(&device);  // the caller's current device, as the reference's cudaMalloc calls use
		int rc = b2ntt_sumcheck_create(&sc, NUM_VARS, COMPOSITION_SIZE, device);
		report(rc, "Sumcheck");
		if (benchmarking) start_before_memcpy = start_before_transpose = std::chrono::high_resolution_clock::now();
		// upload and (for compact data) bit-slice in one call; the split of the two stamps is not observable here
		if (rc == B2NTT_OK) report(b2ntt_sumcheck_load_host(sc, evals_span.data(), DATA_IS_TRANSPOSED ? 1 : 0), "Sumcheck: load");
		if (benchmarking) start_raw = std::chrono::high_resolution_clock::now();
	}
	Sumcheck(const Sumcheck&) = delete;
	Sumcheck& operator=(const Sumcheck&) = delete;
	~Sumcheck() { b2ntt_sumcheck_destroy(sc); }

	void this_round_messages(std::array<uint32_t, INTS_PER_VALUE>& sum_span, std::array<uint32_t, INTERPOLATION_POINTS * INTS_PER_VALUE>& points_span) {
		report(b2ntt_sumcheck_round_messages(sc, sum_span.data(), points_span.data()), "Sumcheck::this_round_messages");
	}

	void move_to_next_round(const std::array<uint32_t, INTS_PER_VALUE>& challenge_span) {
		report(b2ntt_sumcheck_next_round(sc, challenge_span.data()), "Sumcheck::move_to_next_round");
	}
};
