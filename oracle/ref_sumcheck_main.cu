// ref_sumcheck_main.cu -- plain main around the REFERENCE's own Sumcheck<NUM_VARS, COMPOSITION_SIZE, false>
// (sumcheck/sumcheck.cuh, compiled from /root/reference where it lies; see oracle/Makefile) for the B200 box:
// seeded inputs and challenges, prints every round's messages and the wall time of the protocol.
// TEST / BASELINE INFRASTRUCTURE ONLY.  usage: ref_sumcheck <num_vars: 20|24|28> <composition_size: 2|3|4> [quiet]
#include <array>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "ulvt/sumcheck/sumcheck.cuh"

template <uint32_t NV, uint32_t C>
static int run(bool quiet) {
	const size_t words = (size_t)INTS_PER_VALUE * ((size_t)1 << NV) * C;
	std::vector<uint32_t> evals(words);
	std::mt19937 gen(0xdeadbeefu + NV + C);
	for (auto& w : evals) w = (uint32_t)gen();
	std::mt19937 chal(0xc0ffeeu + NV + C);
	const auto t0 = std::chrono::high_resolution_clock::now();
	Sumcheck<NV, C, false> s(evals, true);
	const auto t1 = std::chrono::high_resolution_clock::now();
	for (uint32_t round = 0; round <= NV; ++round) {
		std::array<uint32_t, INTS_PER_VALUE> sum;
		std::array<uint32_t, (C + 1) * INTS_PER_VALUE> points;
		s.this_round_messages(sum, points);
		if (!quiet) {
			printf("round %u sum %08x %08x %08x %08x points", round, sum[0], sum[1], sum[2], sum[3]);
			for (auto p : points) printf(" %08x", p);
			printf("\n");
		}
		if (round < NV) {
			std::array<uint32_t, INTS_PER_VALUE> ch;
			for (auto& w : ch) w = (uint32_t)chal();
			s.move_to_next_round(ch);
		}
	}
	const auto t2 = std::chrono::high_resolution_clock::now();
	printf("{\"impl\": \"reference CUDA sumcheck\", \"num_vars\": %u, \"composition_size\": %u, \"upload_transpose_ms\": %.3f, \"protocol_ms\": %.3f}\n", NV, C,
	       std::chrono::duration<double, std::milli>(t1 - t0).count(), std::chrono::duration<double, std::milli>(t2 - t1).count());
	return 0;
}

int main(int argc, char** argv) {
	const int nv = argc > 1 ? atoi(argv[1]) : 20, c = argc > 2 ? atoi(argv[2]) : 2;
	const bool quiet = argc > 3;
#define CASE(NV, C) \
	if (nv == NV && c == C) return run<NV, C>(quiet);
	CASE(20, 2) CASE(20, 3) CASE(20, 4) CASE(24, 2) CASE(24, 3) CASE(24, 4) CASE(28, 2) CASE(28, 3) CASE(28, 4)
	fprintf(stderr, "unsupported (num_vars, composition_size)\n");
	return 2;
}
