// shim_test.cu -- C++ parity test of the drop-in surface include/ulvt/... over libb2ntt.so.
// TEST CODE.  It reads like the reference's own tests (ntt/tests/test_ntt.cu:191-217 additive cases,
// finite_fields/tests/test_fanpaartower.cu, tests.cu) with Catch2 replaced by a small CHECK macro,
// and uses the oracle (oracle/oracle.h: golden MD5 table, MD5, CPU field restatement) as checker.
//   usage: shim_test [max_log_h_r0 [max_log_h_r2]]     exit code 0 = all checks passed
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../../oracle/oracle.h"
#include "ulvt/finite_fields/binary_tower.cuh"
#include "ulvt/finite_fields/binary_tower_simd.cuh"
#include "ulvt/finite_fields/circuit_generator/utils/utils.hpp"
#include "ulvt/finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh"
#include "ulvt/ntt/additive_ntt.cuh"
#include "ulvt/sumcheck/sumcheck.cuh"
#include "ulvt/utils/bitslicing.cuh"

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                          \
	do {                                                                     \
		g_checks++;                                                          \
		if (!(cond)) {                                                       \
			g_failed++;                                                      \
			printf("CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond);   \
		}                                                                    \
	} while (0)

// test_ntt.cu:191-217
static bool run_and_check_additive_ntt(int log_h, int log_rate) {
	std::mt19937 gen(0xdeadbeef + log_h + log_rate);
	const size_t inp_size = (size_t)1 << log_h, out_size = (size_t)1 << (log_h + log_rate);
	NTTData<uint32_t> ntt_inp(DataOrder::IN_ORDER, inp_size);
	for (size_t i = 0; i < inp_size; i++) ntt_inp.data[i] = gen();
	AdditiveNTTConf<uint32_t, FanPaarTowerField<5>> nttconf(log_h, log_rate);
	AdditiveNTT<uint32_t, FanPaarTowerField<5>> add_ntt(nttconf);
	NTTData<uint32_t> ntt_out(out_size);
	if (!add_ntt.apply(ntt_inp, ntt_out)) return false;
	char hex[33];
	orc_md5_hex(ntt_out.data.get(), ntt_out.byte_len(), hex);
	const char* want = orc_golden_md5(log_rate, log_h);
	return want && strcmp(hex, want) == 0 && ntt_out.order == DataOrder::IN_ORDER;
}

// limb l of a GF(2^128) transform is the GF(2^32) transform of limb l: limb 0 carries the reference stream
static bool run_and_check_additive_ntt128(int log_h, int log_rate) {
	std::mt19937 gen(0xdeadbeef + log_h + log_rate), other(12345 + log_h);
	const size_t inp_size = (size_t)1 << log_h, out_size = (size_t)1 << (log_h + log_rate);
	NTTData<b128> in(DataOrder::IN_ORDER, inp_size);
	for (size_t i = 0; i < inp_size; i++) {
		in.data[i].w[0] = gen();
		for (int l = 1; l < 4; l++) in.data[i].w[l] = other();
	}
	AdditiveNTTConf<b128, FanPaarTowerField<7>> conf(log_h, log_rate);
	AdditiveNTT<b128, FanPaarTowerField<7>> ntt(conf);
	NTTData<b128> out(out_size);
	if (!ntt.apply(in, out)) return false;
	std::vector<uint32_t> limb0(out_size);
	for (size_t i = 0; i < out_size; i++) limb0[i] = out.data[i].w[0];
	char hex[33];
	orc_md5_hex(limb0.data(), limb0.size() * 4, hex);
	const char* want = orc_golden_md5(log_rate, log_h);
	return want && strcmp(hex, want) == 0;
}

// general-basis transform through the C++ shim against the oracle's restatement (oracle/antt_general.c)
static bool run_and_check_general_ntt(int log_h, bool reference_basis) {
	std::mt19937 gen(777 + log_h);
	const size_t n = (size_t)1 << log_h;
	std::vector<b128> basis(log_h);
	for (int k = 0; k < log_h; k++)
		for (int l = 0; l < 4; l++) basis[k].w[l] = gen();  // independent with probability ~ 1 - 2^(log_h - 128)
	const b128* bp = reference_basis ? nullptr : basis.data();
	NTTData<b128> in(DataOrder::IN_ORDER, n), out(n), back(n);
	for (size_t i = 0; i < n; i++)
		for (int l = 0; l < 4; l++) in.data[i].w[l] = gen();
	GeneralAdditiveNTT ntt(log_h, bp);
	if (!ntt.valid() || !ntt.apply(in, out)) return false;
	std::vector<uint32_t> want(4 * n);
	if (orc_antt_general(reinterpret_cast<const uint32_t*>(in.data.get()), want.data(), log_h, reinterpret_cast<const uint32_t*>(bp), 0)) return false;
	if (memcmp(want.data(), out.data.get(), 16 * n)) return false;
	if (reference_basis) {  // equals the stock class
		AdditiveNTTConf<b128, FanPaarTowerField<7>> conf(log_h, 0);
		AdditiveNTT<b128, FanPaarTowerField<7>> stock(conf);
		NTTData<b128> o2(n);
		if (!stock.apply(in, o2) || memcmp(o2.data.get(), out.data.get(), 16 * n)) return false;
	}
	out.order = DataOrder::IN_ORDER;
	return ntt.apply(out, back, true) && memcmp(back.data.get(), in.data.get(), 16 * n) == 0;
}

// sumcheck/test/test.cu:16-96 (test_sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED>) with the verifier of
// sumcheck/test/verifier.cu restated by the oracle; seeded instead of time-seeded
template <uint32_t NUM_VARS, uint32_t COMPOSITION_SIZE, bool DATA_IS_TRANSPOSED>
static void test_sumcheck() {
	constexpr uint32_t interpolation_points = COMPOSITION_SIZE + 1;
	const size_t num_ints_in_evals = INTS_PER_VALUE * ((size_t)1 << NUM_VARS) * COMPOSITION_SIZE;
	std::mt19937 gen(NUM_VARS * 10 + COMPOSITION_SIZE);
	std::vector<uint32_t> multilinear_evals(num_ints_in_evals), compact(num_ints_in_evals);
	for (auto& w : compact) w = gen();
	multilinear_evals = compact;
	if (DATA_IS_TRANSPOSED)
		for (size_t b = 0; b < num_ints_in_evals; b += 128) orc_bitslice(multilinear_evals.data() + b, 128);
	Sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED> s(multilinear_evals, false);
	std::vector<uint32_t> challenges(4 * NUM_VARS);
	std::array<uint32_t, INTS_PER_VALUE> expected_claim{};
	for (uint32_t round = 0; round < NUM_VARS; ++round) {
		std::array<uint32_t, INTS_PER_VALUE> sum;
		std::array<uint32_t, interpolation_points * INTS_PER_VALUE> points;
		s.this_round_messages(sum, points);
		CHECK(round == 0 || sum == expected_claim);
		for (int i = 0; i < 4; i++) CHECK(sum[i] == (points[i] ^ points[4 + i]));
		std::array<uint32_t, INTS_PER_VALUE> challenge;
		for (auto& w : challenge) w = gen();
		memcpy(&challenges[4 * round], challenge.data(), 16);
		orc_sc_interpolate(challenge.data(), points.data(), interpolation_points, expected_claim.data());
		s.move_to_next_round(challenge);
	}
	std::array<uint32_t, INTS_PER_VALUE> sum, brute;
	std::array<uint32_t, interpolation_points * INTS_PER_VALUE> points;
	s.this_round_messages(sum, points);
	CHECK(sum == expected_claim);
	orc_sc_eval_composition(compact.data(), challenges.data(), NUM_VARS, COMPOSITION_SIZE, brute.data());
	CHECK(brute == expected_claim);
}

template <int H>
__global__ void mul_unrolled_kernel(const uint32_t* a, const uint32_t* b, uint32_t* o) {
	multiply_unrolled<H>(a + (size_t)threadIdx.x * (1 << H), b + (size_t)threadIdx.x * (1 << H), o + (size_t)threadIdx.x * (1 << H));
}

template <int H>
static void check_unrolled_on_device(std::mt19937& gen) {
	constexpr int N = 1 << H, T = 8;
	std::vector<uint32_t> a(N * T), b(N * T), got(N * T), want(N * T);
	for (auto& x : a) x = gen();
	for (auto& x : b) x = gen();
	for (int t = 0; t < T; t++) orc_mul_bitsliced(&a[t * N], &b[t * N], &want[t * N], H);
	uint32_t *da, *db;
	cudaMalloc(&da, N * T * 4), cudaMalloc(&db, N * T * 4);
	cudaMemcpy(da, a.data(), N * T * 4, cudaMemcpyHostToDevice);
	cudaMemcpy(db, b.data(), N * T * 4, cudaMemcpyHostToDevice);
	mul_unrolled_kernel<H><<<1, T>>>(da, db, da);  // destination aliases an input (core.cu:21)
	cudaMemcpy(got.data(), da, N * T * 4, cudaMemcpyDeviceToHost);
	CHECK(cudaGetLastError() == cudaSuccess);
	CHECK(got == want);
	std::vector<uint32_t> host(N);
	multiply_unrolled<H>(a.data(), b.data(), host.data());
	CHECK(memcmp(host.data(), want.data(), N * 4) == 0);
	cudaFree(da), cudaFree(db);
}

int main(int argc, char** argv) {
	const int max_r0 = argc > 1 ? atoi(argv[1]) : 24, max_r2 = argc > 2 ? atoi(argv[2]) : 22;
	CHECK(check_gpu_capabilities());  // test_ntt.cu:189

	// ---- additive NTT against the reference's golden digests (test_ntt.cu:52-124) ----
	for (int log_h = 1; log_h <= max_r0; log_h++) CHECK(run_and_check_additive_ntt(log_h, 0));
	for (int log_h = 1; log_h <= max_r2; log_h++) CHECK(run_and_check_additive_ntt(log_h, 2));
	for (int log_h : {3, 9, 16, 20}) CHECK(run_and_check_additive_ntt128(log_h, 0));
	CHECK(run_and_check_additive_ntt128(14, 2));
	for (int log_h : {1, 5, 10, 11, 15}) CHECK(run_and_check_general_ntt(log_h, false));
	for (int log_h : {4, 12}) CHECK(run_and_check_general_ntt(log_h, true));
	{  // a dependent basis is refused
		b128 dep[3] = {};
		dep[0].w[0] = 5, dep[1].w[2] = 9, dep[2].w[0] = 5, dep[2].w[2] = 9;
		GeneralAdditiveNTT bad(3, dep);
		CHECK(!bad.valid());
	}
	{  // apply(): false on wrong size / order, nothing else (additive_ntt.cuh:206-208)
		AdditiveNTTConf<uint32_t, FanPaarTowerField<5>> conf(10, 0);
		AdditiveNTT<uint32_t, FanPaarTowerField<5>> ntt(conf);
		NTTData<uint32_t> bad_size(DataOrder::IN_ORDER, 1000), bad_order(DataOrder::BIT_REVERSED, 1024), out(1024);
		CHECK(!ntt.apply(bad_size, out));
		CHECK(!ntt.apply(bad_order, out));
	}

	// ---- compact tower field against the oracle (test_fanpaartower.cu:55-120 style) ----
	std::mt19937 gen(2024);
	for (int i = 0; i < 2000; i++) {
		const uint32_t a = gen(), b = gen();
		CHECK(FanPaarTowerField<5>::multiply(a, b) == orc_mul32(a, b));
		CHECK(FanPaarTowerField<5>::square(a) == orc_mul32(a, a));
		if (a) CHECK(FanPaarTowerField<5>::multiply(a, FanPaarTowerField<5>::inverse(a)) == 1u);
		CHECK(FanPaarTowerField<4>::multiply(a & 0xffff, b & 0xffff) == orc_mul32(a & 0xffff, b & 0xffff));
		CHECK(FanPaarTowerField<3>::multiply(a & 0xff, b & 0xff) == orc_mul32(a & 0xff, b & 0xff));
		const uint32_t r[6] = {(uint32_t)gen(), (uint32_t)gen(), (uint32_t)gen(), (uint32_t)gen(), (uint32_t)gen(), (uint32_t)gen()};
		b128 x{{a, b, r[0], r[1]}}, y{{r[2], r[3], r[4], r[5]}}, z = FanPaarTowerField<7>::multiply(x, y);
		uint32_t w[4];
		orc_mul128_words(x.w, y.w, w);
		CHECK(memcmp(z.w, w, 16) == 0);
	}
	CHECK(FanPaarTowerField<5>::N_BITS() == 32 && FanPaarTowerField<3>::is_valid(0xff) && !FanPaarTowerField<3>::is_valid(0x100));

	// ---- packed lanes (tests.cu:17-91 style) and the hex helpers (tests.cu:186-188) ----
	for (int i = 0; i < 500; i++) {
		const uint32_t a = gen(), b = gen();
		CHECK(mul_binary_tower_32b_simd<0>(a, b) == orc_mul_packed32(a, b, 0));
		CHECK(mul_binary_tower_32b_simd<2>(a, b) == orc_mul_packed32(a, b, 2));
		CHECK(mul_binary_tower_32b_simd<3>(a, b) == orc_mul_packed32(a, b, 3));
		CHECK(mul_binary_tower_32b_simd<4>(a, b) == orc_mul_packed32(a, b, 4));
		CHECK(mul_binary_tower_32b_simd<5>(a, b) == orc_mul32(a, b));
		uint32_t x = a, y = b;
		orc_interleave32(&x, &y, 3);
		const auto cd = interleave_32b<3>(a, b);
		CHECK(cd.first == x && cd.second == y);
		orc_interleave32(&(x = a), &(y = b), 0);
		const auto cd0 = interleave_32b<0>(a, b);
		CHECK(cd0.first == x && cd0.second == y);
		const uint32_t xa = xor_adjacent_32b<2>(a);
		CHECK(((xa >> 4) & 0x0f0f0f0fu) == (xa & 0x0f0f0f0fu) && (xa & 0x0f0f0f0fu) == ((a ^ (a >> 4)) & 0x0f0f0f0fu));
	}
	{
		uint32_t w[4] = {0, 0, 0, 0};
		write_string_to_int_arr(w, "0x0123456789abcdef0011223344556677");
		CHECK(w[0] == 0x44556677u && w[1] == 0x00112233u && w[2] == 0x89abcdefu && w[3] == 0x01234567u);
		CHECK(int_to_hex(0xbeefu) == "beef");
		CHECK(sizeof(KERNEL_LAUNCH_CONFIGS) == 31 * 3 * 2 * sizeof(dim3));
	}

	// ---- sumcheck (sumcheck/test/test.cu:98-142 at sizes the brute-force check finishes quickly) ----
	test_sumcheck<12, 2, true>();
	test_sumcheck<12, 3, false>();
	test_sumcheck<13, 4, true>();
	test_sumcheck<14, 2, false>();

	// ---- bit-slicing and the bit-sliced products (tests.cu:125-201, test_fanpaartower.cu:127-273 style) ----
	{
		std::array<uint32_t, 128> blk, ref;
		for (auto& x : blk) x = gen();
		ref = blk;
		orc_bitslice(ref.data(), 128);
		BitsliceUtils<128> u(blk);
		u.bitslice_transpose();
		CHECK(blk == ref);
		uint32_t* d;
		std::array<uint32_t, 128> dev = ref;
		cudaMalloc(&d, 512);
		cudaMemcpy(d, dev.data(), 512, cudaMemcpyHostToDevice);
		untranspose_kernel<128><<<1, 32>>>(d, 1);
		cudaMemcpy(dev.data(), d, 512, cudaMemcpyDeviceToHost);
		BitsliceUtils<128>::bitslice_untranspose(blk.data());
		CHECK(dev == blk);
		CHECK(bitslice_device<128>(d, 1));
		cudaMemcpy(dev.data(), d, 512, cudaMemcpyDeviceToHost);
		CHECK(dev == ref);
		cudaFree(d);
	}
	check_unrolled_on_device<2>(gen);
	check_unrolled_on_device<5>(gen);
	check_unrolled_on_device<6>(gen);
	check_unrolled_on_device<7>(gen);

	printf("shim_test: %d checks, %d failed\n", g_checks, g_failed);
	return g_failed ? 1 : 0;
}
