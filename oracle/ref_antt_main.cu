/*
 * ref_antt_main.cu -- harness (ours) around the REFERENCE's own CUDA additive NTT
 * (ntt/additive_ntt.cuh, ntt/nttconf.cu, utils/common.cu compiled from /root/reference, never
 * copied) for sm_100a.  Output binary: oracle/_ref/ref_antt.  It is the "reference's own CUDA
 * kernels" baseline of BASELINE.md section 5 and is run only by bench.py / tests as a baseline.
 * Catch2 is absent, so this replaces ntt/tests/test_ntt.cu:191-217 with a plain main.
 *
 * usage: ref_antt <log_h> <log_rate> [reps]
 * prints one JSON line: md5 of the output, whole-apply() wall ms (as ntt/tests/benchmark_antt.cu:92-97
 * times it, PCIe included) and kernels-only ms (CUDA events around the launch loop of
 * additive_ntt.cuh:233-247, re-issued here on the object's own device buffers).
 */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <random>

#define private public
#include "ulvt/finite_fields/binary_tower.cuh"
#include "ulvt/ntt/additive_ntt.cuh"
#undef private

extern "C" void orc_md5_hex(const void* data, size_t len, char hex[33]);

using P = FanPaarTowerField<5>;

int main(int argc, char** argv) {
	if (argc < 3) {
		fprintf(stderr, "usage: %s log_h log_rate [reps]\n", argv[0]);
		return 2;
	}
	int log_h = atoi(argv[1]), log_rate = atoi(argv[2]);
	int reps = argc > 3 ? atoi(argv[3]) : 5;
	size_t n = (size_t)1 << log_h, m = (size_t)1 << (log_h + log_rate);
	std::mt19937 gen(0xdeadbeef + log_h + log_rate);
	NTTData<uint32_t> in(DataOrder::IN_ORDER, n);
	for (size_t i = 0; i < n; i++) in.data[i] = gen();
	AdditiveNTTConf<uint32_t, P> conf(log_h, log_rate);
	AdditiveNTT<uint32_t, P> ntt(conf);
	NTTData<uint32_t> out(m);
	if (!ntt.apply(in, out)) return 3; /* warm-up + result */
	char hex[33];
	orc_md5_hex(out.data.get(), m * 4, hex);

	double apply_ms = 1e30;
	for (int r = 0; r < reps; r++) {
		auto t0 = std::chrono::high_resolution_clock::now();
		ntt.apply(in, out);
		cudaDeviceSynchronize();
		auto t1 = std::chrono::high_resolution_clock::now();
		double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
		if (ms < apply_ms) apply_ms = ms;
	}

	/* kernels only: same launch loop as apply(), on the data left in the device buffer */
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	const int msk = MAX_STAGES_PER_KERNEL - log_rate;
	AdditiveNTTKernelParams<uint32_t> kp;
	kp.data_io = ntt.data_in_out;
	kp.data_pitch = ntt.out_pitch;
	kp.constants = ntt.pre_computed;
	kp.constants_pitch = ntt.constants_pitch;
	kp.log_h = log_h;
	kp.log_rate = log_rate;
	auto [lc, nk] = ntt.ntt_conf.get_kernel_launch_confs();
	float kern_ms = 1e30f;
	for (int r = 0; r < reps + 1; r++) {
		cudaEventRecord(e0);
		for (int kern = nk - 1; kern >= 0; kern--) {
			kp.start_stage = kern * msk;
			kp.end_stage = std::min(log_h, msk * (kern + 1));
			additive_ntt_kernel<uint32_t, P><<<lc[kern][1], lc[kern][0]>>>(kp);
		}
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (r > 0 && ms < kern_ms) kern_ms = ms;
	}
	cudaError_t err = cudaGetLastError();
	printf("{\"impl\": \"reference-cuda\", \"log_h\": %d, \"log_rate\": %d, \"md5\": \"%s\", \"apply_ms\": %.4f, "
	       "\"kernels_ms\": %.4f, \"kernels\": %d, \"cuda_error\": \"%s\"}\n",
	       log_h, log_rate, hex, apply_ms, kern_ms, nk, cudaGetErrorString(err));
	return 0;
}
