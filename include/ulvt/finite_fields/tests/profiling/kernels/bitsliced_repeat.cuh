// drop-in for finite_fields/tests/profiling/kernels/bitsliced_repeat.cuh: the dependent-chain bit-sliced GF(2^32) multiply
// microbenchmark kernel exported by ulvt_gpu (bitsliced_repeat.cu:6-33; launched by profiling/gpu_benchmarks.cu:8-21 as
// <<<256, 256>>>).  Defined in libulvt_gpu.a / libb2ntt.so (binius-ntt_b200/csrc/b2_compat.cu).
#pragma once
#include <cstdint>

__global__ void bitsliced_repeat(uint32_t* x, uint32_t* y, uint32_t* dst);
