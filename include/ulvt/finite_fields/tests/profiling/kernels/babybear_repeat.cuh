// drop-in for finite_fields/tests/profiling/kernels/babybear_repeat.cuh: dependent-chain BabyBear multiply microbenchmark
// kernel exported by ulvt_gpu (babybear_repeat.cu:3-14).  Defined in libulvt_gpu.a / libb2ntt.so (csrc/b2_compat.cu).
#pragma once
#include "../../../baby_bear.cuh"

__global__ void babybear_repeat(BB31* operands, BB31* results);
