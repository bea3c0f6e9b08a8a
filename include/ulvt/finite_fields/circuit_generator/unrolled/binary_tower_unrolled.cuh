// drop-in for finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh:4-5:
// multiply_unrolled<HEIGHT>(a, b, destination) -- 32 independent tower products in bit-sliced form,
// 2^HEIGHT words per operand, word i = bit i of the 32 elements.  `destination` may alias an input
// (the reference's generated code loads everything before it stores; sumcheck/core/core.cu:21 relies
// on it).  Here the circuit is the template recursion b2::Sliced<H> (b2_field.cuh), which the compiler
// unrolls into LOP3 networks in registers, instead of 4 generated .cu files.
#pragma once
#include <cstdint>

#include "../../../../../binius-ntt_b200/csrc/b2_field.cuh"

template <int HEIGHT>
B2_HD void multiply_unrolled(const uint32_t* field_element_a, const uint32_t* field_element_b, uint32_t* destination) {
	static_assert(HEIGHT >= 1 && HEIGHT <= 7, "multiply_unrolled: HEIGHT 1..7");
	uint32_t product[1 << HEIGHT];
	b2::Sliced<HEIGHT>::mul(field_element_a, field_element_b, product);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
	for (int i = 0; i < (1 << HEIGHT); i++) destination[i] = product[i];
}
