// ulvt/utils/bitslicing.cuh -- drop-in for BitsliceUtils<W>, transpose_kernel<W>, untranspose_kernel<W>
// (utils/bitslicing.cuh:8-104): compact <-> little-endian bit-sliced, blocks of 32 consecutive
// elements of W bits; afterwards word 32*k + b holds bit b of limb k of the 32 elements, element e at
// bit position e (:32-47).  The per-block routines below are for host code and small device use; bulk
// device conversion goes through b2ntt_bitslice (include/b2ntt.h), which the two kernels' launchers
// `bitslice_device` / `unbitslice_device` forward to.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>

#include "../../../binius-ntt_b200/csrc/b2_bundle.cuh"
#include "../../b2ntt.h"

template <size_t BITSLICING_BITS_WIDTH>
class BitsliceUtils {
	static_assert(BITSLICING_BITS_WIDTH % 32 == 0, "width is a multiple of 32 bits");
	static constexpr size_t LIMBS = BITSLICING_BITS_WIDTH / 32;
	std::array<uint32_t, BITSLICING_BITS_WIDTH>& buffer;

public:
	explicit BitsliceUtils(std::array<uint32_t, BITSLICING_BITS_WIDTH>& data) : buffer(data) {}

	// compact (element e = words [e*LIMBS, (e+1)*LIMBS)) -> bit-sliced, in place
	B2_HD static void bitslice_transpose(uint32_t* block) {
		uint32_t t[BITSLICING_BITS_WIDTH];
		for (size_t k = 0; k < LIMBS; k++)
			for (size_t e = 0; e < 32; e++) t[32 * k + e] = block[e * LIMBS + k];
		for (size_t k = 0; k < LIMBS; k++) b2::transpose32(t + 32 * k);
		for (size_t i = 0; i < BITSLICING_BITS_WIDTH; i++) block[i] = t[i];
	}
	// bit-sliced -> compact, in place
	B2_HD static void bitslice_untranspose(uint32_t* block) {
		uint32_t t[BITSLICING_BITS_WIDTH];
		for (size_t i = 0; i < BITSLICING_BITS_WIDTH; i++) t[i] = block[i];
		for (size_t k = 0; k < LIMBS; k++) b2::transpose32(t + 32 * k);
		for (size_t k = 0; k < LIMBS; k++)
			for (size_t e = 0; e < 32; e++) block[e * LIMBS + k] = t[32 * k + e];
	}
	// the same value in all 32 lanes (:66-74)
	static void repeat_value_bitsliced(uint32_t* batch, const uint32_t* value) {
		for (size_t i = 0; i < BITSLICING_BITS_WIDTH; i++) batch[i] = value[i % LIMBS];
		bitslice_transpose(batch);
	}
	void bitslice_transpose() { bitslice_transpose(buffer.data()); }
	// NB the reference's member bitslice_untranspose() calls the forward transpose (:78); kept callable, done right
	void bitslice_untranspose() { bitslice_untranspose(buffer.data()); }
	void repeat_value_bitsliced(const std::array<uint32_t, LIMBS>& value) { repeat_value_bitsliced(buffer.data(), value.data()); }
};

#if defined(__CUDACC__)
// same names and launch contract as utils/bitslicing.cuh:89-104 (grid-stride over the blocks)
template <size_t BITSLICING_BITS_WIDTH>
__global__ static void transpose_kernel(uint32_t* buffer, const uint32_t batches_in_buffer) {
	for (uint32_t b = threadIdx.x + blockIdx.x * blockDim.x; b < batches_in_buffer; b += gridDim.x * blockDim.x)
		BitsliceUtils<BITSLICING_BITS_WIDTH>::bitslice_transpose(buffer + BITSLICING_BITS_WIDTH * (size_t)b);
}
template <size_t BITSLICING_BITS_WIDTH>
__global__ static void untranspose_kernel(uint32_t* buffer, const uint32_t batches_in_buffer) {
	for (uint32_t b = threadIdx.x + blockIdx.x * blockDim.x; b < batches_in_buffer; b += gridDim.x * blockDim.x)
		BitsliceUtils<BITSLICING_BITS_WIDTH>::bitslice_untranspose(buffer + BITSLICING_BITS_WIDTH * (size_t)b);
}
#endif

// bulk device conversion: the coalesced warp-cooperative kernels of libb2ntt.so
template <size_t BITSLICING_BITS_WIDTH>
inline bool bitslice_device(uint32_t* d_buffer, size_t batches, void* stream = nullptr) {
	return b2ntt_bitslice(d_buffer, batches, (int)BITSLICING_BITS_WIDTH, 0, stream) == B2NTT_OK;
}
template <size_t BITSLICING_BITS_WIDTH>
inline bool unbitslice_device(uint32_t* d_buffer, size_t batches, void* stream = nullptr) {
	return b2ntt_bitslice(d_buffer, batches, (int)BITSLICING_BITS_WIDTH, 1, stream) == B2NTT_OK;
}
