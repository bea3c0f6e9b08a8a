// ulvt/finite_fields/binary_tower_simd.cuh -- drop-in for the packed-lane interface of
// finite_fields/binary_tower_simd.cuh:11-188: 32 / 2^H independent GF(2^(2^H)) elements packed in one 32-bit
// word (lane j = bits [j * 2^H, (j+1) * 2^H)).
//   mul_binary_tower_32b_simd<H>(a, b)  lane-wise tower product            (:77-127)
//   interleave_32b<H>(a, b)             swap the odd 2^H-bit blocks of a with the even blocks of b (:129-139)
//   xor_adjacent_32b<H>(a)              every block pair replaced by the XOR of its two blocks, in both (:141-149)
//   PackedBinaryField<LOG_BITS>, interleave<H, LOG_BITS>(a, b)             (:20-33, :151-188; word-wise for H < 5)
// The reference uses this path only in tests and smoke kernels; the product here is the plain per-lane
// recursion b2::Tower<H> (b2_field.cuh), not the reference's interleaved Karatsuba, with identical results.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda/std/utility>

#include "../../../binius-ntt_b200/csrc/b2_field.cuh"

namespace b2_simd_detail {
B2_HD uint32_t even_blocks(int h) { return h == 0 ? 0x55555555u : h == 1 ? 0x33333333u : h == 2 ? 0x0f0f0f0fu : h == 3 ? 0x00ff00ffu : 0x0000ffffu; }
}  // namespace b2_simd_detail

template <size_t HEIGHT>
B2_HD uint32_t mul_binary_tower_32b_simd(uint32_t a, uint32_t b) {
	static_assert(HEIGHT <= 5, "one 32-bit word holds lanes of at most 32 bits");
	if (HEIGHT == 0) return a & b;
	constexpr int W = 1 << HEIGHT;
	const uint32_t lane_mask = W == 32 ? 0xffffffffu : ((1u << (W & 31)) - 1u);
	uint32_t out = 0;
	for (int s = 0; s < 32; s += W) out |= (uint32_t)b2::Tower<HEIGHT>::mul((a >> s) & lane_mask, (b >> s) & lane_mask) << s;
	return out;
}

// (c, d): c keeps a's even blocks and takes b's even blocks into its odd positions; d takes a's odd blocks into its
// even positions and keeps b's odd blocks
template <size_t HEIGHT>
B2_HD cuda::std::pair<uint32_t, uint32_t> interleave_32b(uint32_t a, uint32_t b) {
	static_assert(HEIGHT < 5, "interleave_32b: blocks of at most 16 bits");
	const uint32_t ev = b2_simd_detail::even_blocks((int)HEIGHT);
	const int len = 1 << HEIGHT;
	const uint32_t c = (a & ev) | ((b & ev) << len);
	const uint32_t d = ((a >> len) & ev) | (b & ~ev);
	return cuda::std::make_pair(c, d);
}

template <size_t HEIGHT>
B2_HD uint32_t xor_adjacent_32b(uint32_t a) {
	static_assert(HEIGHT < 5, "xor_adjacent_32b: blocks of at most 16 bits");
	const uint32_t ev = b2_simd_detail::even_blocks((int)HEIGHT);
	const int len = 1 << HEIGHT;
	const uint32_t s = (a ^ (a >> len)) & ev;
	return s | (s << len);
}

template <size_t LOG_BITS>
struct PackedBinaryField {
	static_assert(LOG_BITS >= 5, "at least one 32-bit word");
	uint32_t words[1 << (LOG_BITS - 5)];
	B2_HD bool operator==(const PackedBinaryField& o) const {
		for (size_t i = 0; i < (size_t(1) << (LOG_BITS - 5)); i++)
			if (words[i] != o.words[i]) return false;
		return true;
	}
};

template <size_t HEIGHT, size_t LOG_BITS>
B2_HD cuda::std::pair<PackedBinaryField<LOG_BITS>, PackedBinaryField<LOG_BITS>> interleave(const PackedBinaryField<LOG_BITS>& a, const PackedBinaryField<LOG_BITS>& b) {
	static_assert(HEIGHT < 5, "interleave: word-wise form (blocks of at most 16 bits)");
	PackedBinaryField<LOG_BITS> c, d;
	for (size_t i = 0; i < (size_t(1) << (LOG_BITS - 5)); i++) {
		const auto cd = interleave_32b<HEIGHT>(a.words[i], b.words[i]);
		c.words[i] = cd.first, d.words[i] = cd.second;
	}
	return cuda::std::make_pair(c, d);
}
