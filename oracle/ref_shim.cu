/*
 * ref_shim.cu -- C exports over the REFERENCE's own host-callable field code, compiled from the
 * sources where they lie under /root/reference (never copied).  Output: oracle/_ref/libulvt_ref.so
 * (git-ignored).  Used only by tests/ to validate the oracle restatement.  TEST INFRASTRUCTURE.
 */
#include <cstdint>
#include <cstring>

#include "ulvt/finite_fields/binary_tower.cuh"
#include "ulvt/finite_fields/binary_tower_simd.cuh"
#include "ulvt/finite_fields/circuit_generator/unrolled/binary_tower_unrolled.cuh"
#include "ulvt/utils/bitslicing.cuh"

extern "C" {
uint32_t ref_mul(uint32_t a, uint32_t b, int h) {
	switch (h) {
		case 0: return FanPaarTowerField<0>::multiply(a, b);
		case 1: return FanPaarTowerField<1>::multiply(a, b);
		case 2: return FanPaarTowerField<2>::multiply(a, b);
		case 3: return FanPaarTowerField<3>::multiply(a, b);
		case 4: return FanPaarTowerField<4>::multiply(a, b);
		default: return FanPaarTowerField<5>::multiply(a, b);
	}
}
uint32_t ref_square(uint32_t a, int h) {
	switch (h) {
		case 1: return FanPaarTowerField<1>::square(a);
		case 2: return FanPaarTowerField<2>::square(a);
		case 3: return FanPaarTowerField<3>::square(a);
		case 4: return FanPaarTowerField<4>::square(a);
		default: return FanPaarTowerField<5>::square(a);
	}
}
uint32_t ref_inverse(uint32_t a, int h) {
	switch (h) {
		case 2: return FanPaarTowerField<2>::inverse(a);
		case 3: return FanPaarTowerField<3>::inverse(a);
		case 4: return FanPaarTowerField<4>::inverse(a);
		default: return FanPaarTowerField<5>::inverse(a);
	}
}
uint32_t ref_mul_alpha(uint32_t a, int h) {
	switch (h) {
		case 1: return FanPaarTowerField<1>::multiply_alpha(a);
		case 2: return FanPaarTowerField<2>::multiply_alpha(a);
		case 3: return FanPaarTowerField<3>::multiply_alpha(a);
		case 4: return FanPaarTowerField<4>::multiply_alpha(a);
		default: return FanPaarTowerField<5>::multiply_alpha(a);
	}
}
uint32_t ref_mul_packed(uint32_t a, uint32_t b, int h) {
	switch (h) {
		case 0: return mul_binary_tower_32b_simd<0>(a, b);
		case 1: return mul_binary_tower_32b_simd<1>(a, b);
		case 2: return mul_binary_tower_32b_simd<2>(a, b);
		case 3: return mul_binary_tower_32b_simd<3>(a, b);
		case 4: return mul_binary_tower_32b_simd<4>(a, b);
		default: return mul_binary_tower_32b_simd<5>(a, b);
	}
}
void ref_interleave(uint32_t* a, uint32_t* b, int h) {
	cuda::std::pair<uint32_t, uint32_t> r(0, 0);
	switch (h) {
		case 0: r = interleave_32b<0>(*a, *b); break;
		case 1: r = interleave_32b<1>(*a, *b); break;
		case 2: r = interleave_32b<2>(*a, *b); break;
		case 3: r = interleave_32b<3>(*a, *b); break;
		default: r = interleave_32b<4>(*a, *b); break;
	}
	*a = r.first;
	*b = r.second;
}
void ref_mul_unrolled(const uint32_t* a, const uint32_t* b, uint32_t* out, int h) {
	switch (h) {
		case 2: multiply_unrolled<2>(a, b, out); break;
		case 5: multiply_unrolled<5>(a, b, out); break;
		case 6: multiply_unrolled<6>(a, b, out); break;
		default: multiply_unrolled<7>(a, b, out); break;
	}
}
void ref_bitslice(uint32_t* blk, int width) {
	if (width == 32) BitsliceUtils<32>::bitslice_transpose(blk);
	else if (width == 64) BitsliceUtils<64>::bitslice_transpose(blk);
	else BitsliceUtils<128>::bitslice_transpose(blk);
}
void ref_unbitslice(uint32_t* blk, int width) {
	if (width == 32) BitsliceUtils<32>::bitslice_untranspose(blk);
	else if (width == 64) BitsliceUtils<64>::bitslice_untranspose(blk);
	else BitsliceUtils<128>::bitslice_untranspose(blk);
}
}
