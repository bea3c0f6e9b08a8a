/* ref_shim128.cu -- the reference's compact GF(2^128) product (sumcheck/test/utils/tower_7_mul.cu),
 * in its own translation unit because unbitsliced_mul.cuh and binary_tower.cuh both define
 * FanPaarTowerField.  TEST INFRASTRUCTURE. */
#include <cstdint>
#include "ulvt/sumcheck/test/utils/tower_7_mul.cuh"
extern "C" void ref_mul128_words(const uint32_t* a, const uint32_t* b, uint32_t* out) {
	__uint128_t x = 0, y = 0;
	for (int i = 3; i >= 0; i--) {
		x = (x << 32) | a[i];
		y = (y << 32) | b[i];
	}
	__uint128_t z = tower_height_7_mul(x, y);
	for (int i = 0; i < 4; i++) out[i] = (uint32_t)(z >> (32 * i));
}
