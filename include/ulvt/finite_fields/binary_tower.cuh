// ulvt/finite_fields/binary_tower.cuh -- drop-in for FanPaarTowerField<H> (finite_fields/binary_tower.cuh:19-262):
// the compact Fan-Paar / Wiedemann tower GF(2^(2^H)) = GF(2^(2^(H-1)))[X]/(X^2 + a_{H-1} X + 1),
// element = lo | hi << 2^(H-1).  Same static members (multiply, square, inverse, multiply_alpha, add,
// is_valid, ONE, ZERO, N_BITS), host and device, for H = 0..5 on uint32_t.  The bodies are the
// table-free recursion of b2::Tower<H> (binius-ntt_b200/csrc/b2_field.cuh), not the reference's LUTs.
// New: H = 6 on uint64_t and H = 7 on b128 (four little-endian 32-bit limbs, README.md:43; product as
// sumcheck/test/utils/tower_7_mul.cu:4-20 defines it), so that AdditiveNTT<b128, FanPaarTowerField<7>>
// has a P to name.
#pragma once
#include <cstddef>
#include <cstdint>

#include "../../../binius-ntt_b200/csrc/b2_field.cuh"

struct b128 {
	uint32_t w[4];
	B2_HD bool operator==(const b128& o) const { return w[0] == o.w[0] && w[1] == o.w[1] && w[2] == o.w[2] && w[3] == o.w[3]; }
	B2_HD bool operator!=(const b128& o) const { return !(*this == o); }
};

template <size_t HEIGHT>
class FanPaarTowerField {
	static_assert(HEIGHT <= 5, "uint32_t elements: HEIGHT <= 5 (6 and 7 are specialised)");
	static constexpr uint32_t kMask = HEIGHT == 5 ? 0xffffffffu : ((1u << ((1u << HEIGHT) & 31)) - 1u);

public:
	static constexpr B2_HD uint32_t ONE() { return 1; }
	static constexpr B2_HD uint32_t ZERO() { return 0; }
	static constexpr B2_HD uint32_t N_BITS() { return 1u << HEIGHT; }
	static constexpr B2_HD bool is_valid(uint32_t a) { return (a & ~kMask) == 0; }
	static constexpr B2_HD uint32_t add(uint32_t a, uint32_t b) { return a ^ b; }
	static B2_HD uint32_t multiply(uint32_t a, uint32_t b) { return (uint32_t)b2::Tower<HEIGHT>::mul(a, b); }
	static B2_HD uint32_t square(uint32_t a) { return (uint32_t)b2::Tower<HEIGHT>::square(a); }
	static B2_HD uint32_t inverse(uint32_t a) { return (uint32_t)b2::Tower<HEIGHT>::inverse(a); }
	static B2_HD uint32_t multiply_alpha(uint32_t a) { return (uint32_t)b2::Tower<HEIGHT>::mul_alpha(a); }
};

template <>
class FanPaarTowerField<6> {
public:
	static constexpr B2_HD uint64_t ONE() { return 1; }
	static constexpr B2_HD uint64_t ZERO() { return 0; }
	static constexpr B2_HD uint32_t N_BITS() { return 64; }
	static constexpr B2_HD bool is_valid(uint64_t) { return true; }
	static constexpr B2_HD uint64_t add(uint64_t a, uint64_t b) { return a ^ b; }
	static B2_HD uint64_t multiply(uint64_t a, uint64_t b) { return b2::Tower<6>::mul(a, b); }
	static B2_HD uint64_t square(uint64_t a) { return b2::Tower<6>::square(a); }
	static B2_HD uint64_t inverse(uint64_t a) { return b2::Tower<6>::inverse(a); }
	static B2_HD uint64_t multiply_alpha(uint64_t a) { return b2::Tower<6>::mul_alpha(a); }
};

template <>
class FanPaarTowerField<7> {
public:
	static B2_HD b128 ONE() { return b128{{1, 0, 0, 0}}; }
	static B2_HD b128 ZERO() { return b128{{0, 0, 0, 0}}; }
	static constexpr B2_HD uint32_t N_BITS() { return 128; }
	static constexpr B2_HD bool is_valid(const b128&) { return true; }
	static B2_HD b128 add(const b128& a, const b128& b) { return b128{{a.w[0] ^ b.w[0], a.w[1] ^ b.w[1], a.w[2] ^ b.w[2], a.w[3] ^ b.w[3]}}; }
	static B2_HD b128 multiply(const b128& a, const b128& b) {
		b2::gf128 x, y;
		for (int i = 0; i < 4; i++) x.w[i] = a.w[i], y.w[i] = b.w[i];
		const b2::gf128 z = b2::gf128_mul(x, y);
		return b128{{z.w[0], z.w[1], z.w[2], z.w[3]}};
	}
	static B2_HD b128 square(const b128& a) { return multiply(a, a); }
};
