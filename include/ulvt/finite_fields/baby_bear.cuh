// ulvt/finite_fields/baby_bear.cuh -- drop-in for finite_fields/baby_bear.cuh + risc0_baby_bear.h:40-187: `BB31 = risc0::Fp`, the
// prime field with p = 15 * 2^27 + 1 in Montgomery form (R = 2^32).  Same public surface (constants P, M, R2; construction
// from a uint32_t with wrap-around; fromRaw / asRaw / asUInt32; the arithmetic and comparison operators; pow and inv) so that
// host arrays of BB31 keep their memory image and the reference's tests (ntt/tests/test_ntt.cu:126-187,
// finite_fields/tests/test_kernels.cu:42-70) compile unchanged.  Written for this library; the transform that uses it is
// b2ntt_bb31_* (include/b2ntt.h).
#pragma once
#include <cstddef>
#include <cstdint>

#if defined(__CUDACC__)
#define B2_BB_HD __host__ __device__
#else
#define B2_BB_HD
#endif

namespace risc0 {

class Fp {
public:
	static constexpr uint32_t P = 15 * (uint32_t(1) << 27) + 1;  // the modulus
	static constexpr uint32_t M = 0x88000001;                    // P^-1 mod 2^32
	static constexpr uint32_t R2 = 1172168163;                   // 2^64 mod P

	B2_BB_HD constexpr Fp() : val(0) {}
	B2_BB_HD constexpr Fp(uint32_t v) : val(to_mont(v)) {}  // wraps modulo P
	static B2_BB_HD constexpr Fp one() { return Fp(1); }
	static B2_BB_HD constexpr Fp fromRaw(uint32_t raw) { return Fp(raw, Raw{}); }
	B2_BB_HD constexpr uint32_t asUInt32() const { return mont_mul(val, 1); }
	B2_BB_HD constexpr uint32_t asRaw() const { return val; }
	static B2_BB_HD constexpr Fp maxVal() { return Fp(P - 1); }
	static B2_BB_HD constexpr Fp invalid() { return Fp(0xffffffffu, Raw{}); }

	B2_BB_HD constexpr void operator=(uint32_t rhs) { val = to_mont(rhs); }
	B2_BB_HD constexpr Fp operator+(Fp rhs) const { return Fp(add_mod(val, rhs.val), Raw{}); }
	B2_BB_HD constexpr Fp operator-() const { return Fp(sub_mod(0, val), Raw{}); }
	B2_BB_HD constexpr Fp operator-(Fp rhs) const { return Fp(sub_mod(val, rhs.val), Raw{}); }
	B2_BB_HD constexpr Fp operator*(Fp rhs) const { return Fp(mont_mul(val, rhs.val), Raw{}); }
	B2_BB_HD constexpr Fp operator+=(Fp rhs) { return val = add_mod(val, rhs.val), *this; }
	B2_BB_HD constexpr Fp operator-=(Fp rhs) { return val = sub_mod(val, rhs.val), *this; }
	B2_BB_HD constexpr Fp operator*=(Fp rhs) { return val = mont_mul(val, rhs.val), *this; }
	B2_BB_HD constexpr bool operator==(Fp rhs) const { return val == rhs.val; }
	B2_BB_HD constexpr bool operator!=(Fp rhs) const { return val != rhs.val; }
	B2_BB_HD constexpr bool operator<(Fp rhs) const { return asUInt32() < rhs.asUInt32(); }
	B2_BB_HD constexpr bool operator<=(Fp rhs) const { return asUInt32() <= rhs.asUInt32(); }
	B2_BB_HD constexpr bool operator>(Fp rhs) const { return asUInt32() > rhs.asUInt32(); }
	B2_BB_HD constexpr bool operator>=(Fp rhs) const { return asUInt32() >= rhs.asUInt32(); }
	B2_BB_HD constexpr Fp operator++(int) {
		const Fp before = *this;
		val = add_mod(val, to_mont(1));
		return before;
	}
	B2_BB_HD constexpr Fp operator--(int) {
		const Fp before = *this;
		val = sub_mod(val, to_mont(1));
		return before;
	}
	B2_BB_HD constexpr Fp operator++() { return val = add_mod(val, to_mont(1)), *this; }
	B2_BB_HD constexpr Fp operator--() { return val = sub_mod(val, to_mont(1)), *this; }

	// square-and-multiply, least significant exponent bit first
	static B2_BB_HD constexpr Fp pow(Fp x, size_t n) {
		Fp acc = one();
		for (; n; n >>= 1, x *= x)
			if (n & 1) acc *= x;
		return acc;
	}
	// x^(P-2): Fermat; inv(0) = 0
	static B2_BB_HD constexpr Fp inv(Fp x) { return pow(x, P - 2); }

private:
	struct Raw {};
	uint32_t val;  // Montgomery residue, always < P
	B2_BB_HD constexpr Fp(uint32_t raw, Raw) : val(raw) {}

	static B2_BB_HD constexpr uint32_t add_mod(uint32_t a, uint32_t b) {
		const uint32_t s = a + b;
		return s >= P ? s - P : s;
	}
	static B2_BB_HD constexpr uint32_t sub_mod(uint32_t a, uint32_t b) { return a >= b ? a - b : a + P - b; }
	// a * b / 2^32 mod P: add the multiple of P that clears the low word (q = -t / P mod 2^32 = t * -M)
	static B2_BB_HD constexpr uint32_t mont_mul(uint32_t a, uint32_t b) {
		const uint64_t t = uint64_t(a) * b;
		const uint32_t q = uint32_t(t) * (0u - M);
		const uint32_t r = uint32_t((t + uint64_t(q) * P) >> 32);
		return r >= P ? r - P : r;
	}
	static B2_BB_HD constexpr uint32_t to_mont(uint32_t v) { return mont_mul(v, R2); }
};

}  // namespace risc0

using BB31 = risc0::Fp;
