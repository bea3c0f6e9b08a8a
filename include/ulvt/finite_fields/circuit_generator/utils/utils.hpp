// drop-in for finite_fields/circuit_generator/utils/utils.{hpp,cpp}:8-32 -- the two test helpers that
// finite_field_tests links (tests.cu:186-188): "0x..." hex text -> little-endian 32-bit words, and word -> hex text.
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>

// a[0] receives the LAST 8 hex digits of s, a[1] the 8 before them, ... up to the "0x" prefix
inline void write_string_to_int_arr(uint32_t* a, const std::string& s) {
	size_t end = s.size(), begin = 0;
	if (s.size() >= 2 && s[0] == '0' && (s[1] == 'x' || s[1] == 'X')) begin = 2;
	for (size_t k = 0; end > begin; k++) {
		const size_t from = end - begin > 8 ? end - 8 : begin;
		a[k] = (uint32_t)std::stoul(s.substr(from, end - from), nullptr, 16);
		end = from;
	}
}
inline std::string int_to_hex(unsigned int i) {
	char buf[16];
	snprintf(buf, sizeof buf, "%x", i);
	return buf;
}
