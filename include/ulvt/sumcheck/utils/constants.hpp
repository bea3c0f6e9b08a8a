// ulvt/sumcheck/utils/constants.hpp -- the names user code takes from the reference's sumcheck/utils/constants.hpp,
// as typed constants: the field is GF(2^(2^7)), interpolation nodes live in GF(2^(2^2)), a bit-sliced batch is 128 words
// and a compact value 4 words.  The reference's launch-shape macros have no meaning here and are not provided.
#pragma once
#include <cstddef>
#include <cstdint>

constexpr int TOWER_HEIGHT = 7, INTERPOLATION_TOWER_HEIGHT = 2;
constexpr std::size_t BITS_WIDTH = std::size_t(1) << TOWER_HEIGHT;
constexpr std::size_t INTERPOLATION_BITS_WIDTH = std::size_t(1) << INTERPOLATION_TOWER_HEIGHT;
constexpr std::size_t INTS_PER_VALUE = BITS_WIDTH / 32;
