// ulvt/sumcheck/utils/constants.hpp -- the reference's sumcheck constants (sumcheck/utils/constants.hpp:1-10) that user
// code names; the launch-shape macros (BLOCKS, THREADS_PER_BLOCK) are not consulted by this implementation.
#pragma once
#include <cstddef>
#define TOWER_HEIGHT 7
#define INTERPOLATION_TOWER_HEIGHT 2
#define BITS_WIDTH (size_t(1 << TOWER_HEIGHT))
#define INTERPOLATION_BITS_WIDTH (size_t(1 << INTERPOLATION_TOWER_HEIGHT))
#define INTS_PER_VALUE (BITS_WIDTH / 32)
