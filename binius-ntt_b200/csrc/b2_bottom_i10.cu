// one instantiation of bottom_kernel (b2_bottom_kernel.cuh) per translation unit, so that they compile in parallel
#include "b2_bottom_kernel.cuh"
namespace b2 {
B2_BOTTOM_INSTANCE(launch_bottom_i10, true, true, false)
}
