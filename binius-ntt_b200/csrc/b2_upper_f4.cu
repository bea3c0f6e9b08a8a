// one instantiation of upper_kernel (b2_upper_kernel.cuh) per translation unit, so that they compile in parallel
#include "b2_upper_kernel.cuh"
namespace b2 {
B2_UPPER_INSTANCE(launch_upper_f4, false, 4)
}
