// b2_compat.cu -- the compiled symbols of the reference's `ulvt_gpu` / `unrolled` static libraries that user code links by
// name (CMakeLists.txt:19-30, 80-89), on top of this library:
//   bool check_gpu_capabilities()                        utils/common.cu:6-43
//   dim3 KERNEL_LAUNCH_CONFIGS[31][3][2]                  ntt/nttconf.cu:1-190 (data symbol; not consulted by this library)
//   __global__ bitsliced_repeat(uint32_t*, uint32_t*, uint32_t*)   finite_fields/tests/profiling/kernels/bitsliced_repeat.cu:6-33
//   __global__ babybear_repeat(BB31*, BB31*)              finite_fields/tests/profiling/kernels/babybear_repeat.cu:3-14
// (multiply_unrolled<H>, write_string_to_int_arr and int_to_hex are header-inline in include/ulvt.)
// Built into libb2ntt.so and libulvt_gpu.a.
#include <cuda_runtime.h>

#include "../../include/ulvt/finite_fields/tests/profiling/kernels/babybear_repeat.cuh"
#include "../../include/ulvt/finite_fields/tests/profiling/kernels/bitsliced_repeat.cuh"
#include "../../include/ulvt/ntt/nttconf.cuh"
#include "b2_field.cuh"

#pragma GCC visibility push(default)

bool check_gpu_capabilities() { return b2ntt_check_gpu_capabilities(0) > 0; }

// One butterfly per thread, at most 11 stages per kernel, 1024-thread blocks: kernel 0 walks contiguous tiles with a 1-D
// grid, kernels 1 and 2 put the strided tiles on grid.y / grid.z (the shape convention of nttconf.cu, values computed here).
dim3 KERNEL_LAUNCH_CONFIGS[31][3][2];
namespace {
struct LaunchTableInit {
	LaunchTableInit() {
		for (int lg = 0; lg < 31; lg++)
			for (int kern = 0; kern < 3; kern++) {
				dim3 block(1, 1, 1), grid(1, 1, 1);
				const int first = kern * MAX_STAGES_PER_KERNEL;
				if (lg >= 1 && first < lg) {
					const long long butterflies = 1LL << (lg - 1);
					const int stages = lg - first < MAX_STAGES_PER_KERNEL ? lg - first : MAX_STAGES_PER_KERNEL;
					const int bx = 1 << (stages - 1 < 10 ? stages - 1 : 10);  // threads along the butterfly bits of this kernel
					const int by = kern == 0 ? 1 : 1024 / bx;                 // neighbouring tiles share a block when the tile is small
					block = dim3(bx, by, 1);
					const long long blocks = butterflies / ((long long)bx * by) > 0 ? butterflies / ((long long)bx * by) : 1;
					if (kern == 0)
						grid = dim3((unsigned)blocks, 1, 1);
					else if (kern == 1)
						grid = dim3(1, (unsigned)(blocks < 2048 ? blocks : 2048), (unsigned)(blocks < 2048 ? 1 : blocks / 2048));
					else
						grid = dim3(1, (unsigned)(blocks < 16 ? blocks : 16), (unsigned)(blocks < 16 ? 1 : blocks / 16));
				}
				KERNEL_LAUNCH_CONFIGS[lg][kern][0] = block;
				KERNEL_LAUNCH_CONFIGS[lg][kern][1] = grid;
			}
	}
} launch_table_init;
}  // namespace

// 50 000 dependent bit-sliced GF(2^32) products per thread, x <- x * y: thread t owns words [32 t, 32 t + 32) of x, y and dst
__global__ void bitsliced_repeat(uint32_t* x, uint32_t* y, uint32_t* dst) {
	const uint32_t tid = threadIdx.x + blockIdx.x * blockDim.x;
	uint32_t a[32], b[32];
#pragma unroll
	for (int i = 0; i < 32; i++) a[i] = x[i + 32 * tid], b[i] = y[i + 32 * tid];
#pragma unroll 1
	for (int it = 0; it < 50000; it++) b2::Sliced<5>::mul(a, b, a);
#pragma unroll
	for (int i = 0; i < 32; i++) dst[i + 32 * tid] = a[i];
}

// 32 * 50 000 dependent BabyBear products per thread, a <- b * a with b the thread's operand
__global__ void babybear_repeat(BB31* operands, BB31* results) {
	const uint32_t tid = threadIdx.x + blockIdx.x * blockDim.x;
	const BB31 b = operands[tid];
	BB31 a = b;
#pragma unroll 1
	for (int i = 0; i < 32 * 50000; i++) a = b * a;
	results[tid] = a;
}

#pragma GCC visibility pop
