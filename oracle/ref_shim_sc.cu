/* ref_shim_sc.cu -- C entry points onto the reference's host-callable sumcheck batch helpers
 * (sumcheck/core/core.cu:9-96, compiled from /root/reference where it lies) and BitsliceUtils<128>::
 * repeat_value_bitsliced (utils/bitslicing.cuh:66-74), so that tests can replay Sumcheck's data flow on the CPU
 * and compare it with the compact restatement (oracle/sumcheck.c).  TEST INFRASTRUCTURE. */
#include <cstdint>

#include "ulvt/sumcheck/core/core.cuh"
#include "ulvt/utils/bitslicing.cuh"

extern "C" void ref_sc_compose(const uint32_t* first_batch_of_row, uint32_t* dst, uint32_t composition_size, uint32_t evals_per_col) {
	evaluate_composition_on_batch_row(first_batch_of_row, dst, composition_size, evals_per_col);
}
extern "C" void ref_sc_fold_batch(const uint32_t* lower, const uint32_t* upper, uint32_t* dst, const uint32_t* coefficient, int is_interpolation) {
	fold_batch(lower, upper, dst, coefficient, is_interpolation != 0);
}
extern "C" void ref_sc_fold_small(const uint32_t* source, uint32_t* destination, const uint32_t* coefficient, uint32_t list_len) {
	fold_small(source, destination, coefficient, list_len);
}
extern "C" void ref_sc_compute_sum(uint32_t* sum, uint32_t* bitsliced_batch, uint32_t n_unpadded) { compute_sum(sum, bitsliced_batch, n_unpadded); }
extern "C" void ref_sc_repeat_value(uint32_t* batch, const uint32_t* value) { BitsliceUtils<128>::repeat_value_bitsliced(batch, value); }
