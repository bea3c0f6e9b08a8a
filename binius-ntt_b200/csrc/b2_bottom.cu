// b2_bottom.cu -- launch of the lowest pass: picks the instantiation of bottom_kernel (b2_bottom_kernel.cuh).  sm_100a only.
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the stages below 512 lines.
#include "b2_bottom_kernel.cuh"

namespace b2 {

int launch_bottom(const BottomParams& B, cudaStream_t stream) {
	// the exchanged in-line form needs the compact image on the in-line side of the pass (its element permutation is undone by the
	// chunk-wise loads / stores of the compact words)
	const bool swap = B.swap_inline && (B.up.inverse ? !B.up.src_sliced : !B.up.dst_sliced);
	bool all5 = swap;  // the lean instantiation exists for the exchanged form only
	for (int st = 0; st < B.n_inline + 9 - B.batch_bits && all5; st++) all5 = B.cls[st] == 5;
	if (B.up.inverse) return all5 ? launch_bottom_i11(B, stream) : swap ? launch_bottom_i10(B, stream) : launch_bottom_i00(B, stream);
	return all5 ? launch_bottom_f11(B, stream) : swap ? launch_bottom_f10(B, stream) : launch_bottom_f00(B, stream);
}

}  // namespace b2
