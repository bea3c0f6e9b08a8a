// b2_upper.cu -- launch of the upper passes: picks the instantiation of upper_kernel (b2_upper_kernel.cuh).  sm_100a only.
// Replaces additive_ntt_kernel<T,P> (ntt/additive_ntt.cuh:91-160) for the layers with stride >= 32 lines.
#include "b2_upper_kernel.cuh"

namespace b2 {

int launch_upper(const UpperParams& P, cudaStream_t stream) {
	if (P.k != 2 && P.k != 4 && P.k != 6) return -1;
	if (P.ls != (P.k == 6 ? 3 : 5) || (P.lpitch != kLeafRows5 && P.lpitch != kLeafRows4) || (P.k == 6 && P.lpitch != kLeafRows4)) return -1;
	// lean instantiations for passes (k >= 4) whose layers all have class 4 or all class 5
	int cls = P.k >= 4 ? P.cls[0] : 0;
	for (int i = 1; i < P.k && cls; i++)
		if (P.cls[i] != cls) cls = 0;
	if (cls != 4 && cls != 5) cls = 0;
	if (P.inverse) return cls == 5 ? launch_upper_i5(P, stream) : cls == 4 ? launch_upper_i4(P, stream) : launch_upper_i0(P, stream);
	return cls == 5 ? launch_upper_f5(P, stream) : cls == 4 ? launch_upper_f4(P, stream) : launch_upper_f0(P, stream);
}

}  // namespace b2
