// ro_v2f.cu -- float32-input instantiations of wf_fft2_kernel (MODE_DIRECT_F32), a translation unit of
// their own so that they compile in parallel with ro_v2.cu.
#include "ro_v2.h"
#include "ro_kernels2.cuh"

namespace ro {

namespace {

template <int LOG2N, int P, int R3, int MINB>
bool pick_f32(bool gain, V2Kernel *k)
{
	k->threads = (1 << LOG2N) / P;
	k->log2n = LOG2N;
	k->mode = V2_DIRECT_F32;
	k->gain = gain;
	k->smem = wf2_smem_bytes<LOG2N, P, R3, MODE_DIRECT_F32>();
	k->fn = gain ? (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_DIRECT_F32, true, MINB>
	             : (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_DIRECT_F32, false, MINB>;
	return true;
}

}  // namespace

bool v2_pick_f32(int log2n, bool gain, V2Kernel *out)
{
	switch (log2n) {
	case 10: return pick_f32<10, 16, 4, 12>(gain, out);
	case 11: return pick_f32<11, 16, 8, 6>(gain, out);
	case 12: return pick_f32<12, 16, 16, 3>(gain, out);
	case 13: return pick_f32<13, 32, 8, 2>(gain, out);
	case 14: return pick_f32<14, 32, 16, 1>(gain, out);
	default: return false;
	}
}

bool v2_pick_fused_f32(V2Kernel *out)
{
	out->log2n = 14;
	out->mode = V2_FUSED1_F32;
	out->gain = false;                           // the gain is a kernel parameter here
	out->threads = (1 << 14) / 32;
	out->smem = wf2_smem_bytes<14, 32, 16, MODE_FUSED1_F32>();
	out->fn = (const void *)wf_fft2_kernel<14, 32, 16, MODE_FUSED1_F32, false, 1>;
	return true;
}

}  // namespace ro
