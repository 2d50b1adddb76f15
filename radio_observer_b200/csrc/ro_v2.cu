// ro_v2.cu -- instantiations and launcher of wf_fft2_kernel (see ro_kernels2.cuh).
#include "ro_v2.h"
#include "ro_kernels2.cuh"

namespace ro {

namespace {

template <int LOG2N, int P, int R3, int MINB_STAGE, int MINB_DIRECT>
bool pick_one(bool stage, bool gain, V2Kernel *k)
{
	k->threads = (1 << LOG2N) / P;
	k->log2n = LOG2N;
	k->stage = stage;
	k->gain = gain;
	if (stage) {
		k->smem = wf2_smem_bytes<LOG2N, P, R3, true>();
		k->fn = gain ? (const void *)wf_fft2_kernel<LOG2N, P, R3, true, true, MINB_STAGE>
		             : (const void *)wf_fft2_kernel<LOG2N, P, R3, true, false, MINB_STAGE>;
	} else {
		k->smem = wf2_smem_bytes<LOG2N, P, R3, false>();
		k->fn = gain ? (const void *)wf_fft2_kernel<LOG2N, P, R3, false, true, MINB_DIRECT>
		             : (const void *)wf_fft2_kernel<LOG2N, P, R3, false, false, MINB_DIRECT>;
	}
	return true;
}

}  // namespace

bool v2_pick(int log2n, bool stage, bool gain, V2Kernel *out)
{
	switch (log2n) {
	case 12: return pick_one<12, 16, 16, 3, 3>(stage, gain, out);
	case 13: return pick_one<13, 32, 8, 2, 2>(stage, gain, out);
	case 14: return pick_one<14, 32, 16, 1, 1>(stage, gain, out);
	default: return false;
	}
}

cudaError_t v2_launch(const V2Kernel &k, int grid, cudaStream_t s, const void *d_iq, const float *d_window_half,
                      const float2 *d_tw, float *d_rows, long long row_stride, long long n_frames, long long hop,
                      long long frames_per_cta, float iq_gain)
{
	Wf2Params p;
	p.iq = d_iq;
	p.window_half = d_window_half;
	p.tw = d_tw;
	p.rows = d_rows;
	p.n_frames = n_frames;
	p.hop = hop;
	p.row_stride = row_stride;
	p.frames_per_cta = frames_per_cta;
	p.iq_gain = iq_gain;
	void *args[] = { &p };
	return cudaLaunchKernel(k.fn, dim3(grid), dim3(k.threads), args, k.smem, s);
}

}  // namespace ro
