// ro_v2.cu -- instantiations and launcher of wf_fft2_kernel (see ro_kernels2.cuh).
#include "ro_v2.h"
#include "ro_kernels2.cuh"

namespace ro {

namespace {

template <int LOG2N, int P, int R3, int MINB_STAGE, int MINB_DIRECT>
bool pick_one(int mode, bool gain, V2Kernel *k)
{
	k->threads = (1 << LOG2N) / P;
	k->log2n = LOG2N;
	k->mode = mode;
	k->gain = gain;
	if (mode == V2_STAGED) {
		k->smem = wf2_smem_bytes<LOG2N, P, R3, MODE_STAGED>();
		k->fn = gain ? (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_STAGED, true, MINB_STAGE>
		             : (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_STAGED, false, MINB_STAGE>;
	} else if (mode == V2_DIRECT) {
		k->smem = wf2_smem_bytes<LOG2N, P, R3, MODE_DIRECT>();
		k->fn = gain ? (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_DIRECT, true, MINB_DIRECT>
		             : (const void *)wf_fft2_kernel<LOG2N, P, R3, MODE_DIRECT, false, MINB_DIRECT>;
	} else {
		return false;
	}
	return true;
}

}  // namespace

bool v2_pick(int log2n, int mode, bool gain, V2Kernel *out)
{
	if (mode == V2_FUSED1) {
		if (log2n != 14) return false;
		out->log2n = log2n;
		out->mode = mode;
		out->gain = false;                       // the gain is a kernel parameter here
		out->threads = (1 << 14) / 32;
		out->smem = wf2_smem_bytes<14, 32, 16, MODE_FUSED1>();
		out->fn = (const void *)wf_fft2_kernel<14, 32, 16, MODE_FUSED1, false, 1>;
		return true;
	}
	if (mode == V2_FUSED2) {
		if (log2n != 14) return false;
		out->log2n = log2n;
		out->mode = mode;
		out->gain = false;
		out->threads = (1 << 14) / 32;
		out->smem = wf2_smem_bytes<14, 32, 16, MODE_FUSED2>();
		out->fn = (const void *)wf_fft2_kernel<14, 32, 16, MODE_FUSED2, false, 1>;
		return true;
	}
	if (mode == V2_SPLIT) {
		out->log2n = log2n;
		out->mode = mode;
		out->gain = false;                       // the gain is applied by the split kernel
		switch (log2n) {
#ifndef RO_EXP_ONLY14
		case 12:
			out->threads = (1 << 12) / 16;
			out->smem = wf2_smem_bytes<12, 16, 16, MODE_SPLIT>();
			out->fn = (const void *)wf_fft2_kernel<12, 16, 16, MODE_SPLIT, false, 3>;
			return true;
		case 13:
			out->threads = (1 << 13) / 32;
			out->smem = wf2_smem_bytes<13, 32, 8, MODE_SPLIT>();
			out->fn = (const void *)wf_fft2_kernel<13, 32, 8, MODE_SPLIT, false, 2>;
			return true;
#endif
		case 14:
			out->threads = (1 << 14) / 32;
			out->smem = wf2_smem_bytes<14, 32, 16, MODE_SPLIT>();
			out->fn = (const void *)wf_fft2_kernel<14, 32, 16, MODE_SPLIT, false, 1>;
			return true;
		default:
			return false;
		}
	}
	switch (log2n) {
#ifndef RO_EXP_ONLY14
#ifndef RO_MINB10
#define RO_MINB10 12
#define RO_MINB11 6
#define RO_MINB12 3
#endif
	case 10: return pick_one<10, 16, 4, RO_MINB10, RO_MINB10>(mode, gain, out);
	case 11: return pick_one<11, 16, 8, RO_MINB11, RO_MINB11>(mode, gain, out);
	case 12: return pick_one<12, 16, 16, RO_MINB12, RO_MINB12>(mode, gain, out);
	case 13: return pick_one<13, 32, 8, 2, 2>(mode, gain, out);
#endif
	case 14: return pick_one<14, 32, 16, 1, 1>(mode, gain, out);
	default: return false;
	}
}

int v2_tmem_cols(const V2Kernel &k)
{
	return wf2_tmem_cols_of(k.log2n, k.threads, k.mode == V2_SPLIT || k.mode == V2_FUSED1 || k.mode == V2_FUSED2 || k.mode == V2_FUSED1_F32);
}

cudaError_t v2_launch(const V2Kernel &k, int grid, cudaStream_t s, const void *d_iq, const float *d_window_half,
                      const float2 *d_tw, float *d_rows, long long row_stride, long long n_frames, long long hop,
                      long long frames_per_cta, float iq_gain, int log2na, int pair, int stagger)
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
	p.log2na = log2na;
	p.pair = pair;
	p.stagger = stagger;
	p.raw = NULL;
	p.window_full = NULL;
	p.tw_big = NULL;
	p.scratch = NULL;
	void *args[] = { &p };
	return cudaLaunchKernel(k.fn, dim3(grid), dim3(k.threads), args, k.smem, s);
}

cudaError_t v2_launch_fused(const V2Kernel &k, int grid, cudaStream_t s, const void *d_raw, const float *d_window_full,
                            const float2 *d_tw_big, const float2 *d_tw_nb, void *d_scratch, float *d_rows,
                            long long row_stride, long long n_frames, long long hop, long long frames_per_cta,
                            float iq_gain, int pair, int stagger, int log2na)
{
	const int la = log2na;
	Wf2Params p;
	p.iq = NULL;
	p.window_half = NULL;
	p.tw = d_tw_nb;
	p.rows = d_rows;
	p.n_frames = n_frames << la;                 // items
	p.hop = hop;
	p.row_stride = row_stride;
	p.frames_per_cta = frames_per_cta << la;     // a CTA owns whole frames
	p.iq_gain = iq_gain;
	p.log2na = la;
	p.pair = pair;
	p.stagger = stagger;
	p.raw = d_raw;
	p.window_full = d_window_full;
	p.tw_big = d_tw_big;
	p.scratch = reinterpret_cast<unsigned long long *>(d_scratch);
	void *args[] = { &p };
	return cudaLaunchKernel(k.fn, dim3(grid), dim3(k.threads), args, k.smem, s);
}

namespace {

template <int LOG2NA>
cudaError_t split_one(cudaStream_t s, dim3 grid, bool f32in, const SplitParams &p)
{
	if (f32in) wf_split_kernel<LOG2NA, true><<<grid, 256, 0, s>>>(p);
	else wf_split_kernel<LOG2NA, false><<<grid, 256, 0, s>>>(p);
	return cudaGetLastError();
}

}  // namespace

cudaError_t v2_launch_split(cudaStream_t s, int log2n, int log2na, bool f32in, const void *d_iq,
                            const float *d_window, const float2 *d_tw, void *d_y, long long frame0,
                            long long n_frames, long long hop, float iq_gain)
{
	SplitParams p;
	p.iq = d_iq;
	p.window = d_window;
	p.tw = d_tw;
	p.y = reinterpret_cast<c64 *>(d_y);
	p.frame0 = frame0;
	p.hop = hop;
	p.log2n = log2n;
	p.iq_gain = iq_gain;
	const dim3 grid((unsigned)((1 << (log2n - log2na)) / 256), (unsigned)n_frames);
	switch (log2na) {
	case 1: return split_one<1>(s, grid, f32in, p);
	case 2: return split_one<2>(s, grid, f32in, p);
	case 3: return split_one<3>(s, grid, f32in, p);
	case 4: return split_one<4>(s, grid, f32in, p);
	case 5: return split_one<5>(s, grid, f32in, p);
	default: return cudaErrorInvalidValue;
	}
}

}  // namespace ro

#if RO_TRACE
// tools/trace16k.py: the timeline the instrumented staged 16k kernel left behind (see ro_kernels2.cuh)
extern "C" int ro_trace_dump(unsigned long long *out, size_t n)
{
	const size_t have = sizeof(ro::ro_trace_buf) / sizeof(unsigned long long);
	if (n > have) n = have;
	return (int)cudaMemcpyFromSymbol(out, ro::ro_trace_buf, n * sizeof(unsigned long long));
}
#endif
