// ro_fft64.cuh -- the same waterfall row in DOUBLE precision (cfg.precision = 1).
//
// The reference transforms in double with FFTW (src/FFTBackend.cpp:117-120,236) and takes the
// magnitude in double before the one rounding to float (src/WaterfallBackend.cpp:492-505).
// BASELINE.json's north_star allows "relative 1e-5 in linear power, or 1e-3 dB"; the packed-FP32
// production kernel meets the second (tests/helpers.py), this variant meets the first on every
// bin: it is the arithmetic of the reference, stage for stage, on the GPU.  It is the accuracy
// reference of the product, not its fast path: data stay in a global scratch (L2-resident for the
// batch sizes used) and only the first 11 radix-2 stages run in shared memory.
//
//   wf64_load_kernel    framing, IQ gain, window (src/FFTBackend.cpp:211-232) -> scratch[f][bitrev(n)]
//   wf64_block_kernel   DIT stages 1..min(11, log2N) on 2048-point blocks in shared memory
//   wf64_stage_kernel   one later DIT stage over the whole batch, in place
//   wf64_mag_kernel     sqrt(re*re + im*im) in double, fftshift, float store
#pragma once
#include <cuda_runtime.h>

namespace ro {

struct Wf64Params {
	const void    *iq;        // int16 or float32 interleaved I/Q
	const float   *window;    // N floats
	const double2 *tw;        // N/2 entries exp(-2*pi*i*j/N)
	double2       *x;         // scratch [n_frames][N]
	float         *rows;      // [n_frames][row_stride]
	long long      n_frames, hop, row_stride;
	double         iq_gain;
	int            log2n;
	int            stage;     // wf64_stage_kernel: butterflies span 2^stage points
};

template <bool F32IN>
__global__ void __launch_bounds__(256)
wf64_load_kernel(const Wf64Params p)
{
	const int N = 1 << p.log2n;
	const long long total = p.n_frames << p.log2n;
	for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
	     idx += (long long)gridDim.x * blockDim.x) {
		const long long f = idx >> p.log2n;
		const int n = (int)(idx & (N - 1));
		double re, im;
		if (F32IN) {
			const float2 v = __ldg(reinterpret_cast<const float2 *>(p.iq) + f * p.hop + n);
			re = (double)v.x; im = (double)v.y;
		} else {
			const short2 v = __ldg(reinterpret_cast<const short2 *>(p.iq) + f * p.hop + n);
			re = (double)v.x; im = (double)v.y;
		}
		im = __dadd_rn(im, p.iq_gain);                         // additive on Q (src/FFTBackend.cpp:71-79)
		const double w = (double)__ldg(p.window + n);
		const int r = (int)(__brev((unsigned int)n) >> (32 - p.log2n));
		p.x[(f << p.log2n) + r] = make_double2(__dmul_rn(re, w), __dmul_rn(im, w));
	}
}

// z = a + w*b, a - w*b without contraction surprises: products and sums rounded separately
__device__ __forceinline__ void bfly64(double2 &a, double2 &b, const double2 w)
{
	const double tr = __dsub_rn(__dmul_rn(b.x, w.x), __dmul_rn(b.y, w.y));
	const double ti = __dadd_rn(__dmul_rn(b.x, w.y), __dmul_rn(b.y, w.x));
	const double2 a0 = a;
	a = make_double2(__dadd_rn(a0.x, tr), __dadd_rn(a0.y, ti));
	b = make_double2(__dsub_rn(a0.x, tr), __dsub_rn(a0.y, ti));
}

constexpr int WF64_LOG2B = 11;          // 2048 points * 16 B = 32 KB of shared memory per CTA

__global__ void __launch_bounds__(256)
wf64_block_kernel(const Wf64Params p)
{
	__shared__ double2 s[1 << WF64_LOG2B];
	const int lb = p.log2n < WF64_LOG2B ? p.log2n : WF64_LOG2B;
	const int B = 1 << lb;
	const long long blocks = (p.n_frames << p.log2n) >> lb;
	for (long long blk = blockIdx.x; blk < blocks; blk += gridDim.x) {
		double2 *g = p.x + (blk << lb);
		for (int i = threadIdx.x; i < B; i += blockDim.x) s[i] = g[i];
		__syncthreads();
		for (int st = 1; st <= lb; st++) {
			const int half = 1 << (st - 1);
			for (int i = threadIdx.x; i < B / 2; i += blockDim.x) {
				const int j = i & (half - 1);
				const int a = ((i >> (st - 1)) << st) + j;
				const double2 w = __ldg(p.tw + ((long long)j << (p.log2n - st)));   // W_{2^st}^j
				bfly64(s[a], s[a + half], w);
			}
			__syncthreads();
		}
		for (int i = threadIdx.x; i < B; i += blockDim.x) g[i] = s[i];
		__syncthreads();
	}
}

__global__ void __launch_bounds__(256)
wf64_stage_kernel(const Wf64Params p)
{
	const int st = p.stage;
	const int half = 1 << (st - 1);
	const long long total = (p.n_frames << p.log2n) >> 1;
	for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
	     i += (long long)gridDim.x * blockDim.x) {
		const int j = (int)(i & (half - 1));
		const long long a = ((i >> (st - 1)) << st) + j;       // frames are 2^log2n aligned: no crossing
		const double2 w = __ldg(p.tw + ((long long)j << (p.log2n - st)));
		double2 u = p.x[a], v = p.x[a + half];
		bfly64(u, v, w);
		p.x[a] = u;
		p.x[a + half] = v;
	}
}

__global__ void __launch_bounds__(256)
wf64_mag_kernel(const Wf64Params p)
{
	const int N = 1 << p.log2n;
	const long long total = p.n_frames << p.log2n;
	for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
	     idx += (long long)gridDim.x * blockDim.x) {
		const long long f = idx >> p.log2n;
		const int j = (int)(idx & (N - 1));                    // row index; FFT bin (j + N/2) mod N
		const double2 z = p.x[(f << p.log2n) + ((j + N / 2) & (N - 1))];
		const double m = sqrt(__dadd_rn(__dmul_rn(z.x, z.x), __dmul_rn(z.y, z.y)));
		p.rows[f * p.row_stride + j] = (float)m;
	}
}

}  // namespace ro
