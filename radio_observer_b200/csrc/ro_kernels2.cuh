// ro_kernels2.cuh -- v2 of the fused waterfall kernel: packed f32x2 arithmetic,
// TMA-staged raw samples kept as a ring in shared memory, window table in
// shared memory, compile-time N.
//
// Same stages and the same index mapping as wf_fft_kernel (ro_kernels.cuh):
//   FFTBackend::process          src/FFTBackend.cpp:211-247   framing, IQ gain, window
//   fftw_execute                 src/FFTBackend.cpp:236       N-point forward DFT
//   WaterfallBackend::processFFT src/WaterfallBackend.cpp:492-505  magnitude, fftshift
//
// What changed against v1 and why (profiles/r01_ncu_full_fft16k_v1.txt):
//  * a CTA owns a CONTIGUOUS run of frames; the raw int16 samples of the current
//    frame live in a shared-memory ring of N samples, and only the `hop` new
//    samples of the next frame are fetched, by one elected thread, with
//    cp.async.bulk (TMA, completion on an mbarrier) while the current frame is
//    transformed.  v1 re-read the whole frame from L2 with 32 scalar LDG per thread
//    and stalled on them (lg_throttle, long_scoreboard);
//  * the Nuttall table is symmetric (verified on the host at create time), so its
//    first half sits in shared memory: no per-frame table traffic at all;
//  * int16 -> float goes through PRMT + FADD2 instead of 64 cvt per thread on the
//    16-lane XU pipe;
//  * all butterflies and twiddles use f32x2 (ro_fft2.cuh): half the issue slots;
//  * N is a template parameter, so store addresses are immediates.
#pragma once
#include <stdint.h>
#include "ro_fft2.cuh"

namespace ro {

struct Wf2Params {
	const void   *iq;           // device input, interleaved int16 I/Q; 16-byte aligned when STAGE
	const float  *window_half;  // w[0 .. N/2)   (w[N-1-i] == w[i])
	const float2 *tw;           // N entries exp(-2*pi*i*j/N)
	float        *rows;         // [n_frames][row_stride]
	long long     n_frames;
	long long     hop;
	long long     row_stride;
	long long     frames_per_cta;
	float         iq_gain;
};

__device__ __forceinline__ unsigned int smem_u32(const void *p)
{
	return (unsigned int)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned int count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned int bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned int parity)
{
	asm volatile(
		"{\n\t.reg .pred p;\n\t"
		"WAIT_%=:\n\t"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
		"@p bra DONE_%=;\n\t"
		"bra WAIT_%=;\n\t"
		"DONE_%=:\n\t}"
		:: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA), bytes a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned int bytes, unsigned long long *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
	             :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// bytes of dynamic shared memory the kernel needs
template <int LOG2N, int P, int R3, bool STAGE>
constexpr size_t wf2_smem_bytes()
{
	constexpr size_t N = (size_t)1 << LOG2N;
	constexpr size_t PAD1 = (R3 < 16) ? R3 : 0;
	return sizeof(c64) * (N + P * PAD1) + sizeof(float) * (N / 2) + 32 + (STAGE ? 4 * N : 0);
}

template <int LOG2N, int P, int R3, bool STAGE, bool GAIN, int MINB>
__global__ void __launch_bounds__((1 << LOG2N) / P, MINB)
wf_fft2_kernel(const Wf2Params p)
{
	constexpr int N = 1 << LOG2N;
	constexpr int R1 = P, R2 = P;
	constexpr int T = N / P;
	constexpr int M = N / R1;                  // = R2*R3
	static_assert(R1 * R2 * R3 == N, "factorisation");
	static_assert(R3 <= 16 && R3 >= 2 && P % R3 == 0, "R3");
	constexpr int PAD1 = (R3 < 16) ? R3 : 0;
	constexpr int ROW1 = M + PAD1;
	constexpr int LP = ilog2(P);
	constexpr int L3 = ilog2(R3);
	constexpr int SW = 16 / R3;
	constexpr int NMASK = N - 1;
	constexpr unsigned int CHUNK = 16384;      // bytes per bulk copy

	extern __shared__ __align__(128) unsigned char smem_raw[];
	c64 *S = reinterpret_cast<c64 *>(smem_raw);
	float *win = reinterpret_cast<float *>(S + N + P * PAD1);
	// [0] TMA completion, [1] "pass-2 inputs read", [2] "pass-3 inputs read" (split barriers)
	unsigned long long *mbar = reinterpret_cast<unsigned long long *>(win + N / 2);
	unsigned long long *bar_rd2 = mbar + 1, *bar_rd3 = mbar + 2;
	unsigned int *ring = reinterpret_cast<unsigned int *>(mbar + 4);

	const int t = threadIdx.x;
	const long long f0 = (long long)blockIdx.x * p.frames_per_cta;
	long long f1 = f0 + p.frames_per_cta;
	if (f1 > p.n_frames) f1 = p.n_frames;
	if (f0 >= f1) return;

	const unsigned char *iq8 = reinterpret_cast<const unsigned char *>(p.iq);
	const long long hop = p.hop;

	if (t == 0) {
		mbar_init(mbar, 1);
		mbar_init(bar_rd2, T);
		mbar_init(bar_rd3, T);
	}
	for (int i = t; i < N / 2; i += T) win[i] = __ldg(p.window_half + i);
	__syncthreads();
	if (STAGE && t == 0) {
		mbar_expect_tx(mbar, 4u * N);
		const unsigned char *src = iq8 + 4 * f0 * hop;
#pragma unroll 1
		for (unsigned int o = 0; o < 4u * N; o += CHUNK)
			bulk_g2s(reinterpret_cast<unsigned char *>(ring) + o, src + o, CHUNK < 4u * N - o ? CHUNK : 4u * N - o, mbar);
	}
	unsigned int parity = 0;

	// per-thread twiddle bases, fixed for the life of the CTA
	c64 b1[LP], b2[LP];
	{
		const unsigned long long *tw64 = reinterpret_cast<const unsigned long long *>(p.tw);
		const int n3 = t % R3;
#pragma unroll
		for (int i = 0; i < LP; i++) {
			b1[i] = __ldg(tw64 + ((t << i) & NMASK));
			b2[i] = __ldg(tw64 + (((n3 * P) << i) & NMASK));      // W_{R2*R3} = W_N^{N/M} = W_N^P
		}
	}
	const c64 gain2 = pk(0.f, p.iq_gain);

#pragma unroll 1
	for (long long f = f0; f < f1; f++) {
		c64 v[P];
		// ---- load (shared-memory ring or global) + window ----
		if (STAGE) {
			const int base = (int)(((f - f0) * hop) & NMASK);
			mbar_wait(mbar, parity);
			parity ^= 1u;
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				const int s = n1 * M + t;
				const c64 x = short2_to_c64(ring[(base + s) & NMASK]);
				const float w = (n1 < R1 / 2) ? win[n1 * M + t] : win[(R1 - 1 - n1) * M + (M - 1 - t)];
				v[n1] = mul2(GAIN ? add2(x, gain2) : x, pk(w, w));
			}
		} else {
			const unsigned int *src = reinterpret_cast<const unsigned int *>(p.iq) + f * hop;
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				const int s = n1 * M + t;
				const c64 x = short2_to_c64(__ldg(src + s));
				const float w = (n1 < R1 / 2) ? win[n1 * M + t] : win[(R1 - 1 - n1) * M + (M - 1 - t)];
				v[n1] = mul2(GAIN ? add2(x, gain2) : x, pk(w, w));
			}
		}

		// ---- pass 1 ----
		fft_dif2<R1>(v);
		twiddle_powers2<R1>(v, b1);
		// S still holds the previous frame's pass-3 inputs until every thread has read them
		if (f > f0) mbar_wait(bar_rd3, (unsigned int)(f - 1 - f0) & 1u);
#pragma unroll
		for (int k1 = 0; k1 < R1; k1++) S[k1 * ROW1 + t] = v[bitrev(k1, LP)];
		__syncthreads();

		// every thread has consumed the ring for this frame: fetch the next frame's new samples
		if (STAGE && t == 0 && f + 1 < f1) {
			const long long q0 = (f - f0) * hop + N;          // first new sample, relative to the CTA's first
			const unsigned int bytes = 4u * (unsigned int)hop;  // hop <= N
			const unsigned int pos = 4u * (unsigned int)(q0 & NMASK);
			const unsigned char *src = iq8 + 4 * (f0 * hop + q0);
			unsigned char *dst = reinterpret_cast<unsigned char *>(ring);
			mbar_expect_tx(mbar, bytes);
			unsigned int done = 0;
#pragma unroll 1
			while (done < bytes) {
				const unsigned int at = (pos + done) & (4u * N - 1u);
				unsigned int n = bytes - done;
				if (n > CHUNK) n = CHUNK;
				if (n > 4u * N - at) n = 4u * N - at;         // wrap of the ring
				bulk_g2s(dst + at, src + done, n, mbar);
				done += n;
			}
		}

		// ---- pass 2 ----
		{
			const int k1 = t / R3, n3 = t % R3;
#pragma unroll
			for (int n2 = 0; n2 < R2; n2++) v[n2] = S[k1 * ROW1 + n2 * R3 + n3];
			mbar_arrive(bar_rd2);                       // my pass-2 inputs are in registers
			fft_dif2<R2>(v);
			twiddle_powers2<R2>(v, b2);
			mbar_wait(bar_rd2, (unsigned int)(f - f0) & 1u);   // ... and so are everyone else's
			const int col = k1 ^ (n3 * SW);
#pragma unroll
			for (int k2 = 0; k2 < R2; k2++) S[(k2 * R3 + n3) * R1 + col] = v[bitrev(k2, LP)];
		}
		__syncthreads();

		// ---- pass 3 + magnitude + fftshift store ----
		constexpr int G = P / R3;                  // 16-point (R3) transforms per thread
		if (G >= 2) {
			// thread owns G ADJACENT outputs pidx = G*t + c: their pass-3 inputs are G adjacent
			// (XOR-permuted) slots -> 128-bit shared loads; their magnitudes G adjacent floats
			// -> one 64/128-bit global store per k3
			const int pbase = G * t;
			const int k1b = pbase % R1, k2 = pbase / R1;
			c64 (&u)[P] = v;                       // u[c*R3 + n3]
#pragma unroll
			for (int n3 = 0; n3 < R3; n3++) {
				constexpr int GM = G - 1;
				const int sw = n3 * SW;
				const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(S + (k2 * R3 + n3) * R1 + (k1b ^ (sw & ~GM)));
#pragma unroll
				for (int h2 = 0; h2 < G / 2; h2++) {
					const ulonglong2 q = src[h2];
					u[((2 * h2) ^ (sw & GM)) * R3 + n3] = q.x;
					u[((2 * h2 + 1) ^ (sw & GM)) * R3 + n3] = q.y;
				}
			}
			mbar_arrive(bar_rd3);                      // my pass-3 inputs are in registers
			float mag[R3][G];
#pragma unroll
			for (int c = 0; c < G; c++) {
				c64 w[R3];
#pragma unroll
				for (int n3 = 0; n3 < R3; n3++) w[n3] = u[c * R3 + n3];
				fft_dif2<R3>(w);
#pragma unroll
				for (int k3 = 0; k3 < R3; k3++) {
					float zr, zi;
					upk(w[bitrev(k3, L3)], zr, zi);
					mag[k3][c] = fast_sqrt(fmaf(zr, zr, zi * zi));
				}
			}
			float *row = p.rows + f * p.row_stride + G * t;
#pragma unroll
			for (int k3 = 0; k3 < R3; k3++) {
				// bins G*t + c + k3*(N/R3) go to row index (bin + N/2) mod N; G*t + c < N/R3, no wrap inside
				const int off = (k3 * (N / R3) + N / 2) & NMASK;
				if (G == 2) *reinterpret_cast<float2 *>(row + off) = make_float2(mag[k3][0], mag[k3][1]);
				else *reinterpret_cast<float4 *>(row + off) = make_float4(mag[k3][0], mag[k3][1], mag[k3][2 % G], mag[k3][3 % G]);
			}
		} else {
			float *row = p.rows + f * p.row_stride + t;
			c64 u[R3];
#pragma unroll
			for (int n3 = 0; n3 < R3; n3++) u[n3] = S[((t / R1) * R3 + n3) * R1 + ((t % R1) ^ (n3 * SW))];
			mbar_arrive(bar_rd3);
			fft_dif2<R3>(u);
#pragma unroll
			for (int k3 = 0; k3 < R3; k3++) {
				const int off = (k3 * (N / R3) + N / 2) & NMASK;   // t < N/R3 = T
				float zr, zi;
				upk(u[bitrev(k3, L3)], zr, zi);
				row[off] = fast_sqrt(fmaf(zr, zr, zi * zi));
			}
		}
	}
}

}  // namespace ro
