// ro_fft.cuh -- register-resident radix-2^k FFT building blocks (sm_100a).
//
// The reference computes X[k] = sum_n x[n] exp(-2*pi*i*k*n/N) with FFTW
// (reference src/FFTBackend.cpp:117-120,236).  Here the transform is a
// three-pass Cooley-Tukey decomposition N = R1*R2*R3 whose passes run
// entirely in registers (radix 4..32 per thread) and exchange data through
// shared memory; these helpers are the per-thread pieces.
#pragma once
#include <cuda_runtime.h>

namespace ro {

// cos(2*pi*j/32), sin(2*pi*j/32) for j = 0..16, rounded once from double.
__host__ __device__ constexpr float cw32(int j)
{
	switch (j) {
	case 0:  return 1.000000000e+00f;
	case 1:  return 9.807852804e-01f;
	case 2:  return 9.238795325e-01f;
	case 3:  return 8.314696123e-01f;
	case 4:  return 7.071067812e-01f;
	case 5:  return 5.555702330e-01f;
	case 6:  return 3.826834324e-01f;
	case 7:  return 1.950903220e-01f;
	case 8:  return 0.0f;
	case 9:  return -1.950903220e-01f;
	case 10: return -3.826834324e-01f;
	case 11: return -5.555702330e-01f;
	case 12: return -7.071067812e-01f;
	case 13: return -8.314696123e-01f;
	case 14: return -9.238795325e-01f;
	case 15: return -9.807852804e-01f;
	default: return -1.0f;
	}
}
__host__ __device__ constexpr float sw32(int j)
{
	// sin(t) = cos(pi/2 - t), and cos is even
	return cw32(j <= 8 ? 8 - j : j - 8);
}

// cos(2*pi*j/64), sin(2*pi*j/64) for j = 0..32
__host__ __device__ constexpr float cw64(int j)
{
	constexpr float T[17] = {1.000000000e+00f, 9.951847267e-01f, 9.807852804e-01f, 9.569403357e-01f, 9.238795325e-01f,
	                         8.819212643e-01f, 8.314696123e-01f, 7.730104534e-01f, 7.071067812e-01f, 6.343932842e-01f,
	                         5.555702330e-01f, 4.713967368e-01f, 3.826834324e-01f, 2.902846773e-01f, 1.950903220e-01f,
	                         9.801714033e-02f, 0.0f};
	return j <= 16 ? T[j] : -T[32 - j];
}
__host__ __device__ constexpr float sw64(int j) { return j <= 16 ? cw64(16 - j) : cw64(j - 16); }

__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }

__host__ __device__ constexpr int bitrev(int v, int bits)
{
	int r = 0;
	for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
	return r;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
	return make_float2(fmaf(a.x, b.x, -(a.y * b.y)), fmaf(a.x, b.y, a.y * b.x));
}

// In-place forward DIF FFT of R points held in registers.  Output k is left
// at v[bitrev(k)].  All indices are compile-time after unrolling, so v[] never
// leaves the register file.
template <int R>
__device__ __forceinline__ void fft_dif(float2 (&v)[R])
{
	static_assert(R >= 2 && R <= 32 && (R & (R - 1)) == 0, "radix");
	constexpr int L = ilog2(R);
#pragma unroll
	for (int lvl = 0; lvl < L; lvl++) {
		const int half = R >> (lvl + 1);
#pragma unroll
		for (int s = 0; s < R; s += 2 * half) {
#pragma unroll
			for (int j = 0; j < half; j++) {
				const float2 a = v[s + j];
				const float2 b = v[s + j + half];
				v[s + j] = make_float2(a.x + b.x, a.y + b.y);
				const float dx = a.x - b.x, dy = a.y - b.y;
				const int e = j * (16 / half);        // W_{2*half}^j = W_32^e
				if (e == 0) {
					v[s + j + half] = make_float2(dx, dy);
				} else if (e == 8) {                  // * (-i)
					v[s + j + half] = make_float2(dy, -dx);
				} else {
					const float wr = cw32(e), wi = -sw32(e);
					v[s + j + half] = make_float2(fmaf(dx, wr, -(dy * wi)),
					                              fmaf(dx, wi, dy * wr));
				}
			}
		}
	}
}

// Multiply output k (at v[bitrev(k)]) by w^k, k = 1..R-1, where
// base[i] = w^(2^i) were read from an accurately rounded table.  Powers are
// composed as w^(hi) * w^(lo) with at most ilog2(R)-1 chained products.
template <int R>
__device__ __forceinline__ void twiddle_powers(float2 (&v)[R], const float2 (&base)[ilog2(R)])
{
	constexpr int L = ilog2(R);
	constexpr int LB = (L + 1) / 2;       // low bits
	constexpr int NLO = 1 << LB;
	constexpr int NHI = 1 << (L - LB);
	float2 lo[NLO];
	lo[0] = make_float2(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < LB; msb++) {
		lo[1 << msb] = base[msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			lo[(1 << msb) + rest] = cmul(base[msb], lo[rest]);
	}
	float2 hi[NHI];
	hi[0] = make_float2(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < L - LB; msb++) {
		hi[1 << msb] = base[LB + msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			hi[(1 << msb) + rest] = cmul(base[LB + msb], hi[rest]);
	}
#pragma unroll
	for (int h = 0; h < NHI; h++) {
#pragma unroll
		for (int j = 0; j < NLO; j++) {
			const int k = h * NLO + j;
			if (k == 0) continue;
			const float2 w = h == 0 ? lo[j] : (j == 0 ? hi[h] : cmul(hi[h], lo[j]));
			v[bitrev(k, L)] = cmul(v[bitrev(k, L)], w);
		}
	}
}

__device__ __forceinline__ float fast_sqrt(float x)
{
	float r;
	asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
	return r;
}

}  // namespace ro
