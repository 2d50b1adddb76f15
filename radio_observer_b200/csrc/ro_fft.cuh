// ro_fft.cuh -- register-resident radix-2^k FFT building blocks (sm_100a).
//
// The reference computes X[k] = sum_n x[n] exp(-2*pi*i*k*n/N) with FFTW
// (reference src/FFTBackend.cpp:117-120,236).  Here the transform is a
// three-pass Cooley-Tukey decomposition N = R1*R2*R3 whose passes run
// entirely in registers (radix 4..32 per thread) and exchange data through
// shared memory; the constants and index helpers of the per-thread pieces (ro_fft2.cuh).
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

// cos(2*pi*j/128), sin(2*pi*j/128) for j = 0..127 (table for the first quadrant, rounded once)
__host__ __device__ constexpr float cw128(int j)
{
	constexpr float T[33] = {1.000000000e+00f, 9.987954497e-01f, 9.951847196e-01f, 9.891765118e-01f, 9.807852507e-01f,
	                         9.700312614e-01f, 9.569403529e-01f, 9.415440559e-01f, 9.238795042e-01f, 9.039893150e-01f,
	                         8.819212914e-01f, 8.577286005e-01f, 8.314695954e-01f, 8.032075167e-01f, 7.730104327e-01f,
	                         7.409511209e-01f, 7.071067691e-01f, 6.715589762e-01f, 6.343932748e-01f, 5.956993103e-01f,
	                         5.555702448e-01f, 5.141027570e-01f, 4.713967443e-01f, 4.275550842e-01f, 3.826834261e-01f,
	                         3.368898630e-01f, 2.902846634e-01f, 2.429801822e-01f, 1.950903237e-01f, 1.467304677e-01f,
	                         9.801714122e-02f, 4.906767607e-02f, 0.0f};
	const int k = j & 127;
	return k <= 32 ? T[k] : k <= 64 ? -T[64 - k] : k <= 96 ? -T[k - 64] : T[128 - k];
}
__host__ __device__ constexpr float sw128(int j) { return cw128(j + 96); }      // sin(x) = cos(x - pi/2) = cos(x + 3*pi/2)

__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }

__host__ __device__ constexpr int bitrev(int v, int bits)
{
	int r = 0;
	for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
	return r;
}

__device__ __forceinline__ float fast_sqrt(float x)
{
	float r;
	asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
	return r;
}

}  // namespace ro
