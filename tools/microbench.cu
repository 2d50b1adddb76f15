// microbench.cu -- pipe-rate probes that size the FFT kernel's FP32 / shared-memory roof on B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/microbench tools/microbench.cu
// Each probe runs 148 CTAs (one per SM) of T threads; every thread owns CH independent dependency
// chains and executes U*CH operations per loop trip, so loop overhead is < 3 %.  Rates are computed
// from CTA 0's clock64() span: warp-instructions per cycle per SM.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

#define ITER 512
#define CH 12
#define U 8

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { float2 v = make_float2(a, b); return *reinterpret_cast<u64 *>(&v); }

template <int OP>
__global__ void __launch_bounds__(1024, 1) k_fp(float *out, const float *in, long long *cyc)
{
	float a[CH], b[CH], c[CH];
	u64 a2[CH], b2[CH], c2[CH];
	for (int i = 0; i < CH; i++) {
		a[i] = threadIdx.x + i; b[i] = in[i]; c[i] = in[i + CH];
		a2[i] = pk(a[i], a[i] + 1); b2[i] = pk(b[i], b[i]); c2[i] = pk(c[i], c[i]);
	}
	__syncthreads();
	long long t0 = clock64();
#pragma unroll 1
	for (int it = 0; it < ITER; it++) {
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < CH; i++) {
				if (OP == 0) a[i] = fmaf(a[i], b[i], c[i]);
				if (OP == 1) a[i] = a[i] + b[i];
				if (OP == 2) a[i] = a[i] * b[i];
				if (OP == 3) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a2[i]) : "l"(b2[i]), "l"(c2[i]));
				if (OP == 4) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a2[i]) : "l"(b2[i]));
				if (OP == 5) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(a2[i]) : "l"(b2[i]));
				if (OP == 6) { if (i & 1) a[i] = a[i] + b[i]; else a[i] = fmaf(a[i], b[i], c[i]); }
				if (OP == 7) { if (i & 1) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a2[i]) : "l"(b2[i]));
				               else a[i] = fmaf(a[i], b[i], c[i]); }
				if (OP == 8) a[i] = fmaf(a[i], 1.0001f, 0.5f);
			}
		}
	}
	long long t1 = clock64();
	float s = 0;
	for (int i = 0; i < CH; i++) { float2 x = *reinterpret_cast<float2 *>(&a2[i]); s += a[i] + x.x + x.y; }
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// shared memory: conflict-free loads of W floats per lane, address fixed per thread (no dependence)
template <int W>
__global__ void __launch_bounds__(1024, 1) k_lds(float *out, long long *cyc)
{
	extern __shared__ float sm[];
	for (int i = threadIdx.x; i < 16384; i += blockDim.x) sm[i] = i;
	__syncthreads();
	float s[CH];
	for (int i = 0; i < CH; i++) s[i] = 0;
	const float *p = sm + (threadIdx.x & 255) * W;
	long long t0 = clock64();
#pragma unroll 1
	for (int it = 0; it < ITER; it++) {
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < CH; i++) {
				const volatile float *q = p + ((i * 1024 + u * 256 * W) & 8191);
				if (W == 1) s[i] += q[0];
				if (W == 2) { float2 v; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared((const void *)q)) : "memory"); s[i] += v.x; }
				if (W == 4) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared((const void *)q)) : "memory"); s[i] += v.x; }
			}
		}
	}
	long long t1 = clock64();
	float r = 0; for (int i = 0; i < CH; i++) r += s[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = r;
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int W>
__global__ void __launch_bounds__(1024, 1) k_sts(float *out, long long *cyc)
{
	extern __shared__ float sm[];
	__syncthreads();
	float *p = sm + (threadIdx.x & 255) * W;
	const float x = threadIdx.x;
	long long t0 = clock64();
#pragma unroll 1
	for (int it = 0; it < ITER; it++) {
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < CH; i++) {
				float *q = p + ((i * 1024 + u * 256 * W) & 8191);
				const unsigned a = (unsigned)__cvta_generic_to_shared(q);
				if (W == 1) asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(x));
				if (W == 2) asm volatile("st.shared.v2.f32 [%0], {%1,%2};" :: "r"(a), "f"(x), "f"(x));
				if (W == 4) asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" :: "r"(a), "f"(x), "f"(x), "f"(x), "f"(x));
			}
		}
	}
	long long t1 = clock64();
	__syncthreads();
	out[blockIdx.x * blockDim.x + threadIdx.x] = sm[threadIdx.x];
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>   // 0: cvt.rn.f32.s16, 1: magic-number PRMT+FADD
__global__ void __launch_bounds__(1024, 1) k_cvt(float *out, const int *in, long long *cyc)
{
	int a[CH]; float f[CH];
	for (int i = 0; i < CH; i++) { a[i] = in[i] + threadIdx.x; f[i] = 0; }
	long long t0 = clock64();
#pragma unroll 1
	for (int it = 0; it < ITER; it++) {
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int i = 0; i < CH; i++) {
				if (MODE == 0) { f[i] += (float)(short)(a[i] & 0xffff); }
				else { f[i] += __uint_as_float(__byte_perm(a[i] ^ 0x8000, 0x4b000000, 0x7410)) - 8421376.0f; }
				a[i] = __float_as_int(f[i]);
			}
		}
	}
	long long t1 = clock64();
	float s = 0; for (int i = 0; i < CH; i++) s += f[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <typename F>
static void run(const char *name, F launch, int threads, double ops_per_thread)
{
	long long *d_cyc; cudaMalloc(&d_cyc, 8);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	launch(148, threads, d_cyc);
	cudaDeviceSynchronize();
	cudaEventRecord(e0);
	launch(148, threads, d_cyc);
	cudaEventRecord(e1);
	cudaDeviceSynchronize();
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	long long cyc; cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
	const double warp_instr = threads / 32.0 * ops_per_thread;
	printf("%-34s warps/SM %2d  cycles %8lld  warp-instr/cyc/SM %6.3f  (%.3f ms)  %s\n", name,
	       threads / 32, cyc, warp_instr / (double)cyc, ms, cudaGetErrorString(cudaGetLastError()));
	cudaFree(d_cyc);
}

int main()
{
	float *out, *in; int *iin;
	cudaMalloc(&out, 148 * 1024 * 4);
	cudaMalloc(&in, 4096); cudaMalloc(&iin, 4096);
	float h[64]; for (int i = 0; i < 64; i++) h[i] = 1.0f + 1e-6f * i;
	cudaMemcpy(in, h, sizeof h, cudaMemcpyHostToDevice);
	cudaMemset(iin, 0, 4096);
	const double ops = (double)ITER * U * CH;
	for (int th = 128; th <= 1024; th *= 2) {
#define FP(OPN, NAME) run(NAME, [&](int g, int t, long long *c) { k_fp<OPN><<<g, t>>>(out, in, c); }, th, ops)
		FP(0, "FFMA r,r,r"); FP(8, "FFMA r,imm,imm"); FP(1, "FADD"); FP(2, "FMUL");
		FP(3, "FFMA2 (fma.rn.f32x2)"); FP(4, "FADD2 (add.rn.f32x2)"); FP(5, "FMUL2 (mul.rn.f32x2)");
		FP(6, "FADD/FFMA alternating"); FP(7, "FADD2/FFMA alternating");
		run("cvt.f32.s16 + FADD", [&](int g, int t, long long *c) { k_cvt<0><<<g, t>>>(out, iin, c); }, th, ops);
		run("PRMT magic + 2 FADD", [&](int g, int t, long long *c) { k_cvt<1><<<g, t>>>(out, iin, c); }, th, ops);
		run("LDS.32", [&](int g, int t, long long *c) { k_lds<1><<<g, t, 65536>>>(out, c); }, th, ops);
		run("LDS.64", [&](int g, int t, long long *c) { k_lds<2><<<g, t, 65536>>>(out, c); }, th, ops);
		run("LDS.128", [&](int g, int t, long long *c) { k_lds<4><<<g, t, 65536>>>(out, c); }, th, ops);
		run("STS.32", [&](int g, int t, long long *c) { k_sts<1><<<g, t, 65536>>>(out, c); }, th, ops);
		run("STS.64", [&](int g, int t, long long *c) { k_sts<2><<<g, t, 65536>>>(out, c); }, th, ops);
		run("STS.128", [&](int g, int t, long long *c) { k_sts<4><<<g, t, 65536>>>(out, c); }, th, ops);
	}
	return 0;
}
