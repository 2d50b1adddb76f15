// ro_fft2.cuh -- packed-FP32 (f32x2) radix-2^k FFT building blocks for sm_100a.
//
// Blackwell's FMA pipe executes add/mul/fma.f32x2 on a 64-bit register pair:
// one issue slot, two lanes.  A complex number is kept as one packed pair
// (re in the low lane, im in the high lane), so a complex add is ONE FADD2 and
// a complex multiply is FMUL2 + FFMA2; ptxas folds the lane swap, the per-lane
// sign and scalar broadcasts into operand modifiers (.LO_HI, .NP, .F32), so no
// data-movement instruction is emitted (checked with cuobjdump: a 32-point
// transform is 130 FADD2 + 64 FFMA2 + 34 FMUL2 = 228 instructions against 456
// scalar ones).  Measured on B200 (profiles/r01_microbench_pipe_rates.txt):
// packed ops issue at 2.0 warp-instr/clk/SM = the full 128 lanes/clk, whereas
// three-register scalar FFMA reaches only 2.6.
//
// Same transform as ro_fft.cuh (reference: fftw_execute, src/FFTBackend.cpp:236).
#pragma once
#include <cuda_runtime.h>
#include "ro_fft.cuh"

namespace ro {

typedef unsigned long long c64;   // packed complex: {re (low 32 bits), im (high 32 bits)}

__device__ __forceinline__ c64 pk(float lo, float hi)
{
	c64 r;
	asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}
__device__ __forceinline__ void upk(c64 v, float &lo, float &hi)
{
	asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ c64 swp(c64 v) { float a, b; upk(v, a, b); return pk(b, a); }
__device__ __forceinline__ c64 add2(c64 a, c64 b) { c64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 sub2(c64 a, c64 b) { c64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 mul2(c64 a, c64 b) { c64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 fma2(c64 a, c64 b, c64 c) { c64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// z * (wr + i*wi):  (zr*wr - zi*wi, zi*wr + zr*wi)
__device__ __forceinline__ c64 cmul2(c64 z, float wr, float wi)
{
	return fma2(swp(z), pk(-wi, wi), mul2(z, pk(wr, wr)));
}
__device__ __forceinline__ c64 cmul2(c64 z, c64 w)
{
	float wr, wi;
	upk(w, wr, wi);
	return cmul2(z, wr, wi);
}

// In-place forward DIF FFT of R packed points; output k is left at v[bitrev(k)].
template <int R>
__device__ __forceinline__ void fft_dif2(c64 (&v)[R])
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
				const c64 a = v[s + j], b = v[s + j + half];
				const int e = j * (16 / half);            // W_{2*half}^j = W_32^e
				v[s + j] = add2(a, b);
				const c64 d = sub2(a, b);
				if (e == 0) v[s + j + half] = d;
				else if (e == 8) v[s + j + half] = mul2(swp(d), pk(1.f, -1.f));   // * (-i); folded by ptxas
				else v[s + j + half] = cmul2(d, cw32(e), -sw32(e));
			}
		}
	}
}

__device__ __forceinline__ c64 neg2(c64 v) { float a, b; upk(v, a, b); return pk(-a, -b); }

// Same transform, same register convention (input n at v[n], output k at
// v[bitrev(k)]), decimation in TIME: the in-place array a[i] is the view
// v[bitrev(i)], so the bit-reversed input order DIT wants costs nothing.
// Why: a DIT butterfly applies its twiddle BEFORE the add/sub, so the product
// folds into the sum,
//     A' = A + w*B  = fma(swap(B), (s,-s), fma(B, (c,c), A))      w = c - i*s
//     B' = A - w*B  = fma(A, 2, -A')
// three FFMA2 where DIF needs four instructions (add, sub, mul, fma), and a
// multiplication by -i is a lane swap with signs inside the two fmas.  32 points:
// 194 packed instructions instead of 228; 16 points: 74 instead of 84.
template <int R>
__device__ __forceinline__ void fft_dit2(c64 (&v)[R])
{
	static_assert(R >= 2 && R <= 32 && (R & (R - 1)) == 0, "radix");
	constexpr int L = ilog2(R);
	// in v-space level lvl pairs v[i], v[i + (R >> (lvl+1))] like DIF does, but the twiddle is
	// constant over a block: W_{2^(lvl+1)}^j with j = bitrev(block number, lvl)
	constexpr int BR5[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
	                         1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};
#pragma unroll
	for (int lvl = 0; lvl < L; lvl++) {
		const int vhalf = R >> (lvl + 1);
#pragma unroll
		for (int blk = 0; blk < (1 << lvl); blk++) {
			const int e = (BR5[blk] >> (5 - lvl)) * (16 >> lvl);   // W_32^e
#pragma unroll
			for (int off = 0; off < vhalf; off++) {
				const int ia = blk * 2 * vhalf + off, ib = ia + vhalf;
				const c64 a = v[ia], b = v[ib];
				if (e == 0) {
					v[ia] = add2(a, b);
					v[ib] = sub2(a, b);
				} else if (e == 8) {                      // w = -i:  w*b = (b.im, -b.re)
					v[ia] = fma2(swp(b), pk(1.f, -1.f), a);
					v[ib] = fma2(swp(b), pk(-1.f, 1.f), a);
				} else {
					const float c = cw32(e), sn = sw32(e);
					const c64 a2 = fma2(swp(b), pk(sn, -sn), fma2(b, pk(c, c), a));
					v[ia] = a2;
					v[ib] = fma2(a, pk(2.f, 2.f), neg2(a2));
				}
			}
		}
	}
}

// The transform the kernels call.  Measured on B200 (gpurun_out/r01k_sweep_*.jsonl): DIT is 2-5 %
// faster in the one-kernel path (1k..16k), DIF 2.5 % faster in the long-frame path, whose 16k stage
// waits on global loads, not on the FMA pipe.  RO_DIT=0 forces DIF everywhere (A/B measurements).
#ifndef RO_DIT
#define RO_DIT 1
#endif
template <int R, bool DIT>
__device__ __forceinline__ void fft_fwd2(c64 (&v)[R])
{
	if (RO_DIT && DIT) fft_dit2<R>(v);
	else fft_dif2<R>(v);
}

// Multiply output k (at v[bitrev(k)]) by w^k, k = 1..R-1; base[i] = w^(2^i)
// read from the once-rounded table.  Same composition as twiddle_powers().
template <int R>
__device__ __forceinline__ void twiddle_powers2(c64 (&v)[R], const c64 (&base)[ilog2(R)])
{
	constexpr int L = ilog2(R);
	constexpr int LB = (L + 1) / 2;
	constexpr int NLO = 1 << LB;
	constexpr int NHI = 1 << (L - LB);
	c64 lo[NLO];
	lo[0] = pk(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < LB; msb++) {
		lo[1 << msb] = base[msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			lo[(1 << msb) + rest] = cmul2(lo[rest], base[msb]);
	}
	c64 hi[NHI];
	hi[0] = pk(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < L - LB; msb++) {
		hi[1 << msb] = base[LB + msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			hi[(1 << msb) + rest] = cmul2(hi[rest], base[LB + msb]);
	}
#pragma unroll
	for (int h = 0; h < NHI; h++) {
#pragma unroll
		for (int j = 0; j < NLO; j++) {
			const int k = h * NLO + j;
			if (k == 0) continue;
			const c64 w = h == 0 ? lo[j] : (j == 0 ? hi[h] : cmul2(lo[j], hi[h]));
			v[bitrev(k, L)] = cmul2(v[bitrev(k, L)], w);
		}
	}
}

// The R factors twiddle_powers2 multiplies by, w[k] = base-composed w^k (w[0] = 1), with exactly its composition
// (same products in the same order, so a kernel that keeps them and multiplies later gives the same bits).
template <int R>
__device__ __forceinline__ void twiddle_table2(c64 (&w)[R], const c64 (&base)[ilog2(R)])
{
	constexpr int L = ilog2(R);
	constexpr int LB = (L + 1) / 2;
	constexpr int NLO = 1 << LB;
	constexpr int NHI = 1 << (L - LB);
	c64 lo[NLO];
	lo[0] = pk(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < LB; msb++) {
		lo[1 << msb] = base[msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			lo[(1 << msb) + rest] = cmul2(lo[rest], base[msb]);
	}
	c64 hi[NHI];
	hi[0] = pk(1.f, 0.f);
#pragma unroll
	for (int msb = 0; msb < L - LB; msb++) {
		hi[1 << msb] = base[LB + msb];
#pragma unroll
		for (int rest = 1; rest < (1 << msb); rest++)
			hi[(1 << msb) + rest] = cmul2(hi[rest], base[LB + msb]);
	}
#pragma unroll
	for (int h = 0; h < NHI; h++)
#pragma unroll
		for (int j = 0; j < NLO; j++)
			w[h * NLO + j] = h == 0 ? lo[j] : (j == 0 ? hi[h] : cmul2(lo[j], hi[h]));
}

// ---- int16 I/Q -> packed float without the XU pipe --------------------------
// cvt.rn.f32.s16 runs on the 16-lane XU pipe (measured 0.5 warp-instr/clk/SM).
// Instead: flip the sign bits (x + 32768 as unsigned), splice the 16 bits into
// the mantissa of 2^23 with PRMT, subtract 2^23 + 32768: exact, ALU + FMA pipes.
// (Round 2 tried the XU pipe again now that the kernels sit on the FMA pipe -- two I2F.S16, the second with the
// .H1 selector, no other instruction: 1k -5 %, 8k -7 %, 16k -1.4 %, the rest +-0.5 %, profiles/r02_cvt_xu_ab.txt.)
__device__ __forceinline__ c64 short2_to_c64(unsigned int v)
{
	const unsigned int u = v ^ 0x80008000u;
	const float re = __uint_as_float(__byte_perm(u, 0x4b000000u, 0x7410));
	const float im = __uint_as_float(__byte_perm(u, 0x4b000000u, 0x7432));
	return add2(pk(re, im), pk(-8421376.0f, -8421376.0f));
}

}  // namespace ro
