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

// -DRO_TRACE=1 (tools/trace16k.py, never the product build): lane 0 of every warp of CTAs 0 and 73 of the
// staged 16k kernel stamps clock64() at the phase boundaries of frames RO_TRACE_F0 .. +RO_TRACE_NF-1 of the CTA
// into a file-scope array that ro_trace_dump (ro_v2.cu) copies out: the per-warp timeline of a frame.
#ifndef RO_TRACE
#define RO_TRACE 0
#endif
#if RO_TRACE
#define RO_TRACE_F0 16
#define RO_TRACE_NF 8
#define RO_TRACE_NP 20
static __device__ unsigned long long ro_trace_buf[2 * RO_TRACE_NF * 16 * RO_TRACE_NP];
#define RO_TP(i) do { \
	if ((MODE == MODE_STAGED || MODE == MODE_SPLIT || MODE == MODE_FUSED2) && LOG2N == 14 && (t & 31) == 0 && f - f0 >= RO_TRACE_F0 && f - f0 < RO_TRACE_F0 + RO_TRACE_NF && \
	    (blockIdx.x == 0 || blockIdx.x == 73)) \
		ro_trace_buf[(((blockIdx.x ? 1 : 0) * RO_TRACE_NF + (int)(f - f0) - RO_TRACE_F0) * 16 + (t >> 5)) * RO_TRACE_NP + (i)] = clock64(); \
} while (0)
#else
#define RO_TP(i) do { } while (0)
#endif

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
	int           log2na;       // MODE_SPLIT: the frame has Na = 2^log2na residue classes of N points each
	int           stagger;      // serialise the shared-memory load bursts of the warps that share a scheduler
	int           pair;         // MODE_SPLIT: rows are 16-byte aligned and every CTA starts on an even item:
	                            // classes r, r+1 are stored together (see the store stage); 2: every CTA starts on
	                            // a multiple of four and classes 4j .. 4j+3 are stored together (Na >= 4)
	// MODE_FUSED*: the first radix-Na step happens in this kernel (see the load stage)
	const void   *raw;          // int16 I/Q of the whole batch
	const float  *window_full;  // Na*N floats
	const float2 *tw_big;       // Na*N entries exp(-2*pi*i*j/(Na*N))
	unsigned long long *scratch;// [gridDim.x][Na-1][N] packed complex, private to each CTA (stays in L2)
};

// where the N points of one item come from / where its magnitudes go
enum { MODE_DIRECT = 0,   // int16 I/Q read from global memory, row = item
       MODE_STAGED = 1,   // int16 I/Q through the TMA-staged shared-memory ring, row = item
       MODE_SPLIT  = 2,   // packed complex float from wf_split_kernel (already windowed, first
                          // radix-Na step done): item = frame*Na + r, output k is bin Na*k + r
       MODE_FUSED1 = 3,   // like MODE_SPLIT with Na = 2, but the radix-2 step is done here, per frame, by the
                          // thread that needs its results: no scratch round trip through HBM
       MODE_FUSED2 = 4,   // the same for Na = 4 (64k frames): radix-4 step here, classes 1 and 2 wait in tensor memory,
                          // class 3 in a CTA-private slice of global memory, four classes stored together
       MODE_DIRECT_F32 = 5,   // MODE_DIRECT for float32 I/Q (the reference's raw-TCP / JACK encoding)
       MODE_FUSED1_F32 = 6 }; // MODE_FUSED1 for float32 I/Q
__host__ __device__ constexpr bool mode_is_fused(int m) { return m == MODE_FUSED1 || m == MODE_FUSED1_F32 || m == MODE_FUSED2; }
__host__ __device__ constexpr bool mode_is_split(int m) { return m == MODE_SPLIT || mode_is_fused(m); }
__host__ __device__ constexpr bool mode_is_f32(int m) { return m == MODE_DIRECT_F32 || m == MODE_FUSED1_F32; }
// MODE_FUSED*: bytes of the raw-sample / window staging ring beside the 64 KB stash, and of its barriers
constexpr size_t FUSED_RING_BYTES = 32768, FUSED_BAR_BYTES = 512;
// ... and of the table of pass-2 twiddle bases of the 32k kernels (16 values of n3 x 5 bases), see pass 2
constexpr size_t FUSED_B2_BYTES = 16 * 5 * 8;
template <bool F32IN> struct raw_word { typedef unsigned int type; };          // one int16 I/Q pair
template <> struct raw_word<true> { typedef unsigned long long type; };       // one float32 I/Q pair

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
__device__ __forceinline__ bool mbar_test(unsigned long long *bar, unsigned int parity)
{
	unsigned int ok;
	asm volatile(
		"{\n\t.reg .pred p;\n\t"
		"mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
		"selp.u32 %0, 1, 0, p;\n\t}"
		: "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
	return ok != 0;
}
// Warps w, w+4, w+8, ... share one of the SM's four schedulers.  Left alone they issue their
// shared-memory load bursts in lock step, receive their data together, then fight for the FMA
// pipe together: load time and FMA time add up.  With the chain below the warp of level l starts
// its burst only when the warp of level l-1 on the same scheduler has issued all of its own, so
// the first warp computes while the next ones load.  Named barriers 3*q + l (q = scheduler,
// l = level 1..3), 64 threads each: the producer arrives without blocking, the consumer syncs.
// Used in passes 2 and 3 only (stagger bits 1 and 2): two uses of one barrier are always separated
// by a __syncthreads() there.  Pass 1 follows pass 3 of the previous frame without a CTA-wide
// barrier, its warps arrive staggered already, and chaining it measured 5 % slower.
__device__ __forceinline__ void chain_wait(int q, int l, int dist = 1)
{
	if (l >= dist) asm volatile("bar.sync %0, 64;" :: "r"(3 * q + l) : "memory");
}
__device__ __forceinline__ void chain_pass(int q, int l, int levels, int dist = 1)
{
	if (l + dist < levels) asm volatile("bar.arrive %0, 64;" :: "r"(3 * q + l + dist) : "memory");
}
// Arrival on a split barrier ("my inputs of this pass are in registers"), every thread for itself.
// -DRO_WARP_ARRIVE=1 lets lane 0 arrive for its warp after a __syncwarp() (16 arrivals instead of 512 on
// one mbarrier): measured SLOWER on B200 at every size (16k 26.6 -> 25.7, 8k 51.2 -> 46.5, 1k 426 -> 403
// M frames/s): the __syncwarp() costs more than the arrivals, which the hardware evidently combines.
#ifndef RO_WARP_ARRIVE
#define RO_WARP_ARRIVE 0
#endif
__device__ __forceinline__ void split_arrive(unsigned long long *bar, int t)
{
	if (RO_WARP_ARRIVE) {
		__syncwarp();
		if ((t & 31) == 0) mbar_arrive(bar);
	} else {
		mbar_arrive(bar);
	}
}
// A fence for ptxas.  Nothing at run time (`never` is a kernel parameter that is always 0; if it
// ever ran it would be a warp-wide barrier on id 15, which is harmless), but a barrier instruction
// in the stream keeps ptxas from interleaving a load burst with the arithmetic around it.  Found by
// accident: with the burst of pass 1 fenced the 16k kernel spills 0 bytes instead of 104 and a
// frame takes 21 % fewer cycles.
__device__ __forceinline__ void ptxas_fence(int never)
{
	if (never) asm volatile("bar.sync 15, 32;" ::: "memory");
}
// Further fence sites, selected at compile time (-DRO_FENCE_SITES=mask for the 8k and 16k kernels,
// -DRO_FENCE_SITES_SMALL=mask for the others).  Bit 4 = after each R3-point transform of pass 3:
// measured +3 % on the 16k kernel (25.2 -> 26.1 M frames/s) and +2.5 % on the 8k one, -4 % at 1k;
// every other site within noise.
#ifndef RO_FENCE_SITES
#define RO_FENCE_SITES 16
#endif
#ifndef RO_FENCE_SITES_SMALL
#define RO_FENCE_SITES_SMALL 0
#endif
#define RO_SITE(bit, never) do { if ((LOG2N >= 13 ? RO_FENCE_SITES : RO_FENCE_SITES_SMALL) & (1 << (bit))) ptxas_fence(never); } while (0)
// global -> shared bulk copy (TMA), bytes a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned int bytes, unsigned long long *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
	             :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- tensor memory as a per-thread parking space -------------------------------------------------
// The 256 KB of TMEM of an SM (512 columns x 128 lanes x 32 bit) are idle in these kernels: no tcgen05.mma
// anywhere.  tcgen05.st / tcgen05.ld with shape 32x32b move N registers of every thread of a warp to / from N
// consecutive columns of "its" lane (lane quarter = warp % 4), so a 512-thread CTA has 128 private words per
// thread there -- room for the magnitudes of three residue classes (3 x 32 floats), which shared memory
// (128 KB exchange buffer + 64 KB stash) cannot hold.
__device__ __forceinline__ void tmem_st8(unsigned int taddr, const float (&v)[8])
{
	asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
	             :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
	                "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
	                "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])) : "memory");
}
// N packed complex values (2N words) of this thread to columns taddr .. taddr + 2N - 1
__device__ __forceinline__ void tmem_st_c64(unsigned int taddr, const c64 (&v)[1])
{
	asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};"
	             :: "r"(taddr), "r"((unsigned int)v[0]), "r"((unsigned int)(v[0] >> 32)) : "memory");
}
__device__ __forceinline__ void tmem_st_c64(unsigned int taddr, const c64 (&v)[2])
{
	asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
	             :: "r"(taddr), "r"((unsigned int)v[0]), "r"((unsigned int)(v[0] >> 32)),
	                "r"((unsigned int)v[1]), "r"((unsigned int)(v[1] >> 32)) : "memory");
}
// 16 packed complex values back from columns taddr .. taddr + 31 (no wait: tmem_wait_ld() before the first use)
__device__ __forceinline__ void tmem_ld_c64x16(unsigned int taddr, c64 *v)
{
	unsigned int r[32];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
	             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	             "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
	               "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
	               "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
	             : "r"(taddr) : "memory");
#pragma unroll
	for (int i = 0; i < 16; i++) v[i] = (c64)r[2 * i] | ((c64)r[2 * i + 1] << 32);
}
// 8 packed complex values to / from columns taddr .. taddr + 15
__device__ __forceinline__ void tmem_st_c64x8(unsigned int taddr, const c64 *v)
{
	asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
	             "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
	             :: "r"(taddr),
	                "r"((unsigned int)v[0]), "r"((unsigned int)(v[0] >> 32)), "r"((unsigned int)v[1]), "r"((unsigned int)(v[1] >> 32)),
	                "r"((unsigned int)v[2]), "r"((unsigned int)(v[2] >> 32)), "r"((unsigned int)v[3]), "r"((unsigned int)(v[3] >> 32)),
	                "r"((unsigned int)v[4]), "r"((unsigned int)(v[4] >> 32)), "r"((unsigned int)v[5]), "r"((unsigned int)(v[5] >> 32)),
	                "r"((unsigned int)v[6]), "r"((unsigned int)(v[6] >> 32)), "r"((unsigned int)v[7]), "r"((unsigned int)(v[7] >> 32))
	             : "memory");
}
__device__ __forceinline__ void tmem_ld_c64x8(unsigned int taddr, c64 (&v)[8])
{
	unsigned int r[16];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
	             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	             : "r"(taddr) : "memory");
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
	for (int i = 0; i < 8; i++) v[i] = (c64)r[2 * i] | ((c64)r[2 * i + 1] << 32);
}
// N floats from columns taddr .. taddr + N - 1 (N = 2, 4; no wait)
__device__ __forceinline__ void tmem_ld_f(unsigned int taddr, float (&v)[2])
{
	unsigned int a, b;
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(taddr) : "memory");
	v[0] = __uint_as_float(a); v[1] = __uint_as_float(b);
}
__device__ __forceinline__ void tmem_ld_f(unsigned int taddr, float (&v)[4])
{
	unsigned int a, b, c, d;
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
	v[0] = __uint_as_float(a); v[1] = __uint_as_float(b); v[2] = __uint_as_float(c); v[3] = __uint_as_float(d);
}
__device__ __forceinline__ void tmem_ld_f(unsigned int taddr, float (&v)[8])
{
	unsigned int r[8];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
#pragma unroll
	for (int i = 0; i < 8; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st_f(unsigned int taddr, const float (&v)[8])
{
	tmem_st8(taddr, v);
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(unsigned int taddr, float (&v)[32])
{
	unsigned int r[32];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
	             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
	             "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
	               "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
	               "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
	             : "r"(taddr) : "memory");
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
	for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

// one raw I/Q pair -> packed complex float
__device__ __forceinline__ c64 raw_to_c64(unsigned int w) { return short2_to_c64(w); }
__device__ __forceinline__ c64 raw_to_c64(unsigned long long w) { return w; }

// columns of tensor memory a CTA of the kernel allocates (0: none); see TW_TMEM / SPLIT_TMEM / FUSED_TMEM in the kernel
__host__ __device__ constexpr int wf2_tmem_cols_of(int log2n, int threads, bool split_mode)
{
	// every user is a 16k kernel: 512 threads, one CTA per SM, all 512 columns, 128 per thread
	(void)split_mode;
	return (log2n == 14 && threads == 512) ? 512 : 0;
}
template <int LOG2N, int P, int MODE>
__host__ __device__ constexpr int wf2_tmem_cols()
{
	return wf2_tmem_cols_of(LOG2N, (1 << LOG2N) / P, mode_is_split(MODE));
}

// MODE_SPLIT, 16k items: rows n1 = 0 .. SPLIT_LROWS-1 (4 KB each) of the NEXT item are brought into shared memory by
// cp.async.bulk while this one is transformed (see the load stage); 16 rows fit the magnitude stash, which the
// four-class stores leave idle, 8 more take the 32 KB that were free.  -DRO_SPLIT_LROWS=0: off (A/B).
#ifndef RO_SPLIT_LROWS
#define RO_SPLIT_LROWS 24
#endif
template <int LOG2N, int P, int R3, int MODE>
__host__ __device__ constexpr int wf2_split_lrows()
{
	return (MODE == MODE_SPLIT && LOG2N == 14 && P == 32 && R3 == 16) ? RO_SPLIT_LROWS : 0;
}

// bytes of dynamic shared memory the kernel needs
template <int LOG2N, int P, int R3, int MODE>
constexpr size_t wf2_smem_bytes()
{
	constexpr size_t N = (size_t)1 << LOG2N;
	constexpr size_t PAD1 = (R3 < 16 && !(R3 == 8 && P == 32)) ? R3 : 0;   // 8k: swizzled instead of padded, see the kernel
	return sizeof(c64) * (N + P * PAD1) + (mode_is_split(MODE) ? 0 : sizeof(float) * (N / 2)) + 32 +
	       (MODE == MODE_STAGED ? 4 * N : 0) + (mode_is_split(MODE) && P / R3 >= 2 ? 4 * N : 0) +
	       (mode_is_fused(MODE) ? FUSED_RING_BYTES + FUSED_BAR_BYTES + FUSED_B2_BYTES : 0) +
	       (wf2_split_lrows<LOG2N, P, R3, MODE>() > 16 ? sizeof(c64) * (N / P) * (wf2_split_lrows<LOG2N, P, R3, MODE>() - 16) : 0);
}

template <int LOG2N, int P, int R3, int MODE, bool GAIN, int MINB>
__global__ void __launch_bounds__((1 << LOG2N) / P, MINB)
wf_fft2_kernel(const Wf2Params p)
{
	constexpr bool STAGE = MODE == MODE_STAGED;
	// prefetch.global.L2 of what the NEXT frame / item will load from HBM, issued while this one is
	// transformed (the kernels that read through the LSU have nothing else in flight at that point):
	// float32 16k at 75 % overlap 22.8 -> 25.3 M frames/s, fused 32k 7.21 -> 7.54
#ifndef RO_L2_PREFETCH
#define RO_L2_PREFETCH 1
#endif
	constexpr bool FIN32 = MODE == MODE_FUSED1_F32;   // fused, float32 input
	constexpr bool FUSED = mode_is_fused(MODE);
	constexpr bool F32 = MODE == MODE_DIRECT_F32;
	constexpr int FLA = MODE == MODE_FUSED2 ? 2 : FUSED ? 1 : 0;   // log2(Na) of the fused modes
	constexpr bool SPLIT = MODE == MODE_SPLIT || FUSED;
	// MODE_SPLIT, 16k items, Na >= 4: the magnitudes of three residue classes wait in tensor memory for the
	// fourth, see the store stage.  -DRO_SPLIT_TMEM=0 for the A/B.
#ifndef RO_SPLIT_TMEM
#define RO_SPLIT_TMEM 1
#endif
	constexpr bool SPLIT_TMEM = RO_SPLIT_TMEM && MODE == MODE_SPLIT && LOG2N == 14 && R3 == 16;
	// MODE_FUSED1*: the odd class y_1 of a frame (32 packed complex per thread) waits in tensor memory for its turn
	// instead of in a CTA-private slice of global memory (128 KB out to L2 and 128 KB back per frame, of the 576 KB
	// a frame moved in all).  -DRO_FUSED_TMEM=0 for the A/B.
#ifndef RO_FUSED_TMEM
#define RO_FUSED_TMEM 1
#endif
	constexpr bool FUSED_TMEM = RO_FUSED_TMEM && FUSED && LOG2N == 14 && R3 == 16;
	// ... and next to it (columns 64 .. 127) the 2 x 32 window values the thread ever uses, which used to travel with
	// every TMA burst.  -DRO_FUSED_WIN_TMEM=0: they do again (A/B).
#ifndef RO_FUSED_WIN_TMEM
#define RO_FUSED_WIN_TMEM 1
#endif
	constexpr bool WIN_TMEM = RO_FUSED_WIN_TMEM && FUSED_TMEM && MODE != MODE_FUSED2;   // Na = 4: both halves hold parked classes
	// One-kernel modes, 16k: the 31 + 31 twiddle factors a thread multiplies its pass-1 and pass-2 outputs by are
	// the same for every frame (W_N^(t*k1), W_512^(n3*k2)).  They used to be rebuilt from five table entries each
	// in every frame (26 complex multiplications per pass before the 31 that apply them: 104 of the 824 packed
	// instructions of a frame); now they are composed once, the same way, parked in the thread's 124 words of
	// tensor memory and read back eight at a time.  -DRO_TW_TMEM=0 for the A/B.
#ifndef RO_TW_TMEM
#define RO_TW_TMEM 1
#endif
	// 16k only.  The same for the smaller kernels (4*P words per thread, 4*P*T/128 columns per CTA: 256 at 8k,
	// 128 at 4k, 64 at 2k) measured SLOWER on B200: 8k +0.5 % (int16) / -8 % (float32), 4k -13 %, 2k -40 %
	// (profiles/r02_tw_tmem_ab.txt): their transforms are too short to cover the tcgen05.ld round trips.
	// The host never asks for more resident CTAs than 512 columns allow (v2_tmem_cols, ro_cuda.cu).
	constexpr bool TW_TMEM = RO_TW_TMEM && !SPLIT && wf2_tmem_cols<LOG2N, P, MODE>() > 0;
	constexpr bool USE_TMEM = SPLIT_TMEM || FUSED_TMEM || TW_TMEM;
	// The split path's kernel has 24 words of tensor memory to spare (columns 56..63, 88..95, 120..127 are used
	// neither by parked pairs nor by parked fours): its ten twiddle bases wait there instead of in registers that
	// ptxas spills.  -DRO_SPLIT_BASES_TMEM=0: registers (A/B).
#ifndef RO_SPLIT_BASES_TMEM
#define RO_SPLIT_BASES_TMEM 1
#endif
	// rows of the next item prefetched into shared memory (needs the four-class stores: p.pair == 2)
	constexpr int LROWS = SPLIT_TMEM ? wf2_split_lrows<LOG2N, P, R3, MODE>() : 0;
	static_assert(SPLIT_TMEM || wf2_split_lrows<LOG2N, P, R3, MODE>() == 0, "the landing area is the stash: parked magnitudes must be in tensor memory");
	constexpr int TMEM_COLS = wf2_tmem_cols<LOG2N, P, MODE>();
	constexpr int TMEM_T = (1 << LOG2N) / P;                                                   // threads of the CTA
	constexpr int TMEM_PER_LEVEL = (TMEM_T >= 128) ? TMEM_COLS / (TMEM_T / 128) : TMEM_COLS;  // columns of one thread
#ifndef RO_FUSED_DIT
#define RO_FUSED_DIT 1
#endif
	constexpr bool USE_DIT = !SPLIT || (FUSED && RO_FUSED_DIT);
	constexpr int WINF = SPLIT ? 0 : (1 << LOG2N) / 2;      // floats of window table in shared memory
	constexpr int N = 1 << LOG2N;
	constexpr int R1 = P, R2 = P;
	constexpr int T = N / P;
	constexpr int M = N / R1;                  // = R2*R3
	static_assert(R1 * R2 * R3 == N, "factorisation");
	static_assert(R3 <= 16 && R3 >= 2 && P % R3 == 0, "R3");
	// Pass 2 reads row k1 of S with the R3 lanes (k1, n3): a half-warp covers 16 / R3 rows, which must fall
	// into different banks.  R3 = 16: one row per half-warp.  R3 = 4: rows padded by R3 elements.  R3 = 8:
	// bit 3 of the position inside the row is XORed with the row's parity instead -- the 2 KB of padding
	// were what kept the 8k kernel at ONE CTA per SM (116 768 B of shared memory each, ncu
	// launch__occupancy_limit_shared_mem = 1; 114 720 B fit twice).
	constexpr bool SWZ1 = R3 == 8 && P == 32;        // the 2k kernel (P = 16) is not short of shared memory: padded, 1.7 % faster
	constexpr int PAD1 = (R3 < 16 && !SWZ1) ? R3 : 0;
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
	unsigned long long *mbar = reinterpret_cast<unsigned long long *>(win + WINF);
	unsigned long long *bar_rd2 = mbar + 1, *bar_rd3 = mbar + 2;
	unsigned int *ring = reinterpret_cast<unsigned int *>(mbar + 4);

	const int t = threadIdx.x;
	constexpr int LEVELS = T / 128;            // warps per scheduler
	constexpr bool CHAIN = LEVELS >= 2 && LEVELS <= 4;
	const int wq = (t >> 5) & 3, wl = t >> 7;
	const int cdist = (p.stagger & 8) ? 2 : 1, cdist3 = (p.stagger & 16) ? 2 : 1;
	const long long f0 = (long long)blockIdx.x * p.frames_per_cta;
	long long f1 = f0 + p.frames_per_cta;
	if (f1 > p.n_frames) f1 = p.n_frames;
	if (f0 >= f1) return;

	const unsigned char *iq8 = reinterpret_cast<const unsigned char *>(p.iq);
	const long long hop = p.hop;

	if (t == 0) {
		mbar_init(mbar, 1);
		mbar_init(bar_rd2, RO_WARP_ARRIVE ? T / 32 : T);
		mbar_init(bar_rd3, RO_WARP_ARRIVE ? T / 32 : T);
	}
	for (int i = t; i < WINF; i += T) win[i] = __ldg(p.window_half + i);
	unsigned int tmem = 0;
	if constexpr (USE_TMEM) {
		// the long-frame modes take all 512 columns: one CTA per SM (shared memory), nobody else on the SM wants any
		unsigned int *slot = reinterpret_cast<unsigned int *>(mbar + 3);
		if (t < 32) {
			asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(slot)), "n"(TMEM_COLS) : "memory");
			asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
		}
		asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
		__syncthreads();
		asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
		// this thread's 128 columns: lane quarter of its warp, column block of its warp's level
		tmem = *slot + ((unsigned int)(((t >> 5) & 3) * 32) << 16) + (unsigned int)((t >> 7) * TMEM_PER_LEVEL);
	}
	__syncthreads();
	if (STAGE && t == 0) {
		mbar_expect_tx(mbar, 4u * N);
		const unsigned char *src = iq8 + 4 * f0 * hop;
#pragma unroll 1
		for (unsigned int o = 0; o < 4u * N; o += CHUNK)
			bulk_g2s(reinterpret_cast<unsigned char *>(ring) + o, src + o, CHUNK < 4u * N - o ? CHUNK : 4u * N - o, mbar);
	}
	unsigned int parity = 0;
	// MODE_SPLIT, 16k: rows 0 .. LROWS-1 of item f into the landing area behind mbar + 4 (the magnitude stash of the
	// other split instantiations: here parked magnitudes live in tensor memory, pairs and fours alike)
	constexpr bool prefetch_rows = LROWS > 0;
	// Measured and not kept (B200, profiles/r02_split_late_rows_ab.txt, r02_split_deferred_stores_ab.txt): the last
	// R1 - LROWS rows by a second bulk copy straight into S once every thread holds its pass-3 inputs of the item
	// before (-13 %); the fourth class's row stores deferred into the next item, half in front of each CTA barrier (-4 %).
	auto fetch_rows = [&](long long f) {
		if constexpr (LROWS > 0) {
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			constexpr unsigned int BYTES = (unsigned int)LROWS * (unsigned int)M * 8u;
			mbar_expect_tx(mbar, BYTES);
			const unsigned char *src = iq8 + 8 * f * N;
			unsigned char *dst = reinterpret_cast<unsigned char *>(mbar + 4);
#pragma unroll 1
			for (unsigned int o = 0; o < BYTES; o += CHUNK)
				bulk_g2s(dst + o, src + o, CHUNK < BYTES - o ? CHUNK : BYTES - o, mbar);
		}
	};
	if (prefetch_rows && t == 0) fetch_rows(f0);

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
	// Fused 32k kernels: tensor memory is full (parked class + window) and ptxas spills 100 .. 120 bytes; the
	// pass-2 bases depend on n3 = t % 16 only, so they are read from a 640-byte table in shared memory per pass
	// instead of living in ten registers.  -DRO_FUSED_B2_SMEM=0: registers (A/B).
#ifndef RO_FUSED_B2_SMEM
#define RO_FUSED_B2_SMEM 1
#endif
	constexpr bool B2_SMEM = RO_FUSED_B2_SMEM && FUSED && LP == 5 && R3 == 16;        // (64k: pass 1's bases still come by LDG)
	c64 *b2tab = reinterpret_cast<c64 *>(reinterpret_cast<unsigned char *>(mbar + 4) + 4 * N + FUSED_RING_BYTES + FUSED_BAR_BYTES);
	if constexpr (B2_SMEM) {
		if (t < 16 * LP) b2tab[t] = __ldg(reinterpret_cast<const unsigned long long *>(p.tw) + ((((t / LP) * P) << (t % LP)) & NMASK));
		__syncthreads();
	}
	if constexpr (RO_SPLIT_BASES_TMEM && SPLIT_TMEM && LP == 5) {
		float q8[8];
		float lo, hi;
#pragma unroll
		for (int i = 0; i < 4; i++) { upk(b1[i], lo, hi); q8[2 * i] = lo; q8[2 * i + 1] = hi; }
		tmem_st8(tmem + 56u, q8);
#pragma unroll
		for (int i = 0; i < 4; i++) { upk(b2[i], lo, hi); q8[2 * i] = lo; q8[2 * i + 1] = hi; }
		tmem_st8(tmem + 120u, q8);
		const c64 last[2] = {b1[4], b2[4]};
		tmem_st_c64(tmem + 88u, last);          // b1[4] at 88..89, b2[4] at 90..91
		tmem_wait_st();
	}
	if constexpr (TW_TMEM) {
		c64 w[P];
		twiddle_table2<R1>(w, b1);
#pragma unroll
		for (int c = 0; c < P / 8; c++) tmem_st_c64x8(tmem + (unsigned int)(16 * c), &w[8 * c]);
		twiddle_table2<R2>(w, b2);
#pragma unroll
		for (int c = 0; c < P / 8; c++) tmem_st_c64x8(tmem + (unsigned int)(2 * P + 16 * c), &w[8 * c]);
		tmem_wait_st();
	}
	if constexpr (WIN_TMEM) {
		// the 2 x 32 window values this thread ever uses, w[q*N + n1*M + t], to columns 64 + 2*n1 + q: they used to
		// travel with every burst (two of the six copies, a third of its bytes)
#pragma unroll
		for (int c = 0; c < R1 / 4; c++) {
			float wq[8];
#pragma unroll
			for (int i = 0; i < 4; i++) {
				wq[2 * i] = __ldg(p.window_full + (4 * c + i) * M + t);
				wq[2 * i + 1] = __ldg(p.window_full + N + (4 * c + i) * M + t);
			}
			tmem_st_f(tmem + 64u + (unsigned int)(8 * c), wq);
		}
		tmem_wait_st();
	}
	// v[bitrev(k)] *= (parked factor k), k = 1 .. P-1; col0: first column of the pass's table
	auto twiddle_tmem = [&](c64 (&v)[P], unsigned int col0) {
		if constexpr (TW_TMEM) {
#pragma unroll
			for (int c = 0; c < P / 8; c++) {
				c64 w[8];
				tmem_ld_c64x8(tmem + col0 + (unsigned int)(16 * c), w);
#pragma unroll
				for (int j = 0; j < 8; j++) {
					const int k = 8 * c + j;
					if (k == 0) continue;
					v[bitrev(k, LP)] = cmul2(v[bitrev(k, LP)], w[j]);
				}
			}
		}
	};
	static_assert(!TW_TMEM || 4 * P <= TMEM_PER_LEVEL, "tensor-memory columns per thread");

	// ---- MODE_FUSED1*: the load phase of a 2N-point frame, staged by TMA ----
	// A frame of 2N points needs, for every position m = n1*M + t of this thread, the samples and window
	// values at m and N + m.  Round 1 fetched them with 4*R1 LDG per thread and frame and stalled on them
	// (long_scoreboard 44 % of the stall samples).  Measured since (B200, round 2): with every load of the
	// frame served by L2 the kernel sits on the L2 -> SM throughput, about 4.9 TB/s over the chip,
	// whatever instruction fetches the bytes: LDG and cp.async.bulk give the same 7.6 M frames/s for the
	// 640 KB a frame moved (samples 128, window 128, parked class 128 out + 128 back, row 128).  So the
	// bytes had to go: the Nuttall table is symmetric (checked on the host), w[N + m] = w[N - 1 - m],
	// and a burst that takes positions from BOTH ends of the range needs each window run once:
	// burst b = positions b*HP .. b*HP + HP-1 and R1 - (b+1)*HP .. R1 - b*HP - 1; its stage holds four
	// sample runs (q = 0, 1 x low, high) and two window runs (low, high), six cp.async.bulk copies that
	// complete on the stage's `full` mbarrier.  Every warp reads its 32 columns with LDS (the q = 1 window
	// value of a low position is element M-1-t of the mirrored segment in the high run, and vice versa)
	// and releases the stage on its `empty` mbarrier (one arrival per warp).  The ring has ND stages of
	// memory of their own and NS that live in the 64 KB magnitude stash, which is idle during the load
	// phase (written in pass 3 of the even class, read in pass 3 of the odd class; the last readers of a
	// frame arrive on `stashfree`): 1 + 2 stages of 24 KB while the window runs travelled with the bursts,
	// 2 + 4 of 16 KB since the window waits in tensor memory (WIN_TMEM).  Burst b uses stage b % NST.
	// Schedule, no shared producer state: bursts 0 .. ND-1 of the next frame are issued right after this
	// frame's load phase (a whole frame ahead), bursts ND .. NST-1 at the start of the load phase, burst
	// b + NST - 1 when warp b % 16 starts burst b (into the stage burst b - 1 has just left); the
	// empty-barrier parity to wait for is the complement of the consumer's own phase bit of that stage
	// (its last use has been consumed).
	static_assert(!FUSED || FLA == 1 || (FLA == 2 && FUSED_TMEM), "the fused kernels do the radix-2 step, or the radix-4 step with tensor memory");
	constexpr int FNA = 1 << FLA;
	constexpr int RAWB = FIN32 ? 8 : 4;
	constexpr int HP = (FIN32 || FLA == 2) ? 1 : 2;                // positions per burst from each end
	constexpr unsigned int RUN_RAW = (unsigned int)(HP * M) * RAWB, RUN_WIN = (unsigned int)(HP * M) * 4u;
	// bytes per stage: four sample runs, and unless the window waits in tensor memory two window runs:
	// 16 KB (24 KB with window) for int16, 16 KB (20 KB) for float32
	// (Na = 4: eight sample runs and four window runs -- q = 0, 1 at either end; q = 2, 3 are their mirror images --
	// of one position each: 24 KB)
	// Na = 4 with WIN_LDG: the window values come by LDG one burst ahead instead (no mirror trick, four loads per
	// position), so that a stage is 16 KB of samples and the ring has 2 + 4 stages like the Na = 2 one.
	// -DRO_FUSED2_WIN_LDG=0: window runs in the bursts, 1 + 2 stages of 24 KB (A/B).
#ifndef RO_FUSED2_WIN_LDG
#define RO_FUSED2_WIN_LDG 1
#endif
	constexpr bool WIN_LDG = RO_FUSED2_WIN_LDG && FLA == 2;
	constexpr bool WIN_RUNS = !WIN_TMEM && !WIN_LDG;               // window runs travel with the bursts
	constexpr unsigned int SB = 2u * FNA * RUN_RAW + (WIN_RUNS ? (unsigned int)FNA * RUN_WIN : 0u);
	// ring: ND stages of memory of their own (filled a whole frame ahead), NS more in the magnitude stash (filled at
	// the start of the load phase); 1 + 2 with the window runs in the bursts, 2 + 4 without.  -DRO_FUSED_RING3=1: 1 + 2.
#ifndef RO_FUSED_RING3
#define RO_FUSED_RING3 0
#endif
	constexpr int ND = RO_FUSED_RING3 ? 1 : (int)(FUSED_RING_BYTES / SB), NS = RO_FUSED_RING3 ? 2 : (int)(4u * N / SB);
	constexpr int NST = ND + NS;
	static_assert(!FUSED || (NST >= 3 && NST <= 8 && (2 * NST + 1) * 8 <= (int)FUSED_BAR_BYTES), "ring stages");
	constexpr int NB = R1 / (2 * HP);                              // bursts per frame
	static_assert(!FUSED || (SB <= FUSED_RING_BYTES && 2 * SB <= 4u * N), "ring");
	unsigned char *stash8 = reinterpret_cast<unsigned char *>(mbar + 4);
	unsigned char *ded8 = stash8 + 4 * N;
	unsigned long long *fullb = reinterpret_cast<unsigned long long *>(ded8 + FUSED_RING_BYTES);
	unsigned long long *emptyb = fullb + NST, *stashfree = emptyb + NST;
	unsigned int fullph = 0, stashph = 1;       // a fresh barrier passes a wait on parity 1
	auto stage_ptr = [&](int s) -> unsigned char * { return s < ND ? ded8 + s * SB : stash8 + (s - ND) * SB; };
	// all lanes of one warp: fill the stage of burst b of (batch-relative) frame fr
	auto issue_burst = [&](long long fr, int b) {
		if constexpr (FUSED) {
			const int lane = t & 31;
			const int s = b % NST;
			mbar_wait(emptyb + s, ((fullph >> s) & 1u) ^ 1u);
			if (lane == 0) mbar_expect_tx(fullb + s, SB);
			__syncwarp();
			if (lane < (WIN_RUNS ? 3 * FNA : 2 * FNA)) {
				const int pos0 = (lane & 1) ? R1 - (b + 1) * HP : b * HP;      // odd lanes: the high end
				unsigned char *dst = stage_ptr(s);
				if (lane < 2 * FNA)      // sample run q = lane / 2
					// (Na = 4 re-reads three quarters of a frame's samples 90 us later, by which time L2 has let them go:
					// ncu 302 KB from HBM per frame for 64 KB of new samples.  Loading them with an evict-last policy
					// measured 2 % SLOWER, profiles/r02_fused64k_ab.txt.)
					bulk_g2s(dst + lane * RUN_RAW,
					         reinterpret_cast<const unsigned char *>(p.raw) + (fr * hop + (long long)(lane >> 1) * N + pos0 * M) * RAWB,
					         RUN_RAW, fullb + s);
				else                     // window run q = (lane - 2*Na) / 2 of the first half of the table
					bulk_g2s(dst + 2 * FNA * RUN_RAW + (lane - 2 * FNA) * RUN_WIN,
					         p.window_full + (long long)((lane - 2 * FNA) >> 1) * N + pos0 * M, RUN_WIN, fullb + s);
			}
			__syncwarp();
		}
	};
	if constexpr (FUSED) {
		if (t == 0) {
			for (int s = 0; s < NST; s++) { mbar_init(fullb + s, 1); mbar_init(emptyb + s, T / 32); }
			mbar_init(stashfree, T / 32);
		}
		__syncthreads();
		if (t < 32) {
#pragma unroll
			for (int b = 0; b < ND && b < NB; b++) issue_burst(f0 >> FLA, b);
		}
	}

#pragma unroll 1
	for (long long f = f0; f < f1; f++) {
		c64 v[P];
		// ---- load (shared-memory ring or global) + window ----
		if (STAGE) {
			const int base = (int)(((f - f0) * hop) & NMASK);
			RO_TP(0);
			mbar_wait(mbar, parity);
			parity ^= 1u;
			RO_TP(1);
			// all loads first, then the conversions, fenced on both sides (see ptxas_fence)
			unsigned int raw[R1];
			float wv[R1];
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				raw[n1] = ring[(base + n1 * M + t) & NMASK];
				wv[n1] = (n1 < R1 / 2) ? win[n1 * M + t] : win[(R1 - 1 - n1) * M + (M - 1 - t)];
			}
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				const c64 x = short2_to_c64(raw[n1]);
				v[n1] = mul2(GAIN ? add2(x, gain2) : x, pk(wv[n1], wv[n1]));
			}
		} else if (FUSED) {
			// Frame F = f >> FLA of Na*N points, residue class r = f & (Na-1):
			//     y_r[m] = W_{Na*N}^{m*r} * sum_q xw[q*N + m] * W_Na^{q*r},   X[Na*k + r] = DFT_N(y_r)[k]
			// Pass 1 of every class needs y_r at m = n1*M + t, n1 = 0..R1-1: the SAME m for every r.  So at
			// r = 0 this thread loads its Na*R1 samples once, does R1 Na-point transforms, keeps y_0 as its
			// pass-1 input and parks y_1 .. y_{Na-1} in a slice of global memory that only this CTA touches
			// (Na-1 x 128 KB, rewritten every frame: it lives in L2); at r > 0 it reads back what it wrote
			// itself, so no barrier is involved.  wf_split_kernel + MODE_SPLIT did the same arithmetic with a
			// frame-sized scratch for the whole batch, i.e. a round trip of 16 B per point through HBM.
			if constexpr (FUSED) {
			constexpr int NA = FNA;
			constexpr int NBIG = N << FLA;
			const int r = (int)(f & (NA - 1));
			RO_TP(0);
			// Na = 2: one class parked (tensor memory, or (Na-1) slices of global memory without it); Na = 4: classes
			// 1 and 2 in the two halves of the thread's tensor memory, class 3 in ONE slice of global memory
			unsigned long long *scr = p.scratch + (size_t)blockIdx.x * (FLA == 2 ? 1 : NA - 1) * N;
			if (r == 0) {
				typedef typename raw_word<FIN32>::type raw_t;
				const raw_t *src = reinterpret_cast<const raw_t *>(p.raw) + (f >> FLA) * hop;
				// The twiddle W_{Na*N}^(m*r), m = n1*M + t, is W_{Na*N}^(t*r) (loaded once per frame) times the
				// compile-time constant W_{Na*N/M}^(n1*r) = W_64^n1 (Na = 2), W_128^(n1*r) (Na = 4).
				const unsigned long long *twb = reinterpret_cast<const unsigned long long *>(p.tw_big);
				const c64 wt1 = __ldg(twb + t);
				c64 wt2 = wt1, wt3 = wt1;
				if (FLA == 2) { wt2 = __ldg(twb + 2 * t); wt3 = __ldg(twb + 3 * t); }
				float wnext[2 * HP][NA];               // WIN_LDG: the window values of the next burst
				if constexpr (WIN_LDG) {
#pragma unroll
					for (int jj = 0; jj < 2 * HP; jj++)
#pragma unroll
						for (int q = 0; q < NA; q++)
							wnext[jj][q] = __ldg(p.window_full + q * N + (jj < HP ? jj : R1 - HP + (jj - HP)) * M + t);
				}
#pragma unroll
				for (int b = 0; b < NB; b++) {
					raw_t raw[2 * HP][NA];
					float wv[2 * HP][NA];
					if constexpr (WIN_LDG) {
#pragma unroll
						for (int jj = 0; jj < 2 * HP; jj++)
#pragma unroll
							for (int q = 0; q < NA; q++) wv[jj][q] = wnext[jj][q];
					}
					if (b == 0) {
						// the stash of the previous frame has been read by every warp: bursts ND .. NST-1 may land in it
						if (t < 32) {
							mbar_wait(stashfree, stashph);
#pragma unroll
							for (int b2 = ND; b2 < NST && b2 < NB; b2++) issue_burst(f >> FLA, b2);
						}
						stashph ^= 1u;
					} else if (b + NST - 1 < NB && (t >> 5) == b % (T / 32)) {
						issue_burst(f >> FLA, b + NST - 1);      // into the stage burst b - 1 has just left
					}
					ptxas_fence(p.stagger & 1);
					const int s = b % NST;
					mbar_wait(fullb + s, (fullph >> s) & 1u);
					fullph ^= 1u << s;
					{
						const unsigned char *st = stage_ptr(s);
						const raw_t *rr = reinterpret_cast<const raw_t *>(st);
						const float *wl = reinterpret_cast<const float *>(st + 2 * NA * RUN_RAW), *wh = wl + HP * M;
						float wlo[2 * HP], whi[2 * HP];         // WIN_TMEM: [position][q] of the low and the high run
						if constexpr (WIN_TMEM) {
							tmem_ld_f(tmem + 64u + (unsigned int)(2 * b * HP), wlo);
							tmem_ld_f(tmem + 64u + (unsigned int)(2 * (R1 - (b + 1) * HP)), whi);
						}
#pragma unroll
						for (int j = 0; j < HP; j++) {
							// j: low position b*HP + j;  HP + j: high position R1 - (b+1)*HP + j
#pragma unroll
							for (int q = 0; q < NA; q++) {
								raw[j][q] = rr[(2 * q) * HP * M + j * M + t];
								raw[HP + j][q] = rr[(2 * q + 1) * HP * M + j * M + t];
							}
							if constexpr (WIN_RUNS) {
								// window runs q = 0 .. Na/2-1 at either end; q >= Na/2 is the mirror image of q' = Na-1-q at
								// the OTHER end: w[q*N + m] = w[q'*N + (N-1-m)]
#pragma unroll
								for (int q = 0; q < NA / 2; q++) {
									const float *wlq = wl + 2 * q * HP * M, *whq = wlq + HP * M;
									wv[j][q] = wlq[j * M + t];
									wv[HP + j][q] = whq[j * M + t];
									wv[j][NA - 1 - q] = whq[(HP - 1 - j) * M + (M - 1 - t)];
									wv[HP + j][NA - 1 - q] = wlq[(HP - 1 - j) * M + (M - 1 - t)];
								}
							}
						}
						if constexpr (WIN_TMEM) {
							tmem_wait_ld();
#pragma unroll
							for (int j = 0; j < HP; j++) {
								wv[j][0] = wlo[2 * j]; wv[j][1] = wlo[2 * j + 1];
								wv[HP + j][0] = whi[2 * j]; wv[HP + j][1] = whi[2 * j + 1];
							}
						}
						if constexpr (WIN_LDG) {
							if (b + 1 < NB) {
#pragma unroll
								for (int jj = 0; jj < 2 * HP; jj++)
#pragma unroll
									for (int q = 0; q < NA; q++)
										wnext[jj][q] = __ldg(p.window_full + q * N + (jj < HP ? (b + 1) * HP + jj : R1 - (b + 2) * HP + (jj - HP)) * M + t);
							}
						}
					}
					ptxas_fence(p.stagger & 1);
					c64 y1[2][HP], y2[2][HP];          // classes 1 (and 2) at the low and the high positions of this burst
#pragma unroll
					for (int jj = 0; jj < 2 * HP; jj++) {
						const int n1 = jj < HP ? b * HP + jj : R1 - (b + 1) * HP + (jj - HP);
						const int m = n1 * M + t;
						c64 xq[NA];
#pragma unroll
						for (int q = 0; q < NA; q++)
							xq[q] = mul2(add2(raw_to_c64(raw[jj][q]), gain2), pk(wv[jj][q], wv[jj][q]));
						fft_dif2<NA>(xq);                     // class rr is left at xq[bitrev(rr)]
						v[n1] = xq[0];
						if constexpr (FLA == 2) {
							y1[jj / HP][jj % HP] = cmul2(cmul2(xq[2], wt1), cw128(n1), -sw128(n1));
							y2[jj / HP][jj % HP] = cmul2(cmul2(xq[1], wt2), cw128(2 * n1), -sw128(2 * n1));
							scr[m] = cmul2(cmul2(xq[3], wt3), cw128(3 * n1), -sw128(3 * n1));
						} else {
							xq[1] = cmul2(cmul2(xq[1], wt1), cw64(n1), -sw64(n1));
							if (FUSED_TMEM) y1[jj / HP][jj % HP] = xq[1];
							else scr[m] = xq[1];
						}
					}
					if constexpr (FUSED_TMEM) {
						// columns 2*n1, 2*n1 + 1 of this thread's 128 (Na = 4: class 2 in the upper 64)
						tmem_st_c64(tmem + (unsigned int)(2 * b * HP), y1[0]);
						tmem_st_c64(tmem + (unsigned int)(2 * (R1 - (b + 1) * HP)), y1[1]);
						if constexpr (FLA == 2) {
							tmem_st_c64(tmem + 64u + (unsigned int)(2 * b * HP), y2[0]);
							tmem_st_c64(tmem + 64u + (unsigned int)(2 * (R1 - (b + 1) * HP)), y2[1]);
						}
					}
					// the values are in registers and used: the stage may be refilled
					__syncwarp();
					if ((t & 31) == 0) mbar_arrive(emptyb + s);
					if (b == NB / 2 - 1) RO_TP(18);
					if (b == NB - 1) RO_TP(19);
				}
#if RO_L2_PREFETCH
				if (f + NA < f1) {     // the hop new samples of the next frame
					const unsigned char *nx = reinterpret_cast<const unsigned char *>(src + NBIG);
					for (long long o = (long long)t * 128; o < (long long)sizeof(raw_t) * hop; o += (long long)T * 128)
						asm volatile("prefetch.global.L2 [%0];" :: "l"(nx + o));
				}
#endif
			} else if (FUSED_TMEM && r < 3) {
				tmem_wait_st();
				const unsigned int half = r == 2 ? 64u : 0u;
				tmem_ld_c64x16(tmem + half, &v[0]);
				tmem_ld_c64x16(tmem + half + 32u, &v[16]);
				tmem_wait_ld();
			} else {
				const unsigned long long *src = scr + (FLA == 2 ? 0 : r - 1) * N;
				ptxas_fence(p.stagger & 1);
#pragma unroll
				for (int n1 = 0; n1 < R1; n1++) v[n1] = __ldcg(src + n1 * M + t);
				ptxas_fence(p.stagger & 1);
			}
			}
		} else if (SPLIT) {
			const unsigned long long *src = reinterpret_cast<const unsigned long long *>(p.iq) + f * N;
			RO_TP(0);
			if constexpr (prefetch_rows) {
				// The first LROWS rows of this item were fetched by cp.async.bulk during the previous item's passes 2
				// and 3; the rest comes by LDG as before, issued first so that its latency passes behind the LDS.
				const c64 *land = reinterpret_cast<const c64 *>(mbar + 4);
				ptxas_fence(p.stagger & 1);
#pragma unroll
				for (int n1 = LROWS; n1 < R1; n1++) v[n1] = __ldg(src + n1 * M + t);
				mbar_wait(mbar, parity);
				parity ^= 1u;
#pragma unroll
				for (int n1 = 0; n1 < LROWS; n1++) v[n1] = land[n1 * M + t];
				ptxas_fence(p.stagger & 1);
				RO_TP(1);
			} else {
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) v[n1] = __ldg(src + n1 * M + t);
			ptxas_fence(p.stagger & 1);
			}
			// (prefetching the next item of the scratch into L2 here measured 12 % SLOWER, and bringing it into S
			// with cp.async.bulk while the previous item is in pass 3 -- S is idle then, and thread t reads and writes
			// the same column in pass 1 -- 10 % slower at 64k, 8 % at 128k: profiles/r02_split_tma_tmem_ab.txt.
			// This path is bound by the bytes it moves through HBM, not by the latency of getting them.)
		} else if (F32) {
			// float32 pairs are packed complex already: straight into the transform's registers
			const unsigned long long *src = reinterpret_cast<const unsigned long long *>(p.iq) + f * hop;
#if RO_L2_PREFETCH
			// the next frame's new samples: ask L2 for them now (one 128-byte line per thread and 4 KB of hop)
			if (f + 1 < f1) {
				const unsigned char *nx = reinterpret_cast<const unsigned char *>(src + N);
				for (long long o = (long long)t * 128; o < 8 * hop; o += (long long)T * 128)
					asm volatile("prefetch.global.L2 [%0];" :: "l"(nx + o));
			}
#endif
			float wv[R1];
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				v[n1] = __ldg(src + n1 * M + t);
				wv[n1] = (n1 < R1 / 2) ? win[n1 * M + t] : win[(R1 - 1 - n1) * M + (M - 1 - t)];
			}
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) v[n1] = mul2(GAIN ? add2(v[n1], gain2) : v[n1], pk(wv[n1], wv[n1]));
		} else {
			const unsigned int *src = reinterpret_cast<const unsigned int *>(p.iq) + f * hop;
#if RO_L2_PREFETCH
			if (f + 1 < f1) {
				const unsigned char *nx = reinterpret_cast<const unsigned char *>(src + N);
				for (long long o = (long long)t * 128; o < 4 * hop; o += (long long)T * 128)
					asm volatile("prefetch.global.L2 [%0];" :: "l"(nx + o));
			}
#endif
			unsigned int raw[R1];
			float wv[R1];
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				raw[n1] = __ldg(src + n1 * M + t);
				wv[n1] = (n1 < R1 / 2) ? win[n1 * M + t] : win[(R1 - 1 - n1) * M + (M - 1 - t)];
			}
			ptxas_fence(p.stagger & 1);
#pragma unroll
			for (int n1 = 0; n1 < R1; n1++) {
				const c64 x = short2_to_c64(raw[n1]);
				v[n1] = mul2(GAIN ? add2(x, gain2) : x, pk(wv[n1], wv[n1]));
			}
		}

		// ---- pass 1 ----
		RO_TP(2);
		// The fused 64k kernel has no registers to spare for ten twiddle bases that live as long as the CTA (ptxas
		// spilled 312 bytes per thread, and pass 1 of class 0 spent 5 k clk behind the parked class's stores waiting
		// for spilled bases, profiles/r02_trace_fused64k_timeline.txt): it fetches the five of a pass again, one
		// transform ahead of their use (36 bytes of spills left; 3.17 -> 3.29 M frames/s at 75 % overlap).  The same
		// in the other long-frame kernels: fused 32k -13 %, split path +-2 %, so not there.  -DRO_RELOAD_BASES=0: A/B.
#ifndef RO_RELOAD_BASES
#define RO_RELOAD_BASES 1
#endif
		constexpr bool RELOAD_B = RO_RELOAD_BASES && FLA == 2;
		constexpr bool TMEM_B = RO_SPLIT_BASES_TMEM && SPLIT_TMEM && LP == 5;
		c64 bb[LP];
		if constexpr (TMEM_B) {
			float q8[8], q2[2];
			tmem_ld_f(tmem + 56u, q8);
			tmem_ld_f(tmem + 88u, q2);
			tmem_wait_ld();
#pragma unroll
			for (int i = 0; i < 4; i++) bb[i] = pk(q8[2 * i], q8[2 * i + 1]);
			bb[4] = pk(q2[0], q2[1]);
		}
		if constexpr (RELOAD_B) {
			const unsigned long long *tw64 = reinterpret_cast<const unsigned long long *>(p.tw);
#pragma unroll
			for (int i = 0; i < LP; i++) bb[i] = __ldg(tw64 + ((t << i) & NMASK));
		}
		fft_fwd2<R1, USE_DIT>(v);
		RO_TP(17);
		RO_SITE(0, p.stagger & 1);
		if (TW_TMEM) twiddle_tmem(v, 0u);
		else if (RELOAD_B || TMEM_B) twiddle_powers2<R1>(v, bb);
		else twiddle_powers2<R1>(v, b1);
		RO_SITE(1, p.stagger & 1);
		RO_TP(3);
		// S still holds the previous frame's pass-3 inputs until every thread has read them
		if (f > f0) mbar_wait(bar_rd3, (unsigned int)(f - 1 - f0) & 1u);
		RO_TP(4);
#pragma unroll
		for (int k1 = 0; k1 < R1; k1++) S[k1 * ROW1 + ((SWZ1 && (k1 & 1)) ? (t ^ 8) : t)] = v[bitrev(k1, LP)];
		RO_TP(5);
		__syncthreads();
		RO_TP(6);
		if constexpr (FUSED) {
			// every warp is through the load phase of this frame: the ring's own stages can take the first
			// steps of the next frame now, a whole frame ahead of their use
			if (t < 32 && (f & (FNA - 1)) == 0 && f + FNA < f1) {
#pragma unroll
				for (int b = 0; b < ND && b < NB; b++) issue_burst((f >> FLA) + 1, b);
			}
		}

		// MODE_SPLIT: every thread has read the landing area: the next item's rows may come
		if (LROWS > 0 && prefetch_rows && t == 0 && f + 1 < f1) fetch_rows(f + 1);

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
		// Row k1 of S (R2*R3 points) is read and transformed by the R3 threads (k1, n3) alone, all in one
		// warp; the outputs go to the transposed, XOR-swizzled layout pass 3 reads with 128-bit loads.
		{
			const int k1 = t / R3, n3 = t % R3;
			if (CHAIN && (p.stagger & 2)) chain_wait(wq, wl, cdist);
			RO_TP(7);
#pragma unroll
			for (int n2 = 0; n2 < R2; n2++) {
				// SWZ1: (n2*8 + n3) ^ (8*(k1 & 1)) = n2*8 + n3 +- 8*(k1 & 1), the sign by the parity of n2
				const int sw = SWZ1 ? ((k1 & 1) << 3) : 0;
				v[n2] = S[k1 * ROW1 + n2 * R3 + n3 + ((n2 & 1) ? -sw : sw)];
			}
			if (CHAIN && (p.stagger & 2)) chain_pass(wq, wl, LEVELS, cdist);
			split_arrive(bar_rd2, t);                   // my pass-2 inputs are in registers
			RO_TP(8);
			if constexpr (TMEM_B) {
				float q8[8], q2[2];
				tmem_ld_f(tmem + 120u, q8);
				tmem_ld_f(tmem + 90u, q2);
				tmem_wait_ld();
#pragma unroll
				for (int i = 0; i < 4; i++) bb[i] = pk(q8[2 * i], q8[2 * i + 1]);
				bb[4] = pk(q2[0], q2[1]);
			}
			if constexpr (RELOAD_B && !B2_SMEM) {
				const unsigned long long *tw64 = reinterpret_cast<const unsigned long long *>(p.tw);
#pragma unroll
				for (int i = 0; i < LP; i++) bb[i] = __ldg(tw64 + (((n3 * P) << i) & NMASK));
			}
			if constexpr (B2_SMEM) {
#pragma unroll
				for (int i = 0; i < LP; i++) bb[i] = b2tab[n3 * LP + i];
			}
			fft_fwd2<R2, USE_DIT>(v);
			RO_SITE(2, p.stagger & 1);
			if (TW_TMEM) twiddle_tmem(v, 2u * P);
			else if (RELOAD_B || TMEM_B || B2_SMEM) twiddle_powers2<R2>(v, bb);
			else twiddle_powers2<R2>(v, b2);
			RO_SITE(3, p.stagger & 1);
			RO_TP(9);
			mbar_wait(bar_rd2, (unsigned int)(f - f0) & 1u);   // ... and so are everyone else's
			RO_TP(10);
			{
				const int col = k1 ^ (n3 * SW);
#pragma unroll
				for (int k2 = 0; k2 < R2; k2++) S[(k2 * R3 + n3) * R1 + col] = v[bitrev(k2, LP)];
			}
			RO_TP(11);
		}
		__syncthreads();
		RO_TP(12);

		// ---- pass 3 + magnitude + fftshift store ----
		constexpr int G = P / R3;                  // 16-point (R3) transforms per thread
		if (G >= 2) {
			// thread owns G ADJACENT outputs pidx = G*t + c: their pass-3 inputs are G adjacent
			// (XOR-permuted) slots -> 128-bit shared loads; their magnitudes G adjacent floats
			// -> one 64/128-bit global store per k3
			const int pbase = G * t;
			const int k1b = pbase % R1, k2 = pbase / R1;
			c64 (&u)[P] = v;                       // u[c*R3 + n3]
			if (CHAIN && (p.stagger & 4)) chain_wait(wq, wl, cdist3);
			RO_TP(13);
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
			if (CHAIN && (p.stagger & 4)) chain_pass(wq, wl, LEVELS, cdist3);
			split_arrive(bar_rd3, t);                  // my pass-3 inputs are in registers
			RO_TP(14);
			float mag[R3][G];
#pragma unroll
			for (int c = 0; c < G; c++) {
				c64 w[R3];
#pragma unroll
				for (int n3 = 0; n3 < R3; n3++) w[n3] = u[c * R3 + n3];
				fft_fwd2<R3, USE_DIT>(w);
#pragma unroll
				for (int k3 = 0; k3 < R3; k3++) {
					float zr, zi;
					upk(w[bitrev(k3, L3)], zr, zi);
					mag[k3][c] = fast_sqrt(fmaf(zr, zr, zi * zi));
				}
				RO_SITE(4, p.stagger & 1);
			}
			RO_SITE(5, p.stagger & 1);
			RO_TP(15);
			if (SPLIT) {
				// output k of residue class r is bin Na*k + r of the Na*N-point frame
				const int la = p.log2na;
				const int r = (int)(f & ((1 << la) - 1));
				float *row = p.rows + (f >> la) * p.row_stride;
				const int big_mask = (N << la) - 1, big_half = (N << la) >> 1;
				if constexpr (FLA == 2) {
					// MODE_FUSED2: the four classes of a frame are stored together, 32-byte sectors whole (see the split
					// path below).  Class 0 waits in the shared-memory stash, classes 1 and 2 in the lower half of the
					// thread's tensor memory (columns 32*(r-1) + 2*k3 + c), which their own parked inputs have left.
					float *stash = reinterpret_cast<float *>(mbar + 4);
					if (r == 0) {
#pragma unroll
						for (int k3 = 0; k3 < R3; k3++)
#pragma unroll
							for (int c = 0; c < G; c++) stash[(k3 * G + c) * T + t] = mag[k3][c];
					} else if (r < 3) {
#pragma unroll
						for (int g = 0; g < R3 / 4; g++) {
							float pv[8];
#pragma unroll
							for (int i = 0; i < 4; i++) { pv[2 * i] = mag[4 * g + i][0]; pv[2 * i + 1] = mag[4 * g + i][1 % G]; }
							tmem_st8(tmem + (unsigned int)(32 * (r - 1) + 8 * g), pv);
						}
						tmem_wait_st();
					} else {
						const int odd = t & 1;
#pragma unroll
						for (int g = 0; g < R3 / 4; g++) {
							float m1[8], m2[8];
							tmem_ld_f(tmem + (unsigned int)(8 * g), m1);
							tmem_ld_f(tmem + 32u + (unsigned int)(8 * g), m2);
							tmem_wait_ld();
#pragma unroll
							for (int i = 0; i < 4; i++) {
								const int k3 = 4 * g + i;
								const float4 a0 = make_float4(stash[(k3 * G) * T + t], m1[2 * i], m2[2 * i], mag[k3][0]);
								const float4 a1 = make_float4(stash[(k3 * G + 1 % G) * T + t], m1[2 * i + 1], m2[2 * i + 1], mag[k3][1 % G]);
								const float4 snd = odd ? a0 : a1;
								float4 rcv;
								rcv.x = __shfl_xor_sync(0xffffffffu, snd.x, 1);
								rcv.y = __shfl_xor_sync(0xffffffffu, snd.y, 1);
								rcv.z = __shfl_xor_sync(0xffffffffu, snd.z, 1);
								rcv.w = __shfl_xor_sync(0xffffffffu, snd.w, 1);
								const int blk = (k3 * (N / R3)) << 2;          // first bin of the k3 block
								float4 *lo = reinterpret_cast<float4 *>(row + ((blk + 8 * (t & ~1) + big_half) & big_mask) + 4 * odd);
								float4 *hi = reinterpret_cast<float4 *>(row + ((blk + 8 * (t | 1) + big_half) & big_mask) + 4 * odd);
								__stcs(lo, odd ? rcv : a0);
								__stcs(hi, odd ? a1 : rcv);
							}
						}
					}
				} else if (SPLIT_TMEM && p.pair == 2 && la >= 2) {
					// Na >= 4.  With pairs a thread wrote 8 bytes of every 4*Na: at Na = 4 half of a 32-byte sector
					// now and the other half two items later, and every sector of the row made two partial trips
					// to HBM (ncu, 64k: 243 KB read back and 121 KB written twice per 256 KB row).  Classes r = 4j ..
					// 4j+2 now park their magnitudes in tensor memory (columns [k3/4][class][k3%4][c]) and class 4j+3
					// stores all four: [c0 c1 c2 c3](k), 16 bytes, and at Na = 4 the thread's two outputs k = 2t,
					// 2t+1 are one whole sector, which lane pairs exchange halves of so that every store
					// instruction covers whole sectors.  The host hands every CTA a multiple of four items.
					const int q = r & 3;
					if (q != 3) {
#pragma unroll
						for (int g = 0; g < R3 / 4; g++) {
							float pv[8];
#pragma unroll
							for (int i = 0; i < 4; i++) { pv[2 * i] = mag[4 * g + i][0]; pv[2 * i + 1] = mag[4 * g + i][1 % G]; }
							tmem_st8(tmem + (unsigned int)(g * 32 + q * 8), pv);
						}
						asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
					} else {
						const int odd = t & 1;
#pragma unroll
						for (int g = 0; g < R3 / 4; g++) {
							float pv[32];
							tmem_ld32(tmem + (unsigned int)(g * 32), pv);
#pragma unroll
							for (int i = 0; i < 4; i++) {
								const int k3 = 4 * g + i;
								const float4 a0 = make_float4(pv[2 * i], pv[8 + 2 * i], pv[16 + 2 * i], mag[k3][0]);
								const float4 a1 = make_float4(pv[2 * i + 1], pv[8 + 2 * i + 1], pv[16 + 2 * i + 1], mag[k3][1 % G]);
								const int k = k3 * (N / R3) + G * t;          // first of this thread's two outputs
								if (la == 2) {
									// sector of this thread: floats [8t, 8t + 8) of the k3 block = a0 | a1
									const float4 snd = odd ? a0 : a1;
									float4 rcv;
									rcv.x = __shfl_xor_sync(0xffffffffu, snd.x, 1);
									rcv.y = __shfl_xor_sync(0xffffffffu, snd.y, 1);
									rcv.z = __shfl_xor_sync(0xffffffffu, snd.z, 1);
									rcv.w = __shfl_xor_sync(0xffffffffu, snd.w, 1);
									const int blk = (k3 * (N / R3)) << la;        // first bin of the k3 block
									float *lo = row + ((blk + 8 * (t & ~1) + big_half) & big_mask) + 4 * odd;   // the even lane's sector
									float *hi = row + ((blk + 8 * (t | 1) + big_half) & big_mask) + 4 * odd;    // the odd lane's sector
									*reinterpret_cast<float4 *>(lo) = odd ? rcv : a0;
									*reinterpret_cast<float4 *>(hi) = odd ? a1 : rcv;
								} else {
									*reinterpret_cast<float4 *>(row + (((k << la) + r - 3 + big_half) & big_mask)) = a0;
									*reinterpret_cast<float4 *>(row + ((((k + 1) << la) + r - 3 + big_half) & big_mask)) = a1;
								}
							}
						}
					}
				} else if (p.pair) {
					// Alone, a class writes 4 bytes of every 4*Na: 32 scattered stores per thread and a
					// queue of partial sectors (lg_throttle was the top stall).  Instead an even class
					// parks its magnitudes in a thread-private slice of shared memory, and the odd class
					// that the same thread transforms next stores both: bins Na*k + r - 1, Na*k + r.
					float *stash = reinterpret_cast<float *>(mbar + 4);
					float evt[SPLIT_TMEM ? 32 : 1];            // SPLIT_TMEM: the even class waits in tensor memory, [k3][c]
					if ((r & 1) == 0) {
						if constexpr (SPLIT_TMEM) {
#pragma unroll
							for (int g = 0; g < R3 / 4; g++) {
								float pv[8];
#pragma unroll
								for (int i = 0; i < 4; i++) { pv[2 * i] = mag[4 * g + i][0]; pv[2 * i + 1] = mag[4 * g + i][1 % G]; }
								tmem_st8(tmem + (unsigned int)(8 * g), pv);
							}
							tmem_wait_st();
						} else {
#pragma unroll
						for (int k3 = 0; k3 < R3; k3++)
#pragma unroll
							for (int c = 0; c < G; c++) stash[(k3 * G + c) * T + t] = mag[k3][c];
						}
					} else {
						if constexpr (SPLIT_TMEM) tmem_ld32(tmem, reinterpret_cast<float (&)[32]>(evt));
#pragma unroll
						for (int k3 = 0; k3 < R3; k3++) {
							float ev[G];
#pragma unroll
							for (int c = 0; c < G; c++) ev[c] = SPLIT_TMEM ? evt[(k3 * G + c) % (SPLIT_TMEM ? 32 : 1)] : stash[(k3 * G + c) * T + t];
							const int k = k3 * (N / R3) + G * t;
							if (G == 2 && la == 1) {
								// fused 32k: evict-first stores (STG.E.EF), so that the rows do not push the overlapped input
								// (re-read by the next frame) and the parked classes out of L2.  Measured on B200, 75 %
								// overlap: int16 7.63 -> 7.81, float32 6.35 -> 6.46 M frames/s (-DRO_ROWS_STCS=0 for the A/B)
#ifndef RO_ROWS_STCS
#define RO_ROWS_STCS 1
#endif
								float4 *dst = reinterpret_cast<float4 *>(row + (((k << 1) + big_half) & big_mask));
								const float4 val = make_float4(ev[0], mag[k3][0], ev[1], mag[k3][1]);
								if (FUSED && RO_ROWS_STCS) __stcs(dst, val);
								else *dst = val;
							} else {
#pragma unroll
								for (int c = 0; c < G; c++)
									*reinterpret_cast<float2 *>(row + ((((k + c) << la) + r - 1 + big_half) & big_mask)) =
										make_float2(ev[c], mag[k3][c]);
							}
						}
					}
				} else {
#pragma unroll
				for (int k3 = 0; k3 < R3; k3++) {
#pragma unroll
					for (int c = 0; c < G; c++) {
						const int k = k3 * (N / R3) + G * t + c;
						row[((k << la) + r + big_half) & big_mask] = mag[k3][c];
					}
				}
				}
				if (MODE == MODE_SPLIT || MODE == MODE_FUSED2) RO_TP(16);
				if constexpr (FUSED) {
					// the last class of the frame has read the stash: its 64 KB may stage the next frame
					if (r == FNA - 1) {
						__syncwarp();
						if ((t & 31) == 0) mbar_arrive(stashfree);
					}
				}
			} else {
			float *row = p.rows + f * p.row_stride + G * t;
#pragma unroll
			for (int k3 = 0; k3 < R3; k3++) {
				// bins G*t + c + k3*(N/R3) go to row index (bin + N/2) mod N; G*t + c < N/R3, no wrap inside
				const int off = (k3 * (N / R3) + N / 2) & NMASK;
#ifndef RO_ROWS_STCS_ALL
#define RO_ROWS_STCS_ALL 0
#endif
				if (RO_ROWS_STCS_ALL) {
					if (G == 2) __stcs(reinterpret_cast<float2 *>(row + off), make_float2(mag[k3][0], mag[k3][1]));
					else __stcs(reinterpret_cast<float4 *>(row + off), make_float4(mag[k3][0], mag[k3][1], mag[k3][2 % G], mag[k3][3 % G]));
				} else
				if (G == 2) *reinterpret_cast<float2 *>(row + off) = make_float2(mag[k3][0], mag[k3][1]);
				else *reinterpret_cast<float4 *>(row + off) = make_float4(mag[k3][0], mag[k3][1], mag[k3][2 % G], mag[k3][3 % G]);
			}
			RO_TP(16);
			}
		} else {
			const int la = SPLIT ? p.log2na : 0;
			const int big_mask = (N << la) - 1, big_half = (N << la) >> 1;
			float *row = SPLIT ? p.rows + (f >> la) * p.row_stride + (int)(f & ((1 << la) - 1))
			                   : p.rows + f * p.row_stride + t;
			c64 u[R3];
#pragma unroll
			for (int n3 = 0; n3 < R3; n3++) u[n3] = S[((t / R1) * R3 + n3) * R1 + ((t % R1) ^ (n3 * SW))];
			split_arrive(bar_rd3, t);
			fft_fwd2<R3, USE_DIT>(u);
#pragma unroll
			for (int k3 = 0; k3 < R3; k3++) {
				const int off = (k3 * (N / R3) + N / 2) & NMASK;   // t < N/R3 = T
				float zr, zi;
				upk(u[bitrev(k3, L3)], zr, zi);
				if (SPLIT) row[(((k3 * (N / R3) + t) << la) + big_half) & big_mask] = fast_sqrt(fmaf(zr, zr, zi * zi));
				else row[off] = fast_sqrt(fmaf(zr, zr, zi * zi));
			}
		}
	}
	if constexpr (USE_TMEM) {
		asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
		__syncthreads();
		if (t < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(*reinterpret_cast<unsigned int *>(mbar + 3)), "n"(TMEM_COLS) : "memory");
	}
}

// ---------------------------------------------------------------------------
// Frames longer than the in-CTA transform (N = Na*Nb, Nb = 16384): the first
// radix-Na decimation-in-frequency step, done ONCE per frame for all Na residue
// classes (v1 redid it in every class's CTA):
//     y_r[m] = W_N^{m*r} * sum_q xw[q*Nb + m] * W_Na^{q*r},   X[Na*k + r] = DFT_Nb(y_r)[k]
// with xw = (sample + i*gain) * window  (src/FFTBackend.cpp:211-232).  One thread
// per m: Na coalesced sample loads, Na-point register FFT, Na coalesced 8-byte
// stores into the scratch y[frame][r][m] that wf_fft2_kernel<MODE_SPLIT> reads back
// (from L2: the host sizes a batch of frames to stay cache-resident).
// ---------------------------------------------------------------------------
struct SplitParams {
	const void   *iq;        // int16 or float32 interleaved I/Q
	const float  *window;    // N floats
	const float2 *tw;        // N entries exp(-2*pi*i*j/N)
	c64          *y;         // [n_frames][Na][Nb]
	long long     frame0;    // first frame of this batch
	long long     hop;
	int           log2n;
	float         iq_gain;
};

template <int LOG2NA, bool F32IN>
__global__ void __launch_bounds__(256)
wf_split_kernel(const SplitParams p)
{
	constexpr int NA = 1 << LOG2NA;
	const int N = 1 << p.log2n;
	const int NB = N >> LOG2NA;
	const int m = blockIdx.x * 256 + threadIdx.x;              // < NB (NB is a multiple of 256)
	const long long base = (p.frame0 + blockIdx.y) * p.hop;
	const c64 gain2 = pk(0.f, p.iq_gain);

	c64 v[NA];
#pragma unroll
	for (int q = 0; q < NA; q++) {
		const int s = q * NB + m;
		c64 x;
		if (F32IN) x = __ldg(reinterpret_cast<const unsigned long long *>(p.iq) + base + s);
		else x = short2_to_c64(__ldg(reinterpret_cast<const unsigned int *>(p.iq) + base + s));
		const float w = __ldg(p.window + s);
		v[q] = mul2(add2(x, gain2), pk(w, w));
	}
	fft_fwd2<NA, false>(v);
	c64 b[LOG2NA];
	const unsigned long long *tw64 = reinterpret_cast<const unsigned long long *>(p.tw);
#pragma unroll
	for (int i = 0; i < LOG2NA; i++) b[i] = __ldg(tw64 + ((m << i) & (N - 1)));
	twiddle_powers2<NA>(v, b);
	c64 *dst = p.y + ((long long)blockIdx.y * NA) * NB + m;
#pragma unroll
	for (int r = 0; r < NA; r++) dst[(long long)r * NB] = v[bitrev(r, LOG2NA)];
}

}  // namespace ro
