// ro_v2.h -- host-side entry to the v2 (packed f32x2, TMA-staged) waterfall kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace ro {

enum { V2_DIRECT = 0, V2_STAGED = 1, V2_SPLIT = 2, V2_FUSED1 = 3, V2_FUSED2 = 4, V2_DIRECT_F32 = 5, V2_FUSED1_F32 = 6 };   // = MODE_* of ro_kernels2.cuh

struct V2Kernel {
	const void *fn;      // kernel symbol (for attribute / occupancy queries)
	int         threads;
	size_t      smem;
	int         log2n;
	int         mode;
	bool        gain;
};

// Picks the instantiation for N = 2^log2n; returns false when v2 has none.
// mode V2_SPLIT exists for log2n = 14 only (the in-CTA part of longer frames).
bool v2_pick(int log2n, int mode, bool gain, V2Kernel *out);

// The same for float32 I/Q input (N = 1024 .. 16384), launched with v2_launch like V2_DIRECT (ro_v2f.cu).
bool v2_pick_f32(int log2n, bool gain, V2Kernel *out);
// V2_FUSED1 for float32 I/Q (frames of 2*16384 points), launched with v2_launch_fused.
bool v2_pick_fused_f32(V2Kernel *out);

// Columns of tensor memory one CTA of the kernel allocates for its whole life (0: none): at most 512 / this many
// CTAs may be resident on an SM, a CTA beyond that waits in tcgen05.alloc until another one exits.
int v2_tmem_cols(const V2Kernel &k);

// Launches it: n_frames items of `hop`-spaced N-sample windows of d_iq (int16 I/Q), or, for
// V2_SPLIT, n_frames = frames*Na items of N packed complex floats each (d_iq = the scratch
// written by v2_launch_split), whose outputs interleave into rows of Na*N floats; `pair` != 0
// promises 16-byte aligned rows and an even frames_per_cta (classes r, r+1 are stored together), `pair` == 2
// a frames_per_cta that is a multiple of four (Na >= 4: classes 4j .. 4j+3 are stored together).
cudaError_t v2_launch(const V2Kernel &k, int grid, cudaStream_t s, const void *d_iq, const float *d_window_half,
                      const float2 *d_tw, float *d_rows, long long row_stride, long long n_frames, long long hop,
                      long long frames_per_cta, float iq_gain, int log2na, int pair, int stagger);

// V2_FUSED1 (frames of 2*16384 points): one launch does the radix-2 step and the two 16k transforms of each
// frame; d_scratch holds grid * 16384 packed complex.  The raw samples and the first half of the window
// (which must be symmetric) reach shared memory by cp.async.bulk: d_raw and every frame start
// (hop * bytes per sample) must be 16-byte aligned, rows 16-byte aligned (pair = 1).
// V2_FUSED2 (frames of 4*16384 points, int16): log2na = 2, d_window_full / d_tw_big of 65536 entries.
cudaError_t v2_launch_fused(const V2Kernel &k, int grid, cudaStream_t s, const void *d_raw, const float *d_window_full,
                            const float2 *d_tw_big, const float2 *d_tw_nb, void *d_scratch, float *d_rows,
                            long long row_stride, long long n_frames, long long hop, long long frames_per_cta,
                            float iq_gain, int pair, int stagger, int log2na = 1);

// First radix-Na step of frames frame0 .. frame0+n_frames-1 (N = 2^log2n points, Na = 2^log2na)
// into d_y[n_frames][Na][N/Na] (packed complex float).  f32in: float32 I/Q instead of int16.
cudaError_t v2_launch_split(cudaStream_t s, int log2n, int log2na, bool f32in, const void *d_iq,
                            const float *d_window, const float2 *d_tw, void *d_y, long long frame0,
                            long long n_frames, long long hop, float iq_gain);

}  // namespace ro
