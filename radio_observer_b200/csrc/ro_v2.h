// ro_v2.h -- host-side entry to the v2 (packed f32x2, TMA-staged) waterfall kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace ro {

struct V2Kernel {
	const void *fn;      // kernel symbol (for attribute / occupancy queries)
	int         threads;
	size_t      smem;
	int         log2n;
	bool        stage;
	bool        gain;
};

// Picks the instantiation for N = 2^log2n; returns false when v2 has none.
bool v2_pick(int log2n, bool stage, bool gain, V2Kernel *out);

// Launches it: n_frames frames of `hop`-spaced N-sample windows of d_iq (int16 I/Q).
cudaError_t v2_launch(const V2Kernel &k, int grid, cudaStream_t s, const void *d_iq, const float *d_window_half,
                      const float2 *d_tw, float *d_rows, long long row_stride, long long n_frames, long long hop,
                      long long frames_per_cta, float iq_gain);

}  // namespace ro
