/*
 * host/shims/fftw3.h -- the five FFTW3 entry points the reference's FFTBackend links against
 * (src/FFTBackend.cpp:117-120 fftw_malloc / fftw_plan_dft_1d, :236 fftw_execute, :134-137
 * fftw_destroy_plan / fftw_free), for building radio-observer-cuda on a machine without libfftw3.
 *
 * The CUDA backend never executes a plan: FFTBackend's constructor creates one (it is the base
 * class) and that is all.  The stock "waterfall" factory stays usable in the same binary through
 * fftw3_min.cpp, a plain double-precision transform -- NOT FFTW, neither in speed nor in rounding.
 * host/Makefile uses the system's <fftw3.h> and -lfftw3 instead whenever they are installed.
 */
#ifndef RO_HOST_FFTW3_MIN_H
#define RO_HOST_FFTW3_MIN_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef double fftw_complex[2];
typedef struct ro_min_fft_plan *fftw_plan;

#define FFTW_FORWARD  (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE  (0U)
#define FFTW_ESTIMATE (1U << 6)

void     *fftw_malloc(size_t n);
void      fftw_free(void *p);
fftw_plan fftw_plan_dft_1d(int n, fftw_complex *in, fftw_complex *out, int sign, unsigned flags);
void      fftw_execute(const fftw_plan p);
void      fftw_destroy_plan(fftw_plan p);

#ifdef __cplusplus
}
#endif

#endif
