/*
 * oracle/shims/fftw3.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Declaration-only stand-in for the five FFTW3 entry points the reference
 * calls (src/FFTBackend.cpp:117-120 fftw_malloc/fftw_plan_dft_1d, :236
 * fftw_execute, :134-137 fftw_destroy_plan/fftw_free).  libfftw3 is not
 * installed in this image, so oracle/_ref links the
 * reference's own translation units against oracle/shims/fftw3_shim.cpp,
 * an independent double-precision FFT.  Bit-exactness with a real FFTW
 * build is NOT claimed (O(1e-15 log N) relative differences).
 */
#ifndef RO_ORACLE_FFTW3_SHIM_H
#define RO_ORACLE_FFTW3_SHIM_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef double fftw_complex[2];
typedef struct ro_shim_plan_s *fftw_plan;

#define FFTW_FORWARD  (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE  (0U)
#define FFTW_ESTIMATE (1U << 6)

void *fftw_malloc(size_t n);
void  fftw_free(void *p);
fftw_plan fftw_plan_dft_1d(int n, fftw_complex *in, fftw_complex *out,
                           int sign, unsigned flags);
void  fftw_execute(const fftw_plan plan);
void  fftw_destroy_plan(fftw_plan plan);

#ifdef __cplusplus
}
#endif

#endif
