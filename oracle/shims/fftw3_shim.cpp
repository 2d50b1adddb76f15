/*
 * oracle/shims/fftw3_shim.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Backs the FFTW3 API subset used by the reference (see fftw3.h) with an
 * independent double-precision transform:
 *   - power-of-two n : Stockham autosort, radix-4 passes (+ one radix-2),
 *                      twiddles generated in long double and rounded once;
 *   - other n        : direct O(n^2) DFT with long double accumulation
 *                      (only meant for small n in tests).
 * X[k] = sum_n x[n] exp(sign * 2*pi*i*k*n/N), unnormalised, out of place,
 * which is what fftw_plan_dft_1d(..., FFTW_FORWARD, ...) computes
 * (reference call site src/FFTBackend.cpp:120,236).
 */
#include "fftw3.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

struct ro_shim_plan_s {
	int n;
	int sign;
	fftw_complex *in;
	fftw_complex *out;
	bool pow2;
	std::vector<double> tw;      // n complex twiddles exp(sign*2*pi*i*k/n)
	std::vector<double> scratch; // n complex
};

static void make_twiddles(std::vector<double> &tw, int n, int sign)
{
	tw.resize(2 * (size_t)n);
	const long double two_pi = 8.0L * atanl(1.0L);
	for (int k = 0; k < n; k++) {
		long double a = two_pi * (long double)k / (long double)n;
		tw[2 * k]     = (double)cosl(a);
		tw[2 * k + 1] = (double)((long double)sign * sinl(a));
	}
}

extern "C" void *fftw_malloc(size_t n)
{
	void *p = NULL;
	if (posix_memalign(&p, 64, n ? n : 64) != 0) return NULL;
	return p;
}

extern "C" void fftw_free(void *p) { free(p); }

extern "C" fftw_plan fftw_plan_dft_1d(int n, fftw_complex *in, fftw_complex *out,
                                      int sign, unsigned /*flags*/)
{
	ro_shim_plan_s *p = new ro_shim_plan_s;
	p->n = n;
	p->sign = sign < 0 ? -1 : 1;
	p->in = in;
	p->out = out;
	p->pow2 = n > 0 && (n & (n - 1)) == 0;
	make_twiddles(p->tw, n, p->sign);
	p->scratch.resize(2 * (size_t)n);
	return p;
}

extern "C" void fftw_destroy_plan(fftw_plan plan) { delete plan; }

/* One Stockham radix-4 pass: length-n sub-transforms at stride s.
 * src/dst hold n*s complex values; tstep = N/n indexes the length-N table. */
static void pass4(int n, int s, int tstep, const double *tw, int sign,
                  const double *x, double *y)
{
	const int m = n / 4;
	for (int p = 0; p < m; p++) {
		const double w1r = tw[2 * (size_t)(p * tstep)],     w1i = tw[2 * (size_t)(p * tstep) + 1];
		const double w2r = tw[2 * (size_t)(2 * p * tstep)], w2i = tw[2 * (size_t)(2 * p * tstep) + 1];
		const double w3r = tw[2 * (size_t)(3 * p * tstep)], w3i = tw[2 * (size_t)(3 * p * tstep) + 1];
		const double *xa = x + 2 * (size_t)s * (p);
		const double *xb = x + 2 * (size_t)s * (p + m);
		const double *xc = x + 2 * (size_t)s * (p + 2 * m);
		const double *xd = x + 2 * (size_t)s * (p + 3 * m);
		double *y0 = y + 2 * (size_t)s * (4 * p);
		double *y1 = y0 + 2 * (size_t)s;
		double *y2 = y1 + 2 * (size_t)s;
		double *y3 = y2 + 2 * (size_t)s;
		for (int q = 0; q < s; q++) {
			const double ar = xa[2 * q], ai = xa[2 * q + 1];
			const double br = xb[2 * q], bi = xb[2 * q + 1];
			const double cr = xc[2 * q], ci = xc[2 * q + 1];
			const double dr = xd[2 * q], di = xd[2 * q + 1];
			const double apcr = ar + cr, apci = ai + ci;
			const double amcr = ar - cr, amci = ai - ci;
			const double bpdr = br + dr, bpdi = bi + di;
			/* (b - d) * (sign*i): forward (sign=-1) multiplies by -i */
			const double tr = br - dr, ti = bi - di;
			const double jr = (sign < 0) ? ti : -ti;
			const double ji = (sign < 0) ? -tr : tr;
			y0[2 * q]     = apcr + bpdr;
			y0[2 * q + 1] = apci + bpdi;
			const double u1r = amcr + jr, u1i = amci + ji;
			y1[2 * q]     = u1r * w1r - u1i * w1i;
			y1[2 * q + 1] = u1r * w1i + u1i * w1r;
			const double u2r = apcr - bpdr, u2i = apci - bpdi;
			y2[2 * q]     = u2r * w2r - u2i * w2i;
			y2[2 * q + 1] = u2r * w2i + u2i * w2r;
			const double u3r = amcr - jr, u3i = amci - ji;
			y3[2 * q]     = u3r * w3r - u3i * w3i;
			y3[2 * q + 1] = u3r * w3i + u3i * w3r;
		}
	}
}

/* Final radix-2 pass (n == 2 at stride s). */
static void pass2(int s, const double *x, double *y)
{
	for (int q = 0; q < s; q++) {
		const double ar = x[2 * q], ai = x[2 * q + 1];
		const double br = x[2 * (size_t)(q + s)], bi = x[2 * (size_t)(q + s) + 1];
		y[2 * q] = ar + br;
		y[2 * q + 1] = ai + bi;
		y[2 * (size_t)(q + s)] = ar - br;
		y[2 * (size_t)(q + s) + 1] = ai - bi;
	}
}

static void dft_direct(const ro_shim_plan_s *p)
{
	const int n = p->n;
	const double *x = (const double *)p->in;
	double *y = (double *)p->out;
	for (int k = 0; k < n; k++) {
		long double sr = 0, si = 0;
		for (int j = 0; j < n; j++) {
			size_t idx = (size_t)(((long long)k * j) % n);
			long double wr = p->tw[2 * idx], wi = p->tw[2 * idx + 1];
			sr += x[2 * j] * wr - x[2 * j + 1] * wi;
			si += x[2 * j] * wi + x[2 * j + 1] * wr;
		}
		y[2 * k] = (double)sr;
		y[2 * k + 1] = (double)si;
	}
}

extern "C" void fftw_execute(const fftw_plan p)
{
	const int N = p->n;
	if (N <= 0) return;
	if (!p->pow2) { dft_direct(p); return; }
	if (N == 1) { p->out[0][0] = p->in[0][0]; p->out[0][1] = p->in[0][1]; return; }

	/* Count passes so that the last one lands in p->out. */
	int passes = 0;
	for (int n = N; n > 1; n = (n >= 4) ? n / 4 : n / 2) passes++;

	const double *src = (const double *)p->in;
	double *bufs[2] = { (double *)p->out, p->scratch.data() };
	int which = (passes & 1) ? 0 : 1; /* first destination */
	int n = N, s = 1;
	while (n > 1) {
		double *dst = bufs[which];
		if (n >= 4) {
			pass4(n, s, N / n, p->tw.data(), p->sign, src, dst);
			n /= 4; s *= 4;
		} else {
			pass2(s, src, dst);
			n /= 2; s *= 2;
		}
		src = dst;
		which ^= 1;
	}
}
