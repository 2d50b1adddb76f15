/*
 * host/shims/fftw3_min.cpp -- see fftw3.h.  Out-of-place complex DFT in double: iterative radix-2
 * with a precomputed twiddle table when n is a power of two, the O(n^2) definition otherwise
 * (FFTW accepts any n; every shipped configuration uses 2^15 .. 2^19).
 */
#include "fftw3.h"

#include <cmath>
#include <cstdlib>
#include <vector>

struct ro_min_fft_plan {
	int n, sign;
	fftw_complex *in, *out;
	std::vector<double> tw;       // cos, sin of 2*pi*k/n, k < n/2 (power of two) or k < n
	std::vector<int> rev;         // bit reversal (power of two)
};

extern "C" void *fftw_malloc(size_t n)
{
	void *p = NULL;
	return posix_memalign(&p, 64, n ? n : 64) == 0 ? p : NULL;
}

extern "C" void fftw_free(void *p) { free(p); }

extern "C" fftw_plan fftw_plan_dft_1d(int n, fftw_complex *in, fftw_complex *out, int sign, unsigned)
{
	if (n < 1 || !in || !out) return NULL;
	ro_min_fft_plan *p = new ro_min_fft_plan;
	p->n = n; p->sign = sign < 0 ? -1 : 1; p->in = in; p->out = out;
	const bool pow2 = (n & (n - 1)) == 0;
	const int nt = pow2 ? n / 2 : n;
	p->tw.resize(2 * (size_t)(nt > 0 ? nt : 1));
	const long double step = 8.0L * atanl(1.0L) / (long double)n;
	for (int k = 0; k < nt; k++) {
		p->tw[2 * k] = (double)cosl(step * k);
		p->tw[2 * k + 1] = (double)sinl(step * k);
	}
	if (pow2) {
		int bits = 0;
		while ((1 << bits) < n) bits++;
		p->rev.resize(n);
		for (int i = 0; i < n; i++) {
			int r = 0;
			for (int b = 0; b < bits; b++) r |= ((i >> b) & 1) << (bits - 1 - b);
			p->rev[i] = r;
		}
	}
	return p;
}

extern "C" void fftw_execute(const fftw_plan p)
{
	const int n = p->n;
	const double sg = (double)p->sign;
	if (p->rev.empty()) {                     // definition, any n
		std::vector<double> tmp(2 * (size_t)n);
		for (int k = 0; k < n; k++) {
			double re = 0, im = 0;
			for (int j = 0; j < n; j++) {
				const size_t e = (size_t)(((long long)k * j) % n);
				const double c = p->tw[2 * e], s = sg * p->tw[2 * e + 1];
				re += p->in[j][0] * c - p->in[j][1] * s;
				im += p->in[j][0] * s + p->in[j][1] * c;
			}
			tmp[2 * k] = re; tmp[2 * k + 1] = im;
		}
		for (int k = 0; k < n; k++) { p->out[k][0] = tmp[2 * k]; p->out[k][1] = tmp[2 * k + 1]; }
		return;
	}
	if (p->in == p->out) {
		for (int i = 0; i < n; i++) {
			const int r = p->rev[i];
			if (r > i) {
				const double a = p->out[i][0], b = p->out[i][1];
				p->out[i][0] = p->out[r][0]; p->out[i][1] = p->out[r][1];
				p->out[r][0] = a; p->out[r][1] = b;
			}
		}
	} else {
		for (int i = 0; i < n; i++) { p->out[p->rev[i]][0] = p->in[i][0]; p->out[p->rev[i]][1] = p->in[i][1]; }
	}
	for (int len = 2; len <= n; len <<= 1) {
		const int half = len / 2, stride = n / len;
		for (int base = 0; base < n; base += len) {
			for (int j = 0; j < half; j++) {
				const double c = p->tw[2 * (size_t)(j * stride)], s = sg * p->tw[2 * (size_t)(j * stride) + 1];
				double *a = p->out[base + j], *b = p->out[base + j + half];
				const double tr = b[0] * c - b[1] * s, ti = b[0] * s + b[1] * c;
				b[0] = a[0] - tr; b[1] = a[1] - ti;
				a[0] += tr; a[1] += ti;
			}
		}
	}
}

extern "C" void fftw_destroy_plan(fftw_plan p) { delete p; }
