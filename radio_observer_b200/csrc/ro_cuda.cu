// ro_cuda.cu -- C ABI (include/ro_cuda.h) over the sm_100a waterfall kernels.
//
// Host-side restatement of the framing rules of FFTBackend::process
// (reference src/FFTBackend.cpp:192-279): frame k covers samples
// [k*hop, k*hop + bins); the first frame needs `bins` samples; a tail shorter
// than a frame is carried to the next call and dropped at end of stream.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdarg.h>
#include <algorithm>
#include <deque>
#include <new>
#include <vector>

#include "../../include/ro_cuda.h"
#include "ro_kernels.cuh"
#include "ro_v2.h"
#include "ro_fft64.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
	return code;
}

#define CU(call)                                                              \
	do {                                                                      \
		cudaError_t e_ = (call);                                              \
		if (e_ != cudaSuccess)                                                \
			return fail(RO_E_CUDA, "%s failed: %s (%s:%d)", #call,            \
			            cudaGetErrorString(e_), __FILE__, __LINE__);          \
	} while (0)

typedef void (*det_kernel_t)(const ro::DetParams);

struct Pending {            // one submitted batch of the streaming path
	uint64_t first_row;
	size_t   n_rows;
	size_t   consumed;
	int      slot;          // -1: already moved to host vectors
	cudaEvent_t done;
	std::vector<float> rows;
	std::vector<ro_det_rec> det;
};

}  // namespace

// the same for the library's other translation units (ro_static.cu)
__attribute__((visibility("hidden"))) int ro_internal_fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
	return code;
}

struct ro_wf {
	ro_wf_config cfg;
	int N, log2n, overlap, hop, fmt;      // fmt: device encoding (int16 / float32)
	int in_bytes;                         // bytes per complex sample on the device
	int row_width;
	int n_sms;
	bool det_on;
	// extension stages (wf_reduce_kernel): R rows x D bins -> one RMS value
	int R, D, log2d;
	bool red_on;
	int out_bins;                         // N / D: floats per device row handed to the detector / D2H
	float *d_red[2];                      // batch path: reduced rows per slot
	float *d_sred[2];                     // streaming path: reduced rows per slot
	float *d_rscratch;                    // resident path: full rows of one chunk
	size_t rscratch_frames;
	// cfg.precision = 1: the reference's arithmetic in double (ro_fft64.cuh)
	bool fp64;
	double2 *d_tw64;                      // N/2 entries exp(-2*pi*i*j/N)
	double2 *d_x64;                       // [x64_frames][N]
	long long x64_frames;
	det_kernel_t det_fn;
	const char *last_kernel;              // what launch_fft ran last (ro_wf_kernel_name)
	char name_staged[96], name_direct[96], name_f32[96];
	// v2 kernels (int16 input, N = 4096..16384, symmetric window): [0] direct loads, [1] TMA-staged ring
	bool v2_ok;
	bool v2_f32;                          // ... or, for float32 input, MODE_DIRECT_F32 in both slots
	ro::V2Kernel v2[2];
	int v2_grid_cap[2];
	// frames longer than 16384 points: wf_split_kernel (first radix-Na step, once per frame) into a
	// cache-sized scratch, then wf_fft2_kernel<MODE_SPLIT> per residue class
	bool split_ok;
	int stagger;                          // chain the load bursts of warps sharing a scheduler (RO_STAGGER)
	int stagger_small;                    // the same for the 256-thread kernels (4k, 8k; RO_STAGGER_SMALL)
	ro::V2Kernel v2s;
	int v2s_grid_cap;
	// 32k and 64k frames of int16 input: radix-Na step fused into the 16k kernel (MODE_FUSED*), the
	// only scratch is a CTA-private slice that stays in L2
	bool fused_ok;
	ro::V2Kernel v2f;
	int v2f_grid_cap;
	void *d_fscratch;
	int log2na;
	float2 *d_tw_nb;                      // exp(-2*pi*i*j/16384)
	void   *d_scratch;                    // [scratch_frames][N] packed complex float
	long long scratch_frames;

	float  *d_window;
	float2 *d_tw;

	cudaStream_t own_stream, h2d_stream, d2h_stream;
	cudaStream_t user_stream;
	bool use_user_stream;

	// profiling
	bool profiling;
	std::vector<cudaEvent_t> ev_pool;
	std::vector<std::pair<cudaEvent_t, cudaEvent_t> > ev_used[2];
	uint64_t launches;

	// batch path (process_host) work space
	size_t chunk_frames;
	void  *d_in[2];
	float *d_rows[2];
	ro_det_rec *d_det[2];
	size_t d_in_bytes, d_rows_bytes;
	cudaEvent_t ev_h2d[2], ev_comp[2], ev_d2h[2];

	// streaming path
	size_t batch_frames;                  // submit threshold
	size_t stage_cap;                     // samples per staging slot
	unsigned char *h_stage[2];            // pinned input staging
	void  *d_sin[2];                      // device input ping-pong (carry + new)
	float *d_srows[2];                    // device rows per slot (full rows)
	ro_det_rec *d_sdet[2];
	float *h_srows[2];                    // pinned host rows per slot (row_width)
	ro_det_rec *h_sdet[2];
	cudaEvent_t ev_slot[2];
	bool   slot_used[2];
	size_t slot_rows_cap;
	int    cur;                           // slot being filled
	size_t staged;                        // new samples in h_stage[cur]
	size_t carry;                         // samples already on device at the front of d_sin[cur]
	uint64_t next_row;                    // global index of the next frame
	std::deque<Pending*> pending;
	bool sticky_error;
};

// ---------------------------------------------------------------------------
// host tables
// ---------------------------------------------------------------------------
extern "C" int ro_window_table(int32_t bins, float *out)
{
	if (bins < 2 || !out) return fail(RO_E_INVALID, "ro_window_table: bad arguments");
	// reference src/FFTBackend.cpp:155-186: float coefficients, double angle and
	// cosine, one rounding to float.  PI = 4*atan(1) (src/FFTBackend.cpp:100).
	const double PI = 4.0 * atan(1.0);
	const float a0 = 0.355768f, a1 = 0.487396f, a2 = 0.144232f, a3 = 0.012604f;
	for (int i = 0; i < bins; i++) {
		out[i] = (float)(a0
		         - a1 * cos(2.0 * PI * (float)i / (float)(bins - 1))
		         + a2 * cos(4.0 * PI * (float)i / (float)(bins - 1))
		         - a3 * cos(6.0 * PI * (float)i / (float)(bins - 1)));
	}
	return RO_OK;
}

extern "C" int ro_twiddle_table(int32_t n, float *out)
{
	if (n < 1 || !out) return fail(RO_E_INVALID, "ro_twiddle_table: bad arguments");
	const long double two_pi = 8.0L * atanl(1.0L);
	for (int j = 0; j < n; j++) {
		// reduce to the first octant so every entry is rounded from an argument <= pi/4
		int q = (int)(((long long)j * 8) / n);          // octant 0..7
		long double a;
		float c, s;
		long long r = (long long)j * 8 - (long long)q * n;   // j/n - q/8 = r/(8n)
		a = two_pi * (long double)r / (8.0L * (long double)n);
		long double ca = cosl(a), sa = sinl(a);
		if (q & 1) {                                   // angle = (q+1)*pi/4 - (pi/4 - a)
			long double b = two_pi / 8.0L - a;
			ca = cosl(b); sa = sinl(b);
			// cos((q+1)pi/4 - b), sin(...) with (q+1) even -> multiples of pi/2
			int m = ((q + 1) / 2) & 3;
			long double cc, ss;
			switch (m) {
			case 0: cc = ca;  ss = -sa; break;         // angle = -b
			case 1: cc = sa;  ss = ca;  break;         // pi/2 - b
			case 2: cc = -ca; ss = sa;  break;         // pi - b
			default: cc = -sa; ss = -ca; break;        // 3pi/2 - b
			}
			c = (float)cc; s = (float)ss;
		} else {
			int m = (q / 2) & 3;                       // angle = m*pi/2 + a
			long double cc, ss;
			switch (m) {
			case 0: cc = ca;  ss = sa;  break;
			case 1: cc = -sa; ss = ca;  break;
			case 2: cc = -ca; ss = -sa; break;
			default: cc = sa; ss = -ca; break;
			}
			c = (float)cc; s = (float)ss;
		}
		out[2 * j] = c;
		out[2 * j + 1] = -s;                           // exp(-i*theta)
	}
	return RO_OK;
}

extern "C" const char *ro_last_error(void) { return g_err; }
extern "C" const char *ro_version(void) { return "radio_observer_b200 0.1 (sm_100a)"; }

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
extern "C" int ro_dev_count(int *n)
{
	if (!n) return fail(RO_E_INVALID, "ro_dev_count: null");
	CU(cudaGetDeviceCount(n));
	return RO_OK;
}
extern "C" int ro_dev_alloc(int device, size_t bytes, void **d_ptr)
{
	if (!d_ptr) return fail(RO_E_INVALID, "ro_dev_alloc: null");
	CU(cudaSetDevice(device));
	CU(cudaMalloc(d_ptr, bytes ? bytes : 1));
	return RO_OK;
}
extern "C" int ro_dev_free(int device, void *d_ptr)
{
	CU(cudaSetDevice(device));
	CU(cudaFree(d_ptr));
	return RO_OK;
}
extern "C" int ro_dev_upload(int device, void *d_dst, const void *src, size_t bytes)
{
	CU(cudaSetDevice(device));
	CU(cudaMemcpy(d_dst, src, bytes, cudaMemcpyHostToDevice));
	return RO_OK;
}
extern "C" int ro_dev_download(int device, void *dst, const void *d_src, size_t bytes)
{
	CU(cudaSetDevice(device));
	CU(cudaMemcpy(dst, d_src, bytes, cudaMemcpyDeviceToHost));
	return RO_OK;
}
extern "C" int ro_dev_memset(int device, void *d_dst, int value, size_t bytes)
{
	CU(cudaSetDevice(device));
	CU(cudaMemset(d_dst, value, bytes));
	return RO_OK;
}
extern "C" int ro_host_alloc_pinned(size_t bytes, void **ptr)
{
	if (!ptr) return fail(RO_E_INVALID, "ro_host_alloc_pinned: null");
	CU(cudaMallocHost(ptr, bytes ? bytes : 1));
	return RO_OK;
}
extern "C" int ro_host_free_pinned(void *ptr)
{
	CU(cudaFreeHost(ptr));
	return RO_OK;
}

// ---------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------
namespace {

cudaStream_t batch_stream(ro_wf *h) { return h->use_user_stream ? h->user_stream : h->own_stream; }

int get_event(ro_wf *h, cudaEvent_t *ev)
{
	if (!h->ev_pool.empty()) {
		*ev = h->ev_pool.back();
		h->ev_pool.pop_back();
		return RO_OK;
	}
	CU(cudaEventCreate(ev));
	return RO_OK;
}

// Launch the fused kernel over n_frames frames of d_iq on stream s.
int launch_fft(ro_wf *h, cudaStream_t s, const void *d_iq, float *d_rows, long long row_stride,
               long long n_frames)
{
	if (n_frames <= 0) return RO_OK;
	if (h->fp64) {
		h->last_kernel = "wf64_load/block/stage/mag_kernel (FP64 variant)";
		cudaEvent_t e0 = 0, e1 = 0;
		if (h->profiling) {
			int rc = get_event(h, &e0); if (rc) return rc;
			rc = get_event(h, &e1); if (rc) return rc;
			CU(cudaEventRecord(e0, s));
		}
		const int blocks = h->n_sms * 8;
		for (long long f0 = 0; f0 < n_frames; f0 += h->x64_frames) {
			ro::Wf64Params p;
			p.iq = (const unsigned char *)d_iq + f0 * h->hop * h->in_bytes;
			p.window = h->d_window;
			p.tw = h->d_tw64;
			p.x = h->d_x64;
			p.rows = d_rows + f0 * row_stride;
			p.n_frames = std::min<long long>(h->x64_frames, n_frames - f0);
			p.hop = h->hop;
			p.row_stride = row_stride;
			p.iq_gain = h->cfg.iq_gain;
			p.log2n = h->log2n;
			p.stage = 0;
			if (h->fmt == ro::FMT_FLOAT32) ro::wf64_load_kernel<true><<<blocks, 256, 0, s>>>(p);
			else ro::wf64_load_kernel<false><<<blocks, 256, 0, s>>>(p);
			ro::wf64_block_kernel<<<blocks, 256, 0, s>>>(p);
			h->launches += 2;
			for (int st = ro::WF64_LOG2B + 1; st <= h->log2n; st++) {
				p.stage = st;
				ro::wf64_stage_kernel<<<blocks, 256, 0, s>>>(p);
				h->launches++;
			}
			ro::wf64_mag_kernel<<<blocks, 256, 0, s>>>(p);
			h->launches++;
			CU(cudaGetLastError());
		}
		if (h->profiling) {
			CU(cudaEventRecord(e1, s));
			h->ev_used[0].push_back(std::make_pair(e0, e1));
		}
		return RO_OK;
	}
	// every FP32 kernel stores 2 or 4 adjacent magnitudes with one vector store and reads float32 pairs
	// with 64-bit loads (ro_cuda.h states the contract)
	if (((uintptr_t)d_rows & 15) != 0 || (row_stride & 3) != 0)
		return fail(RO_E_INVALID, "device rows must be 16-byte aligned with a pitch that is a multiple of 4 floats");
	if (h->fmt == ro::FMT_FLOAT32 && ((uintptr_t)d_iq & 7) != 0)
		return fail(RO_E_INVALID, "float32 device input must be 8-byte aligned");
	// the fused long-frame kernel stages samples with cp.async.bulk: 16-byte aligned frame starts
	if (h->fused_ok && ((uintptr_t)d_iq & 15) == 0 && ((long long)h->hop * h->in_bytes) % 16 == 0) {
		h->last_kernel = h->log2na == 2 ? "wf_fft2_kernel<14,32,16,MODE_FUSED2> (4 x 16k per frame, TMA-staged load phase)"
		               : h->fmt == ro::FMT_FLOAT32 ? "wf_fft2_kernel<14,32,16,MODE_FUSED1_F32> (2 x 16k per frame, TMA-staged load phase)"
		                                           : "wf_fft2_kernel<14,32,16,MODE_FUSED1> (2 x 16k per frame, TMA-staged load phase)";
		const long long cap = h->v2f_grid_cap;
		const long long fpc = (n_frames + cap - 1) / cap;          // contiguous frames per CTA
		const int grid = (int)((n_frames + fpc - 1) / fpc);
		cudaEvent_t e0 = 0, e1 = 0;
		if (h->profiling) {
			int rc = get_event(h, &e0); if (rc) return rc;
			rc = get_event(h, &e1); if (rc) return rc;
			CU(cudaEventRecord(e0, s));
		}
		CU(ro::v2_launch_fused(h->v2f, grid, s, d_iq, h->d_window, h->d_tw, h->d_tw_nb, h->d_fscratch, d_rows,
		                       row_stride, n_frames, h->hop, fpc, (float)h->cfg.iq_gain, 1, h->stagger, h->log2na));
		h->launches++;
		if (h->profiling) {
			CU(cudaEventRecord(e1, s));
			h->ev_used[0].push_back(std::make_pair(e0, e1));
		}
		return RO_OK;
	}
	if (h->split_ok) {
		h->last_kernel = "wf_split_kernel + wf_fft2_kernel<MODE_SPLIT> (radix-Na step through a scratch, then Na transforms per frame)";
		cudaEvent_t e0 = 0, e1 = 0;
		if (h->profiling) {
			int rc = get_event(h, &e0); if (rc) return rc;
			rc = get_event(h, &e1); if (rc) return rc;
			CU(cudaEventRecord(e0, s));
		}
		for (long long f0 = 0; f0 < n_frames; f0 += h->scratch_frames) {
			const long long nf = std::min<long long>(h->scratch_frames, n_frames - f0);
			CU(ro::v2_launch_split(s, h->log2n, h->log2na, h->fmt == ro::FMT_FLOAT32, d_iq, h->d_window, h->d_tw,
			                       h->d_scratch, f0, nf, h->hop, (float)h->cfg.iq_gain));
			const long long items = nf << h->log2na;
			const long long cap = h->v2s_grid_cap;
			long long ipc = (items + cap - 1) / cap;                 // contiguous items per CTA
			int pair = (h->cfg.tuning & RO_TUNE_NO_PAIR) ? 0 : 1;
			// Na >= 4: four classes are stored together (pair = 2) when a CTA's share is a multiple of four or long
			// enough that rounding it up costs little; the short launches of the streaming path keep pairs
			if (pair && h->log2na >= 2 && (ipc % 4 == 0 || ipc >= 16)) {
				pair = 2;
				ipc = (ipc + 3) & ~3LL;
			} else if (pair) {
				ipc += ipc & 1;                                      // classes r, r+1 stay in one CTA
			}
			const int grid = (int)((items + ipc - 1) / ipc);
			CU(ro::v2_launch(h->v2s, grid, s, h->d_scratch, NULL, h->d_tw_nb, d_rows + f0 * row_stride, row_stride,
			                 items, 0, ipc, 0.f, h->log2na, pair, h->v2s.threads == 512 ? h->stagger : 0));
			h->launches += 2;
		}
		if (h->profiling) {
			CU(cudaEventRecord(e1, s));
			h->ev_used[0].push_back(std::make_pair(e0, e1));
		}
		return RO_OK;
	}
	if (h->v2_ok) {
		// the staged variant needs 16-byte aligned frame starts for cp.async.bulk
		const int st = (h->hop % 4 == 0 && ((uintptr_t)d_iq & 15) == 0) ? 1 : 0;
		h->last_kernel = h->v2_f32 ? h->name_f32 : st ? h->name_staged : h->name_direct;
		const long long cap = h->v2_grid_cap[st];
		const long long fpc = (n_frames + cap - 1) / cap;          // contiguous frames per CTA
		const int grid = (int)((n_frames + fpc - 1) / fpc);
		cudaEvent_t e0 = 0, e1 = 0;
		if (h->profiling) {
			int rc = get_event(h, &e0); if (rc) return rc;
			rc = get_event(h, &e1); if (rc) return rc;
			CU(cudaEventRecord(e0, s));
		}
		CU(ro::v2_launch(h->v2[st], grid, s, d_iq, h->d_window, h->d_tw, d_rows, row_stride, n_frames,
		                 h->hop, fpc, (float)h->cfg.iq_gain, 0, 0, h->v2[st].threads == 512 ? h->stagger : h->stagger_small));
		h->launches++;
		if (h->profiling) {
			CU(cudaEventRecord(e1, s));
			h->ev_used[0].push_back(std::make_pair(e0, e1));
		}
		return RO_OK;
	}
	return fail(RO_E_STATE, "no kernel for bins = %d on this handle", h->N);
}

int launch_detect(ro_wf *h, cudaStream_t s, const float *d_rows, long long row_stride,
                  long long n_rows, ro_det_rec *d_det)
{
	if (n_rows <= 0 || !h->det_on || !d_det) return RO_OK;
	ro::DetParams p;
	p.rows = d_rows;
	p.row_stride = row_stride;
	p.n_rows = n_rows;
	p.bins = h->out_bins;
	p.det_lo = h->cfg.det_lo;
	p.det_w = h->cfg.det_w;
	p.noise_lo = h->cfg.noise_lo;
	p.noise_w = h->cfg.noise_w;
	p.avg_bins = h->cfg.avg_bins;
	p.out = reinterpret_cast<float4 *>(d_det);
	const int wpb = 4;                                  // warps per block
	long long blocks = (n_rows + wpb - 1) / wpb;
	const long long cap = (long long)h->n_sms * 16;
	if (blocks > cap) blocks = cap;
	cudaEvent_t e0 = 0, e1 = 0;
	if (h->profiling) {
		int rc = get_event(h, &e0); if (rc) return rc;
		rc = get_event(h, &e1); if (rc) return rc;
		CU(cudaEventRecord(e0, s));
	}
	h->det_fn<<<(int)blocks, wpb * 32, 0, s>>>(p);
	CU(cudaGetLastError());
	h->launches++;
	if (h->profiling) {
		CU(cudaEventRecord(e1, s));
		h->ev_used[1].push_back(std::make_pair(e0, e1));
	}
	return RO_OK;
}

// n_out groups of R full rows at d_rows (pitch N) -> n_out rows of out_bins floats at d_out
int launch_reduce(ro_wf *h, cudaStream_t s, const float *d_rows, long long n_out, float *d_out)
{
	if (n_out <= 0) return RO_OK;
	ro::ReduceParams p;
	p.rows = d_rows;
	p.out = d_out;
	p.in_stride = h->N;
	p.out_stride = h->out_bins;
	p.n_out = n_out;
	p.bins = h->N;
	p.row_avg = h->R;
	p.log2d = h->log2d;
	long long blocks = (n_out * (h->N / 4) + 255) / 256;
	const long long cap = (long long)h->n_sms * 8;
	if (blocks > cap) blocks = cap;
	ro::wf_reduce_kernel<<<(int)blocks, 256, 0, s>>>(p);
	CU(cudaGetLastError());
	h->launches++;
	return RO_OK;
}

void free_all(ro_wf *h)
{
	if (!h) return;
	cudaSetDevice(h->cfg.device);
	for (Pending *p : h->pending) { if (p->done) cudaEventDestroy(p->done); delete p; }
	h->pending.clear();
	for (int i = 0; i < 2; i++) {
		cudaFree(h->d_in[i]); cudaFree(h->d_rows[i]); cudaFree(h->d_det[i]);
		cudaFree(h->d_sin[i]); cudaFree(h->d_srows[i]); cudaFree(h->d_sdet[i]);
		cudaFree(h->d_red[i]); cudaFree(h->d_sred[i]);
		if (h->h_stage[i]) cudaFreeHost(h->h_stage[i]);
		if (h->h_srows[i]) cudaFreeHost(h->h_srows[i]);
		if (h->h_sdet[i]) cudaFreeHost(h->h_sdet[i]);
		if (h->ev_h2d[i]) cudaEventDestroy(h->ev_h2d[i]);
		if (h->ev_comp[i]) cudaEventDestroy(h->ev_comp[i]);
		if (h->ev_d2h[i]) cudaEventDestroy(h->ev_d2h[i]);
		if (h->ev_slot[i]) cudaEventDestroy(h->ev_slot[i]);
		for (auto &pr : h->ev_used[i]) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
	}
	for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
	cudaFree(h->d_window);
	cudaFree(h->d_tw);
	cudaFree(h->d_tw_nb);
	cudaFree(h->d_scratch);
	cudaFree(h->d_rscratch);
	cudaFree(h->d_fscratch);
	cudaFree(h->d_tw64);
	cudaFree(h->d_x64);
	if (h->own_stream) cudaStreamDestroy(h->own_stream);
	if (h->h2d_stream) cudaStreamDestroy(h->h2d_stream);
	if (h->d2h_stream) cudaStreamDestroy(h->d2h_stream);
	delete h;
}

}  // namespace

extern "C" int ro_wf_create(const ro_wf_config *cfg, ro_wf **out)
{
	if (!cfg || !out) return fail(RO_E_INVALID, "ro_wf_create: null argument");
	*out = NULL;
	const int N = cfg->bins;
	if (N < 1024 || N > 524288 || (N & (N - 1)))
		return fail(RO_E_INVALID, "bins = %d: must be a power of two in [1024, 524288]", N);
	if (cfg->iq_phase_shift != 0)
		return fail(RO_E_INVALID, "iq_phase_shift = %d is not supported (the reference path "
		            "src/FFTBackend.cpp:53-92 is ill-defined for non-zero shifts)", cfg->iq_phase_shift);
	if (cfg->in_format < RO_IN_INT16 || cfg->in_format > RO_IN_COMPLEX64)
		return fail(RO_E_INVALID, "in_format = %d", cfg->in_format);
	if (cfg->n_bands < 0 || cfg->n_bands > RO_MAX_BANDS)
		return fail(RO_E_INVALID, "n_bands = %d", cfg->n_bands);
	const int R = cfg->row_avg > 0 ? cfg->row_avg : 1, D = cfg->bin_decim > 0 ? cfg->bin_decim : 1;
	if (cfg->row_avg < 0 || R > 65536)
		return fail(RO_E_INVALID, "row_avg = %d: must be in [1, 65536] (0 = off)", cfg->row_avg);
	if (cfg->bin_decim < 0 || D > 128 || (D & (D - 1)))
		return fail(RO_E_INVALID, "bin_decim = %d: must be a power of two in [1, 128] (0 = off)", cfg->bin_decim);
	if (cfg->precision != RO_PREC_FP32 && cfg->precision != RO_PREC_FP64)
		return fail(RO_E_INVALID, "precision = %d: 0 (FP32) or 1 (FP64)", cfg->precision);
	const int NO = N / D;                     // bins of a device row after decimation
	int width = 0;
	for (int b = 0; b < cfg->n_bands; b++) {
		if (cfg->band_lo[b] < 0 || cfg->band_hi[b] > NO || cfg->band_lo[b] >= cfg->band_hi[b])
			return fail(RO_E_INVALID, "band %d = [%d, %d) outside [0, %d)", b, cfg->band_lo[b], cfg->band_hi[b], NO);
		width += cfg->band_hi[b] - cfg->band_lo[b];
	}
	const bool det_on = cfg->det_w > 0;
	if (det_on) {
		if (cfg->det_lo < 0 || cfg->det_lo + cfg->det_w > NO || cfg->noise_lo < 0 || cfg->noise_w < 1 ||
		    cfg->noise_lo + cfg->noise_w > NO || cfg->avg_bins < 1)
			return fail(RO_E_INVALID, "detector bands outside the row (det [%d,+%d) noise [%d,+%d) avg %d)",
			            cfg->det_lo, cfg->det_w, cfg->noise_lo, cfg->noise_w, cfg->avg_bins);
	}
	int ndev = 0;
	CU(cudaGetDeviceCount(&ndev));
	if (cfg->device < 0 || cfg->device >= ndev)
		return fail(RO_E_INVALID, "device %d of %d", cfg->device, ndev);
	CU(cudaSetDevice(cfg->device));

	ro_wf *h = new (std::nothrow) ro_wf();
	if (!h) return fail(RO_E_NOMEM, "out of memory");
	h->cfg = *cfg;
	h->N = N;
	h->log2n = ro::ilog2(N);
	int ov = cfg->overlap;
	if (ov < 0) ov = 0;                       // src/FFTBackend.cpp:108-109
	if (ov >= N) ov = N - 1;
	h->overlap = ov;
	h->hop = N - ov;
	h->fmt = cfg->in_format == RO_IN_INT16 ? ro::FMT_INT16 : ro::FMT_FLOAT32;
	h->in_bytes = h->fmt == ro::FMT_INT16 ? 4 : 8;
	h->R = R;
	h->D = D;
	h->log2d = ro::ilog2(D);
	h->red_on = R > 1 || D > 1;
	h->out_bins = NO;
	h->row_width = cfg->n_bands ? width : NO;
	h->det_on = det_on;
	cudaDeviceProp prop;
	cudaError_t e = cudaGetDeviceProperties(&prop, cfg->device);
	if (e != cudaSuccess) { free_all(h); return fail(RO_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
	h->n_sms = prop.multiProcessorCount;
	if (prop.major != 10) {
		free_all(h);
		return fail(RO_E_CUDA, "device %d (%s) is sm_%d%d: this library holds sm_100a code only and has no other path",
		            cfg->device, prop.name, prop.major, prop.minor);
	}

	{   // detector variant by noise band width
		const int epl = (cfg->noise_w + 31) / 32;
		if (!det_on) h->det_fn = ro::wf_detect_kernel<0>;
		else if (epl <= 2) h->det_fn = ro::wf_detect_kernel<2>;
		else if (epl <= 4) h->det_fn = ro::wf_detect_kernel<4>;
		else if (epl <= 8) h->det_fn = ro::wf_detect_kernel<8>;
		else if (epl <= 16) h->det_fn = ro::wf_detect_kernel<16>;
		else if (epl <= 32) h->det_fn = ro::wf_detect_kernel<32>;
		else h->det_fn = ro::wf_detect_kernel<0>;
	}

	// tables
	std::vector<float> win(N), tw(2 * (size_t)N);
	ro_window_table(N, win.data());
	ro_twiddle_table(N, tw.data());
#define CUH(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { free_all(h); \
	return fail(RO_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); } } while (0)
	CUH(cudaMalloc(&h->d_window, sizeof(float) * N));
	CUH(cudaMalloc(&h->d_tw, sizeof(float2) * N));
	CUH(cudaMemcpy(h->d_window, win.data(), sizeof(float) * N, cudaMemcpyHostToDevice));
	CUH(cudaMemcpy(h->d_tw, tw.data(), sizeof(float2) * N, cudaMemcpyHostToDevice));
	bool sym = true;                          // w[i] == w[N-1-i] bit for bit
	for (int i = 0; i < N / 2 && sym; i++) sym = memcmp(&win[i], &win[N - 1 - i], sizeof(float)) == 0;
	{   // v2 eligibility: an instantiated size and a symmetric window
		const bool gain = cfg->iq_gain != 0.0;
		h->v2_ok = false;
		if (sym && h->fmt == ro::FMT_INT16 &&
		    ro::v2_pick(h->log2n, ro::V2_DIRECT, gain, &h->v2[0]) && ro::v2_pick(h->log2n, ro::V2_STAGED, gain, &h->v2[1])) {
			h->v2_ok = true;
			for (int st = 0; st < 2; st++) {
				CUH(cudaFuncSetAttribute(h->v2[st].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->v2[st].smem));
				int occ2 = 0;
				CUH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, h->v2[st].fn, h->v2[st].threads, h->v2[st].smem));
				if (ro::v2_tmem_cols(h->v2[st]) > 0) occ2 = std::min(occ2, 512 / ro::v2_tmem_cols(h->v2[st]));   // tensor-memory columns per SM
				if (occ2 < 1) { h->v2_ok = false; break; }
				h->v2_grid_cap[st] = occ2 * h->n_sms;
			}
		}
		// float32 input (raw TCP / JACK / the adapter's narrowed Complex): the same kernel reading packed
		// float pairs from global memory; both slots hold it
		h->v2_f32 = false;
		if (sym && h->fmt == ro::FMT_FLOAT32 && ro::v2_pick_f32(h->log2n, gain, &h->v2[0])) {
			h->v2[1] = h->v2[0];
			CUH(cudaFuncSetAttribute(h->v2[0].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->v2[0].smem));
			int occ2 = 0;
			CUH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, h->v2[0].fn, h->v2[0].threads, h->v2[0].smem));
			if (ro::v2_tmem_cols(h->v2[0]) > 0) occ2 = std::min(occ2, 512 / ro::v2_tmem_cols(h->v2[0]));   // tensor-memory columns per SM
			if (occ2 >= 1) {
				h->v2_grid_cap[0] = h->v2_grid_cap[1] = occ2 * h->n_sms;
				h->v2_ok = h->v2_f32 = true;
			}
		}
		if (h->v2_ok) {
			const int P = h->v2[0].threads ? N / h->v2[0].threads : 0;
			snprintf(h->name_staged, sizeof h->name_staged, "wf_fft2_kernel<%d,%d,%d,MODE_STAGED%s> (TMA-staged sample ring)", h->log2n, P, N / (P * P), gain ? ",GAIN" : "");
			snprintf(h->name_direct, sizeof h->name_direct, "wf_fft2_kernel<%d,%d,%d,MODE_DIRECT%s> (unaligned frame starts)", h->log2n, P, N / (P * P), gain ? ",GAIN" : "");
			snprintf(h->name_f32, sizeof h->name_f32, "wf_fft2_kernel<%d,%d,%d,MODE_DIRECT_F32%s>", h->log2n, P, N / (P * P), gain ? ",GAIN" : "");
		}
	}
	h->fp64 = cfg->precision == RO_PREC_FP64;
	if (h->fp64) {
		std::vector<double> tw64((size_t)N);                       // N/2 complex entries
		const long double two_pi = 8.0L * atanl(1.0L);
		for (int j = 0; j < N / 2; j++) {
			const long double a = two_pi * (long double)j / (long double)N;
			tw64[2 * (size_t)j] = (double)cosl(a);
			tw64[2 * (size_t)j + 1] = (double)-sinl(a);
		}
		CUH(cudaMalloc((void **)&h->d_tw64, sizeof(double) * (size_t)N));
		CUH(cudaMemcpy(h->d_tw64, tw64.data(), sizeof(double) * (size_t)N, cudaMemcpyHostToDevice));
		h->x64_frames = std::max<long long>(1, ((long long)256 << 20) / ((long long)sizeof(double2) * N));
		CUH(cudaMalloc((void **)&h->d_x64, sizeof(double2) * (size_t)N * (size_t)h->x64_frames));
	}
	h->split_ok = false;
	// passes 2 and 3 chained (bits 1, 2): +19 % on the 16k kernel (B200, 21.0 -> 25.0 M frames/s);
	// chaining pass 1 as well (bit 0) costs 5 %: its warps already arrive staggered from pass 3
	// bit 3: the pass-2 chain releases the warp two levels down (0.7 .. 3.9 % over distance 1)
	h->stagger = 14;
	// 8k kernel (two 256-thread CTAs per SM): chaining passes 2 and 3 gives 43.5 -> 47.2 M frames/s;
	// the 4k kernel (three CTAs per SM) gains nothing
	h->stagger_small = h->log2n == 13 ? 6 : 0;
	if (cfg->tuning & RO_TUNE_NO_CHAIN) h->stagger = h->stagger_small = 0;
	int log2nb = 14;                              // in-CTA transform of the split path
	if (h->log2n - log2nb > 5) log2nb = h->log2n - 5;      // the split kernel does radix <= 32
	if (h->log2n > 14 && ro::v2_pick(log2nb, ro::V2_SPLIT, false, &h->v2s)) {
		h->log2na = h->log2n - log2nb;
		const int NBv = 1 << log2nb;
		CUH(cudaFuncSetAttribute(h->v2s.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->v2s.smem));
		int occ3 = 0;
		CUH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3, h->v2s.fn, h->v2s.threads, h->v2s.smem));
		if (ro::v2_tmem_cols(h->v2s) > 0) occ3 = std::min(occ3, 512 / ro::v2_tmem_cols(h->v2s));   // tensor-memory columns per SM
		if (occ3 >= 1) {
			h->v2s_grid_cap = occ3 * h->n_sms;
			std::vector<float> twb(2 * (size_t)NBv);
			ro_twiddle_table(NBv, twb.data());
			CUH(cudaMalloc(&h->d_tw_nb, sizeof(float2) * NBv));
			CUH(cudaMemcpy(h->d_tw_nb, twb.data(), sizeof(float2) * NBv, cudaMemcpyHostToDevice));
			// a batch = whole waves of the persistent grid.  Measured on B200 (32k/64k frames): a 40 MB
			// scratch that stays in the 126 MB L2 gives only 2 items per CTA per launch and loses more to
			// launch gaps and tails than a 256 MB scratch loses to HBM (3.9 vs 4.5 M frames/s at 32k);
			// cfg.scratch_mb overrides
			const long long na = 1LL << h->log2na;
			long long unit = h->v2s_grid_cap, a = unit, b = na;     // frames per wave = cap / gcd(cap, Na)
			while (b) { long long t2 = a % b; a = b; b = t2; }
			unit /= a;
			// 512 MB: 64k frames 2.27 -> 2.40 M frames/s against 256 MB (B200, profiles/r02_scratch_mb_ab.txt)
			long long budget_mb = 512;
			if (cfg->scratch_mb > 0) budget_mb = cfg->scratch_mb;
			const long long frame_bytes = 8LL * N;
			long long units = (budget_mb << 20) / (unit * frame_bytes);
			if (units < 1) units = 1;
			// a CTA takes its items in pairs (Na = 2) or fours (Na >= 4), see the store stage of MODE_SPLIT: whole
			// waves of pairs / fours keep every SM busy (13 items per CTA rounded up to 16 left 20 of 148 SMs idle)
			if (!(cfg->tuning & RO_TUNE_NO_PAIR)) {
				const long long grp = h->log2na >= 2 ? 4 : 2;
				const long long ipu = std::max<long long>(1, unit * na / h->v2s_grid_cap);   // items per CTA and unit
				long long g2 = grp, x = ipu % grp;
				while (x) { long long t3 = g2 % x; g2 = x; x = t3; }                           // gcd(grp, ipu)
				const long long step = grp / g2;
				if (units > step) units -= units % step;
			}
			// the split kernel indexes frames with blockIdx.y: at most 65535 per launch
			if (units * unit > 65535) units = std::max<long long>(1, 65535 / unit);
			h->scratch_frames = std::min<long long>(units * unit, 65535);
			CUH(cudaMalloc(&h->d_scratch, (size_t)(h->scratch_frames * frame_bytes)));
			h->split_ok = true;
		}
	}
	h->fused_ok = false;
	// 32k frames (int16, float32) and 64k frames (int16) run in ONE kernel when the window is symmetric
	if (h->split_ok && sym && log2nb == 14 && !(cfg->tuning & RO_TUNE_NO_FUSED) &&
	    ((h->log2na == 1 && (h->fmt == ro::FMT_INT16 ? ro::v2_pick(14, ro::V2_FUSED1, false, &h->v2f) : ro::v2_pick_fused_f32(&h->v2f))) ||
	     (h->log2na == 2 && h->fmt == ro::FMT_INT16 && ro::v2_pick(14, ro::V2_FUSED2, false, &h->v2f)))) {
		CUH(cudaFuncSetAttribute(h->v2f.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->v2f.smem));
		int occ4 = 0;
		CUH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ4, h->v2f.fn, h->v2f.threads, h->v2f.smem));
		if (ro::v2_tmem_cols(h->v2f) > 0) occ4 = std::min(occ4, 512 / ro::v2_tmem_cols(h->v2f));   // tensor-memory columns per SM
		if (occ4 >= 1) {
			h->v2f_grid_cap = occ4 * h->n_sms;
			const size_t per_cta = (((size_t)1 << h->log2na) - 1) * ((size_t)1 << 14) * sizeof(float2);
			CUH(cudaMalloc(&h->d_fscratch, per_cta * (size_t)h->v2f_grid_cap));
			h->fused_ok = true;
		}
	}
	CUH(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
	CUH(cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking));
	CUH(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
	for (int i = 0; i < 2; i++) {
		CUH(cudaEventCreateWithFlags(&h->ev_h2d[i], cudaEventDisableTiming));
		CUH(cudaEventCreateWithFlags(&h->ev_comp[i], cudaEventDisableTiming));
		CUH(cudaEventCreateWithFlags(&h->ev_d2h[i], cudaEventDisableTiming));
		CUH(cudaEventCreateWithFlags(&h->ev_slot[i], cudaEventDisableTiming));
	}
	h->batch_frames = cfg->max_batch_frames > 0 ? (size_t)cfg->max_batch_frames : 64;
	// a batch is a whole number of row groups
	h->batch_frames = ((h->batch_frames + (size_t)R - 1) / (size_t)R) * (size_t)R;
	*out = h;
	return RO_OK;
}

extern "C" void ro_wf_destroy(ro_wf *h) { free_all(h); }

extern "C" int ro_wf_row_width(const ro_wf *h) { return h ? h->row_width : 0; }
extern "C" int ro_wf_overlap(const ro_wf *h) { return h ? h->overlap : 0; }
extern "C" int ro_wf_out_bins(const ro_wf *h) { return h ? h->out_bins : 0; }

extern "C" size_t ro_wf_frame_count(const ro_wf *h, size_t n_samples)
{
	if (!h || n_samples < (size_t)h->N) return 0;
	return (n_samples - (size_t)h->N) / (size_t)h->hop + 1;
}

extern "C" size_t ro_wf_row_count(const ro_wf *h, size_t n_samples)
{
	return h ? ro_wf_frame_count(h, n_samples) / (size_t)h->R : 0;
}

extern "C" int ro_wf_set_stream(ro_wf *h, void *cuda_stream)
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	h->user_stream = (cudaStream_t)cuda_stream;
	h->use_user_stream = cuda_stream != NULL;
	return RO_OK;
}

extern "C" int ro_wf_sync(ro_wf *h)
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	CU(cudaSetDevice(h->cfg.device));
	CU(cudaStreamSynchronize(batch_stream(h)));
	CU(cudaStreamSynchronize(h->h2d_stream));
	CU(cudaStreamSynchronize(h->d2h_stream));
	return RO_OK;
}

extern "C" int ro_wf_set_profiling(ro_wf *h, int enable)
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	h->profiling = enable != 0;
	return RO_OK;
}

extern "C" int ro_wf_kernel_times(ro_wf *h, double ms_sum[2], uint64_t launches[2])
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	CU(cudaSetDevice(h->cfg.device));
	CU(cudaDeviceSynchronize());
	for (int k = 0; k < 2; k++) {
		double sum = 0.0;
		for (auto &pr : h->ev_used[k]) {
			float ms = 0.f;
			CU(cudaEventElapsedTime(&ms, pr.first, pr.second));
			sum += ms;
			h->ev_pool.push_back(pr.first);
			h->ev_pool.push_back(pr.second);
		}
		if (ms_sum) ms_sum[k] = sum;
		if (launches) launches[k] = h->ev_used[k].size();
		h->ev_used[k].clear();
	}
	return RO_OK;
}

extern "C" uint64_t ro_wf_launch_count(const ro_wf *h) { return h ? h->launches : 0; }

extern "C" const char *ro_wf_kernel_name(const ro_wf *h)
{
	if (!h) return "";
	if (h->last_kernel) return h->last_kernel;
	return h->fp64 ? "wf64_load/block/stage/mag_kernel (FP64 variant)" : "(nothing launched yet)";
}

// ---------------------------------------------------------------------------
// batch paths
// ---------------------------------------------------------------------------
extern "C" int ro_wf_process_resident(ro_wf *h, const void *d_iq, size_t n_samples,
                                      float *d_rows, ro_det_rec *d_det,
                                      size_t max_frames, size_t *n_frames, int sync)
{
	if (!h || !d_iq || !d_rows) return fail(RO_E_INVALID, "ro_wf_process_resident: null argument");
	if (h->cfg.in_format == RO_IN_COMPLEX64)
		return fail(RO_E_INVALID, "device-resident input must be int16 or float32");
	CU(cudaSetDevice(h->cfg.device));
	size_t F = ro_wf_row_count(h, n_samples);
	if (F > max_frames) F = max_frames;
	if (n_frames) *n_frames = F;
	cudaStream_t s = batch_stream(h);
	int rc;
	if (!h->red_on) {
		rc = launch_fft(h, s, d_iq, d_rows, h->N, (long long)F);
		if (rc) return rc;
	} else {
		// full rows of one chunk of row groups go through a scratch (about 256 MiB), then the reduction
		if (!h->d_rscratch) {
			size_t groups = ((size_t)256 << 20) / (sizeof(float) * (size_t)h->N * (size_t)h->R);
			if (groups < 1) groups = 1;
			h->rscratch_frames = groups * (size_t)h->R;
			CU(cudaMalloc((void **)&h->d_rscratch, h->rscratch_frames * (size_t)h->N * sizeof(float)));
		}
		const size_t gpc = h->rscratch_frames / (size_t)h->R;       // groups per chunk
		for (size_t g0 = 0; g0 < F; g0 += gpc) {
			const size_t ng = std::min(gpc, F - g0);
			const unsigned char *src = (const unsigned char *)d_iq + g0 * (size_t)h->R * (size_t)h->hop * (size_t)h->in_bytes;
			rc = launch_fft(h, s, src, h->d_rscratch, h->N, (long long)(ng * (size_t)h->R));
			if (rc) return rc;
			rc = launch_reduce(h, s, h->d_rscratch, (long long)ng, d_rows + g0 * (size_t)h->out_bins);
			if (rc) return rc;
		}
	}
	rc = launch_detect(h, s, d_rows, h->out_bins, (long long)F, d_det);
	if (rc) return rc;
	if (sync) CU(cudaStreamSynchronize(s));
	return RO_OK;
}

extern "C" int ro_wf_reduce_resident(ro_wf *h, const float *d_rows, size_t n_in, float *d_out,
                                     size_t *n_out, int sync)
{
	if (!h || !d_rows || !d_out) return fail(RO_E_INVALID, "ro_wf_reduce_resident: null argument");
	CU(cudaSetDevice(h->cfg.device));
	const size_t n = n_in / (size_t)h->R;
	if (n_out) *n_out = n;
	cudaStream_t s = batch_stream(h);
	int rc = launch_reduce(h, s, d_rows, (long long)n, d_out);
	if (rc) return rc;
	if (sync) CU(cudaStreamSynchronize(s));
	return RO_OK;
}

extern "C" int ro_wf_detect_resident(ro_wf *h, const float *d_rows, size_t n_rows,
                                     ro_det_rec *d_det, int sync)
{
	if (!h || !d_rows || !d_det) return fail(RO_E_INVALID, "ro_wf_detect_resident: null argument");
	if (!h->det_on) return fail(RO_E_STATE, "detector not configured (det_w = 0)");
	CU(cudaSetDevice(h->cfg.device));
	cudaStream_t s = batch_stream(h);
	int rc = launch_detect(h, s, d_rows, h->out_bins, (long long)n_rows, d_det);
	if (rc) return rc;
	if (sync) CU(cudaStreamSynchronize(s));
	return RO_OK;
}

namespace {

// D2H of the configured bands (or full rows) of n rows starting at d_rows.
int copy_rows_out(ro_wf *h, cudaStream_t s, float *dst, const float *d_rows, size_t n)
{
	if (!dst || n == 0) return RO_OK;
	if (h->cfg.n_bands == 0) {
		CU(cudaMemcpyAsync(dst, d_rows, sizeof(float) * (size_t)h->out_bins * n, cudaMemcpyDeviceToHost, s));
		return RO_OK;
	}
	size_t off = 0;
	for (int b = 0; b < h->cfg.n_bands; b++) {
		const size_t w = (size_t)(h->cfg.band_hi[b] - h->cfg.band_lo[b]);
		CU(cudaMemcpy2DAsync(dst + off, sizeof(float) * (size_t)h->row_width,
		                     d_rows + h->cfg.band_lo[b], sizeof(float) * (size_t)h->out_bins,
		                     sizeof(float) * w, n, cudaMemcpyDeviceToHost, s));
		off += w;
	}
	return RO_OK;
}

int ensure_batch_buffers(ro_wf *h)
{
	if (h->d_in[0]) return RO_OK;
	// about 256 MiB of rows per slot, at least 1 frame, at most 8192 frames
	size_t cf = ((size_t)256 << 20) / (sizeof(float) * (size_t)h->N);
	if (cf < 1) cf = 1;
	if (cf > 8192) cf = 8192;
	cf = std::max((size_t)1, cf / (size_t)h->R) * (size_t)h->R;       // whole row groups
	h->chunk_frames = cf;
	h->d_in_bytes = ((cf - 1) * (size_t)h->hop + (size_t)h->N) * (size_t)h->in_bytes;
	h->d_rows_bytes = cf * (size_t)h->N * sizeof(float);
	for (int i = 0; i < 2; i++) {
		CU(cudaMalloc(&h->d_in[i], h->d_in_bytes));
		CU(cudaMalloc((void **)&h->d_rows[i], h->d_rows_bytes));
		CU(cudaMalloc((void **)&h->d_det[i], sizeof(ro_det_rec) * cf));
		if (h->red_on) CU(cudaMalloc((void **)&h->d_red[i], (cf / (size_t)h->R) * (size_t)h->out_bins * sizeof(float)));
	}
	return RO_OK;
}

// narrow reference `Complex` {double,double} to float32 pairs
void narrow_complex64(float *dst, const double *src, size_t n_samples)
{
	for (size_t i = 0; i < 2 * n_samples; i++) dst[i] = (float)src[i];
}

}  // namespace

extern "C" int ro_wf_process_host(ro_wf *h, const void *iq, size_t n_samples,
                                  float *rows, ro_det_rec *det,
                                  size_t max_frames, size_t *n_frames)
{
	if (!h || !iq) return fail(RO_E_INVALID, "ro_wf_process_host: null argument");
	CU(cudaSetDevice(h->cfg.device));
	const size_t R = (size_t)h->R;
	size_t G = ro_wf_row_count(h, n_samples);        // output rows
	if (G > max_frames) G = max_frames;
	if (n_frames) *n_frames = G;
	if (G == 0) return RO_OK;
	const size_t F = G * R;                          // frames transformed
	int rc = ensure_batch_buffers(h);
	if (rc) return rc;

	std::vector<float> narrowed;
	const unsigned char *src = (const unsigned char *)iq;
	if (h->cfg.in_format == RO_IN_COMPLEX64) {
		const size_t used = (F - 1) * (size_t)h->hop + (size_t)h->N;
		narrowed.resize(2 * used);
		narrow_complex64(narrowed.data(), (const double *)iq, used);
		src = (const unsigned char *)narrowed.data();
	}

	cudaStream_t sc = batch_stream(h);
	const size_t cf = h->chunk_frames;
	size_t c = 0;
	for (size_t f0 = 0; f0 < F; f0 += cf, c++) {
		const int s = (int)(c & 1);
		const size_t nf = std::min(cf, F - f0);
		const size_t ns = (nf - 1) * (size_t)h->hop + (size_t)h->N;
		// input slot is free once the kernels of chunk c-2 are done
		if (c >= 2) CU(cudaStreamWaitEvent(h->h2d_stream, h->ev_comp[s], 0));
		CU(cudaMemcpyAsync(h->d_in[s], src + f0 * (size_t)h->hop * (size_t)h->in_bytes,
		                   ns * (size_t)h->in_bytes, cudaMemcpyHostToDevice, h->h2d_stream));
		CU(cudaEventRecord(h->ev_h2d[s], h->h2d_stream));
		CU(cudaStreamWaitEvent(sc, h->ev_h2d[s], 0));
		// rows slot is free once the D2H of chunk c-2 is done
		if (c >= 2) CU(cudaStreamWaitEvent(sc, h->ev_d2h[s], 0));
		rc = launch_fft(h, sc, h->d_in[s], h->d_rows[s], h->N, (long long)nf);
		if (rc) return rc;
		const size_t ng = nf / R, g0 = f0 / R;       // chunk_frames is a multiple of R
		const float *d_out = h->d_rows[s];
		if (h->red_on) {
			rc = launch_reduce(h, sc, h->d_rows[s], (long long)ng, h->d_red[s]);
			if (rc) return rc;
			d_out = h->d_red[s];
		}
		rc = launch_detect(h, sc, d_out, h->out_bins, (long long)ng, det ? h->d_det[s] : NULL);
		if (rc) return rc;
		CU(cudaEventRecord(h->ev_comp[s], sc));
		CU(cudaStreamWaitEvent(h->d2h_stream, h->ev_comp[s], 0));
		rc = copy_rows_out(h, h->d2h_stream, rows ? rows + g0 * (size_t)h->row_width : NULL, d_out, ng);
		if (rc) return rc;
		if (det && h->det_on)
			CU(cudaMemcpyAsync(det + g0, h->d_det[s], sizeof(ro_det_rec) * ng, cudaMemcpyDeviceToHost, h->d2h_stream));
		CU(cudaEventRecord(h->ev_d2h[s], h->d2h_stream));
	}
	CU(cudaStreamSynchronize(h->d2h_stream));
	CU(cudaStreamSynchronize(sc));
	CU(cudaStreamSynchronize(h->h2d_stream));
	return RO_OK;
}

// ---------------------------------------------------------------------------
// streaming path: Backend::process / endStream
// ---------------------------------------------------------------------------
namespace {

int ensure_stream_buffers(ro_wf *h)
{
	if (h->h_stage[0]) return RO_OK;
	const size_t N = (size_t)h->N, hop = (size_t)h->hop;
	h->stage_cap = h->batch_frames * hop + N;
	// up to R-1 frames wait on the device for their row group to complete: room for them in the
	// carry and in the row slots
	const size_t held = ((size_t)h->R - 1) * hop;
	h->slot_rows_cap = h->stage_cap / hop + 2 + ((size_t)h->R - 1);
	const size_t host_in_bytes = h->stage_cap * (size_t)h->in_bytes;
	for (int i = 0; i < 2; i++) {
		CU(cudaMallocHost((void **)&h->h_stage[i], host_in_bytes));
		CU(cudaMalloc(&h->d_sin[i], (h->stage_cap + N + held) * (size_t)h->in_bytes));
		CU(cudaMalloc((void **)&h->d_srows[i], h->slot_rows_cap * N * sizeof(float)));
		CU(cudaMalloc((void **)&h->d_sdet[i], h->slot_rows_cap * sizeof(ro_det_rec)));
		if (h->red_on) CU(cudaMalloc((void **)&h->d_sred[i], (h->slot_rows_cap / (size_t)h->R + 1) * (size_t)h->out_bins * sizeof(float)));
		CU(cudaMallocHost((void **)&h->h_srows[i], h->slot_rows_cap * (size_t)h->row_width * sizeof(float)));
		CU(cudaMallocHost((void **)&h->h_sdet[i], h->slot_rows_cap * sizeof(ro_det_rec)));
	}
	return RO_OK;
}

// wait for a batch; with `move_out` its results leave the pinned slot (the slot is about to be reused)
int retire(ro_wf *h, Pending *p, bool move_out)
{
	if (p->slot < 0) return RO_OK;
	CU(cudaEventSynchronize(h->ev_slot[p->slot]));
	if (!move_out) return RO_OK;
	p->rows.assign(h->h_srows[p->slot], h->h_srows[p->slot] + p->n_rows * (size_t)h->row_width);
	if (h->det_on) p->det.assign(h->h_sdet[p->slot], h->h_sdet[p->slot] + p->n_rows);
	p->slot = -1;
	return RO_OK;
}

// One batch of the streaming path, three streams (north_star: "pinned double-buffered staging with async
// H2D copies on a side stream"):
//   h2d_stream   new samples of the slot, once the kernels that last read this slot's device input are done
//   own_stream   transform, (reduce,) detector, then the device-to-device carry of the overlap into the
//                other slot
//   d2h_stream   band rows and detector records into the slot's pinned buffers; ev_slot marks them ready
// so the upload of batch k+1 and the download of batch k-1 run beside the kernels of batch k.
int submit(ro_wf *h)
{
	const size_t total = h->carry + h->staged;
	const size_t G = ro_wf_row_count(h, total);           // whole row groups only; the rest stays staged
	const size_t F = G * (size_t)h->R;
	if (F == 0) return RO_OK;
	const int s = h->cur;
	cudaStream_t sc = h->own_stream;
	unsigned char *d_in = (unsigned char *)h->d_sin[s];
	if (h->staged) {
		if (h->slot_used[s]) CU(cudaStreamWaitEvent(h->h2d_stream, h->ev_comp[s], 0));
		CU(cudaMemcpyAsync(d_in + h->carry * (size_t)h->in_bytes, h->h_stage[s],
		                   h->staged * (size_t)h->in_bytes, cudaMemcpyHostToDevice, h->h2d_stream));
		CU(cudaEventRecord(h->ev_h2d[s], h->h2d_stream));
		CU(cudaStreamWaitEvent(sc, h->ev_h2d[s], 0));
	}
	// the slot's device rows and records are free once their previous download is done
	if (h->slot_used[s]) CU(cudaStreamWaitEvent(sc, h->ev_slot[s], 0));
	int rc = launch_fft(h, sc, d_in, h->d_srows[s], h->N, (long long)F);
	if (rc) return rc;
	const float *d_out = h->d_srows[s];
	if (h->red_on) {
		rc = launch_reduce(h, sc, h->d_srows[s], (long long)G, h->d_sred[s]);
		if (rc) return rc;
		d_out = h->d_sred[s];
	}
	rc = launch_detect(h, sc, d_out, h->out_bins, (long long)G, h->d_sdet[s]);
	if (rc) return rc;
	const size_t left = total - F * (size_t)h->hop;       // overlap carried to the next frame
	if (left) {
		// the other slot's input was last read by the kernels of the previous batch: same stream, in order;
		// its upload region starts behind the carry, so the two copies do not meet
		CU(cudaMemcpyAsync(h->d_sin[s ^ 1], d_in + F * (size_t)h->hop * (size_t)h->in_bytes,
		                   left * (size_t)h->in_bytes, cudaMemcpyDeviceToDevice, sc));
	}
	CU(cudaEventRecord(h->ev_comp[s], sc));
	CU(cudaStreamWaitEvent(h->d2h_stream, h->ev_comp[s], 0));
	rc = copy_rows_out(h, h->d2h_stream, h->h_srows[s], d_out, G);
	if (rc) return rc;
	if (h->det_on)
		CU(cudaMemcpyAsync(h->h_sdet[s], h->d_sdet[s], sizeof(ro_det_rec) * G, cudaMemcpyDeviceToHost, h->d2h_stream));
	CU(cudaEventRecord(h->ev_slot[s], h->d2h_stream));
	h->slot_used[s] = true;

	Pending *p = new (std::nothrow) Pending();
	if (!p) return fail(RO_E_NOMEM, "out of memory");
	p->first_row = h->next_row;
	p->n_rows = G;
	p->consumed = 0;
	p->slot = s;
	p->done = 0;
	h->pending.push_back(p);
	h->next_row += G;
	h->carry = left;
	h->staged = 0;
	h->cur = s ^ 1;
	// the slot we are about to fill must be idle -- its upload done (ev_slot lies behind it) -- and results
	// still waiting in its pinned buffers are moved out
	for (Pending *q : h->pending)
		if (q->slot == h->cur) { rc = retire(h, q, true); if (rc) return rc; }
	return RO_OK;
}

}  // namespace

extern "C" int ro_wf_push(ro_wf *h, const void *iq, size_t n_samples)
{
	if (!h || (!iq && n_samples)) return fail(RO_E_INVALID, "ro_wf_push: null argument");
	if (h->sticky_error) return fail(RO_E_CUDA, "handle is in a failed state");
	CU(cudaSetDevice(h->cfg.device));
	int rc = ensure_stream_buffers(h);
	if (rc) { h->sticky_error = true; return rc; }
	const size_t host_bytes = h->cfg.in_format == RO_IN_COMPLEX64 ? 16 : (size_t)h->in_bytes;
	const unsigned char *src = (const unsigned char *)iq;
	while (n_samples) {
		const size_t space = h->stage_cap - h->staged;
		const size_t take = std::min(space, n_samples);
		unsigned char *dst = h->h_stage[h->cur] + h->staged * (size_t)h->in_bytes;
		if (h->cfg.in_format == RO_IN_COMPLEX64) narrow_complex64((float *)dst, (const double *)src, take);
		else memcpy(dst, src, take * host_bytes);
		h->staged += take;
		src += take * host_bytes;
		n_samples -= take;
		const size_t F = ro_wf_frame_count(h, h->carry + h->staged);
		if (F >= h->batch_frames || h->staged == h->stage_cap) {
			rc = submit(h);
			if (rc) { h->sticky_error = true; return rc; }
		}
	}
	return RO_OK;
}

extern "C" int ro_wf_flush(ro_wf *h)
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	if (h->sticky_error) return fail(RO_E_CUDA, "handle is in a failed state");
	CU(cudaSetDevice(h->cfg.device));
	int rc = ensure_stream_buffers(h);
	if (rc) return rc;
	rc = submit(h);
	if (rc) h->sticky_error = true;
	return rc;
}

extern "C" int ro_wf_poll(ro_wf *h, float *rows, size_t max_rows, ro_det_rec *det,
                          size_t *n_rows, uint64_t *first_row)
{
	if (!h || !n_rows) return fail(RO_E_INVALID, "ro_wf_poll: null argument");
	*n_rows = 0;
	if (h->sticky_error) return fail(RO_E_CUDA, "handle is in a failed state");
	CU(cudaSetDevice(h->cfg.device));
	size_t got = 0;
	while (got < max_rows && !h->pending.empty()) {
		Pending *p = h->pending.front();
		int rc = retire(h, p, false);
		if (rc) { h->sticky_error = true; return rc; }
		if (got == 0 && first_row) *first_row = p->first_row + p->consumed;
		const size_t take = std::min(max_rows - got, p->n_rows - p->consumed);
		// straight from the pinned slot unless the slot had to be vacated for a later batch
		const float *src_rows = p->slot >= 0 ? h->h_srows[p->slot] : p->rows.data();
		const ro_det_rec *src_det = p->slot >= 0 ? h->h_sdet[p->slot] : p->det.data();
		if (rows)
			memcpy(rows + got * (size_t)h->row_width, src_rows + p->consumed * (size_t)h->row_width,
			       take * (size_t)h->row_width * sizeof(float));
		if (det && h->det_on) memcpy(det + got, src_det + p->consumed, take * sizeof(ro_det_rec));
		p->consumed += take;
		got += take;
		if (p->consumed == p->n_rows) { h->pending.pop_front(); delete p; }
	}
	*n_rows = got;
	return RO_OK;
}

extern "C" int ro_wf_reset(ro_wf *h)
{
	if (!h) return fail(RO_E_INVALID, "null handle");
	CU(cudaSetDevice(h->cfg.device));
	CU(cudaStreamSynchronize(h->h2d_stream));
	CU(cudaStreamSynchronize(h->own_stream));
	CU(cudaStreamSynchronize(h->d2h_stream));
	h->slot_used[0] = h->slot_used[1] = false;
	for (Pending *p : h->pending) delete p;
	h->pending.clear();
	h->carry = h->staged = 0;
	h->next_row = 0;
	h->cur = 0;
	h->sticky_error = false;
	return RO_OK;
}
