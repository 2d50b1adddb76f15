// ro_static.cu -- ro_wf_long_rows_static (include/ro_cuda.h): the long-frame transform with no
// device allocation anywhere on its path.
//
// The long-frame path of ro_cuda.cu (launch_fft, split branch) runs two kernels per batch of frames:
// wf_split_kernel does the first radix-Na step of every frame into a scratch of packed complex
// floats, wf_fft2_kernel<14,32,16,MODE_SPLIT> then runs the Na 16k transforms of each frame out of
// it.  There the scratch and the three tables (window, twiddles of N, twiddles of 16384) are
// cudaMalloc'ed by ro_wf_create.  Here both live in file-scope __device__ arrays of a fixed size,
// which every GPU that loads this module holds until the process exits; the kernels are the same
// instantiations, launched through the same launchers (ro_v2.h), and are given the arrays' DEVICE
// addresses -- in host code the array names are host-side stand-ins, so the addresses are asked for
// once per device with cudaGetSymbolAddress while that device is current.
//
// Nothing in the scratch has to start at zero: the split kernel writes every element of
// [n_frames][Na][16384] before the transform kernel reads it, so the previous call's contents are
// simply overwritten.  Two calls that overlap on one GPU (two streams, two threads) share the one
// array and race.
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include <mutex>
#include <vector>

#include "../../include/ro_cuda.h"
#include "ro_v2.h"

__attribute__((visibility("hidden"))) int ro_internal_fail(int code, const char *fmt, ...);   // ro_cuda.cu: sets ro_last_error()

// the scratch: RO_STATIC_SCRATCH_BYTES / (8 * bins) frames per call
static __device__ __align__(256) unsigned char ro_static_scratch[RO_STATIC_SCRATCH_BYTES];

// the tables of ONE frame size at a time: window [bins] | exp(-2*pi*i*j/bins) [bins] float2 |
// exp(-2*pi*i*j/16384) [16384] float2
#define RO_STATIC_TABLE_FLOATS (3 * RO_STATIC_MAX_BINS + 2 * 16384)
static __device__ __align__(256) float ro_static_tables[RO_STATIC_TABLE_FLOATS];

namespace {

const int MAX_DEVICES = 64;

struct StaticDev {
	bool          ready;
	void         *scratch;       // device address of ro_static_scratch on this device
	float        *tables;        // device address of ro_static_tables on this device
	int           table_bins;    // frame size the tables hold, 0 = none
	ro::V2Kernel  k;             // wf_fft2_kernel<14,32,16,MODE_SPLIT>
	long long     grid_cap;      // resident CTAs of it on this device
};

StaticDev  g_dev[MAX_DEVICES];
std::mutex g_lock;

#define CUS(call)                                                                  \
	do {                                                                           \
		cudaError_t e_ = (call);                                                   \
		if (e_ != cudaSuccess)                                                     \
			return ro_internal_fail(RO_E_CUDA, "%s failed: %s (%s:%d)", #call,     \
			                        cudaGetErrorString(e_), __FILE__, __LINE__);   \
	} while (0)

// once per device, with that device current
int setup_device(int device, StaticDev *d)
{
	cudaDeviceProp prop;
	CUS(cudaGetDeviceProperties(&prop, device));
	if (prop.major != 10)
		return ro_internal_fail(RO_E_CUDA, "device %d (%s) is sm_%d%d: this library holds sm_100a code only and has no other path",
		                        device, prop.name, prop.major, prop.minor);
	void *p = NULL;
	CUS(cudaGetSymbolAddress(&p, ro_static_scratch));
	d->scratch = p;
	CUS(cudaGetSymbolAddress(&p, ro_static_tables));
	d->tables = static_cast<float *>(p);
	if (!d->scratch || !d->tables || ((uintptr_t)d->scratch & 255) != 0 || ((uintptr_t)d->tables & 255) != 0)
		return ro_internal_fail(RO_E_CUDA, "cudaGetSymbolAddress returned an unusable address on device %d", device);
	if (!ro::v2_pick(14, ro::V2_SPLIT, false, &d->k))
		return ro_internal_fail(RO_E_STATE, "no 16k split-path kernel in this build");
	CUS(cudaFuncSetAttribute(d->k.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d->k.smem));
	int occ = 0;
	CUS(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, d->k.fn, d->k.threads, d->k.smem));
	if (occ < 1) return ro_internal_fail(RO_E_CUDA, "the 16k split-path kernel does not fit an SM of device %d", device);
	d->grid_cap = (long long)occ * prop.multiProcessorCount;
	d->table_bins = 0;
	d->ready = true;
	return RO_OK;
}

// the tables of frame size N into the static array (host tables: the same two functions ro_wf_create uses)
int load_tables(StaticDev *d, int N, cudaStream_t s)
{
	std::vector<float> t((size_t)3 * N + 2 * 16384);
	int rc = ro_window_table(N, t.data());
	if (rc) return rc;
	rc = ro_twiddle_table(N, t.data() + N);
	if (rc) return rc;
	rc = ro_twiddle_table(16384, t.data() + (size_t)3 * N);
	if (rc) return rc;
	// whatever this stream still runs out of the old tables finishes first; the copy is complete on return
	CUS(cudaStreamSynchronize(s));
	d->table_bins = 0;
	CUS(cudaMemcpyAsync(d->tables, t.data(), t.size() * sizeof(float), cudaMemcpyHostToDevice, s));
	CUS(cudaStreamSynchronize(s));
	d->table_bins = N;
	return RO_OK;
}

}  // namespace

extern "C" int ro_wf_long_rows_static(int32_t device, int32_t bins, int32_t overlap, int32_t in_format,
                                      double iq_gain, const void *d_iq, size_t n_samples, float *d_rows,
                                      size_t max_frames, size_t *n_frames, void *cuda_stream, int sync)
{
	if (n_frames) *n_frames = 0;
	if (!d_iq || !d_rows) return ro_internal_fail(RO_E_INVALID, "ro_wf_long_rows_static: null argument");
	if (device < 0 || device >= MAX_DEVICES) return ro_internal_fail(RO_E_INVALID, "ro_wf_long_rows_static: device %d out of range", device);
	int log2n = 0;
	while ((1 << log2n) < bins && log2n < 30) log2n++;
	if (bins < 32768 || bins > RO_STATIC_MAX_BINS || (1 << log2n) != bins)
		return ro_internal_fail(RO_E_INVALID, "ro_wf_long_rows_static: bins = %d; this entry point takes the long frames only, "
		                        "a power of two from 32768 to %d (the static tables hold no more)", bins, RO_STATIC_MAX_BINS);
	if (in_format != RO_IN_INT16 && in_format != RO_IN_FLOAT32)
		return ro_internal_fail(RO_E_INVALID, "ro_wf_long_rows_static: device-resident input must be int16 or float32");
	if (overlap < 0) overlap = 0;                       // clamped like FFTBackend::FFTBackend (src/FFTBackend.cpp:108-109)
	if (overlap >= bins) overlap = bins - 1;
	const long long hop = bins - overlap;
	if (((uintptr_t)d_rows & 15) != 0) return ro_internal_fail(RO_E_INVALID, "device rows must be 16-byte aligned");
	if (in_format == RO_IN_FLOAT32 && ((uintptr_t)d_iq & 7) != 0)
		return ro_internal_fail(RO_E_INVALID, "float32 device input must be 8-byte aligned");
	size_t F = n_samples >= (size_t)bins ? (n_samples - (size_t)bins) / (size_t)hop + 1 : 0;
	if (F > max_frames) F = max_frames;
	const size_t cap_frames = (size_t)RO_STATIC_SCRATCH_BYTES / ((size_t)8 * (size_t)bins);
	if (F > cap_frames)
		return ro_internal_fail(RO_E_INVALID, "ro_wf_long_rows_static: %zu frames of %d bins need %zu bytes of scratch, the static array "
		                        "holds %zu (%zu frames of this size): pass fewer samples per call or a smaller max_frames",
		                        F, bins, F * (size_t)8 * (size_t)bins, (size_t)RO_STATIC_SCRATCH_BYTES, cap_frames);
	if (F == 0) return RO_OK;

	std::lock_guard<std::mutex> guard(g_lock);
	CUS(cudaSetDevice(device));
	StaticDev *d = &g_dev[device];
	if (!d->ready) {
		int rc = setup_device(device, d);
		if (rc) return rc;
	}
	cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
	if (d->table_bins != bins) {
		int rc = load_tables(d, bins, s);
		if (rc) return rc;
	}
	const float  *d_window = d->tables;
	const float2 *d_tw = reinterpret_cast<const float2 *>(d->tables + bins);
	const float2 *d_tw_nb = reinterpret_cast<const float2 *>(d->tables + (size_t)3 * bins);
	const int log2na = log2n - 14;

	// the two launches of launch_fft's split branch (ro_cuda.cu), on the static scratch
	CUS(ro::v2_launch_split(s, log2n, log2na, in_format == RO_IN_FLOAT32, d_iq, d_window, d_tw, d->scratch,
	                        0, (long long)F, hop, (float)iq_gain));
	const long long items = (long long)F << log2na;
	long long ipc = (items + d->grid_cap - 1) / d->grid_cap;      // contiguous items per CTA
	int pair = 1;                                                 // classes r, r+1 stay in one CTA ...
	if (log2na >= 2 && (ipc % 4 == 0 || ipc >= 16)) {             // ... or four of them (launch_fft, ro_cuda.cu)
		pair = 2;
		ipc = (ipc + 3) & ~3LL;
	} else {
		ipc += ipc & 1;
	}
	const int grid = (int)((items + ipc - 1) / ipc);
	CUS(ro::v2_launch(d->k, grid, s, d->scratch, NULL, d_tw_nb, d_rows, bins, items, 0, ipc, 0.f, log2na, pair,
	                  d->k.threads == 512 ? 14 : 0));
	if (n_frames) *n_frames = F;
	if (sync) CUS(cudaStreamSynchronize(s));
	return RO_OK;
}
