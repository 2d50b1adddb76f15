/*
 * ro_cuda.h -- C ABI of the B200-native waterfall backend (libro_cuda.so).
 *
 * This is the drop-in boundary for the spectral hot path of radio-observer.
 * The reference has no C FFI for this path: its boundary is the C++ virtual
 * interface `Backend` (reference src/Backend.h:83-102) implemented by
 * FFTBackend / WaterfallBackend and the per-row statistics of BolidRecorder.
 * Each entry point below names the reference code it replaces; the C++ adapter
 * (host/, factory "waterfall_cuda") and the ctypes binding used by the tests
 * call exactly these functions.  See INTEGRATION.md for the reference-side stub.
 *
 * Conventions: every function returning int returns 0 on success and a
 * negative RO_E_* code on failure; ro_last_error() gives the message of the
 * last failure on the calling thread.  No C++ exception crosses this ABI.
 * A handle is single-producer: call it from one thread at a time.
 *
 * Device resources beyond global memory: the kernels for frames of 16384 points and more (one CTA per SM,
 * about 229 KB of shared memory each) allocate all 512 columns of the SM's tensor memory for their lifetime
 * and use them as thread-private storage (twiddle tables, parked residue classes and magnitudes; no
 * tcgen05.mma).  Another kernel that wants tensor memory on the same SM at the same time (a cuBLAS / CUTLASS
 * GEMM of the same process on another stream) waits in its tcgen05.alloc until such a CTA exits, and the
 * other way round: it costs overlap, never progress.
 */
#ifndef RO_CUDA_H
#define RO_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RO_OK              0
#define RO_E_INVALID      -1   /* bad argument / unsupported configuration       */
#define RO_E_CUDA         -2   /* CUDA runtime error (sticky per handle)         */
#define RO_E_NOMEM        -3
#define RO_E_STATE        -4   /* call not valid in the handle's current state   */
#define RO_E_OVERFLOW     -5   /* caller buffer too small                        */

/* input encodings (reference src/WAVStream.cpp:115-121, src/RawStream.cpp:59-63,
 * src/Backend.h:26-41) */
#define RO_IN_INT16        0   /* interleaved little-endian int16 [I,Q], unscaled */
#define RO_IN_FLOAT32      1   /* interleaved float32 [I,Q]                       */
#define RO_IN_COMPLEX64    2   /* reference `Complex` {double re, im}; narrowed to
                                  float32 on the host before upload               */

#define RO_MAX_BANDS       4

#define RO_PREC_FP32       0
#define RO_PREC_FP64       1

typedef struct ro_wf ro_wf;   /* opaque: one stream on one GPU */

/* Configuration.  bins/overlap/iq_gain/iq_phase_shift carry the meaning of the
 * "waterfall" factory keys (reference src/WaterfallBackend.cpp:620-644,
 * clamps src/FFTBackend.cpp:108-109); det_* are BIN values as computed by
 * BolidRecorder::start (reference src/BolidRecorder.cpp:80-103). */
typedef struct ro_wf_config {
	int32_t bins;            /* N: power of two, 1024 .. 524288                  */
	int32_t overlap;         /* clamped to [0, bins-1] like the reference         */
	int32_t sample_rate;     /* Hz; bookkeeping only                              */
	int32_t in_format;       /* RO_IN_*                                           */
	double  iq_gain;         /* added to Q before the window (src/FFTBackend.cpp:71-79) */
	int32_t iq_phase_shift;  /* must be 0 (reference path is ill-defined otherwise) */
	int32_t n_bands;         /* 0 = host rows are full rows; else row-index bands */
	int32_t band_lo[RO_MAX_BANDS];   /* [lo, hi) in fftshifted row indices        */
	int32_t band_hi[RO_MAX_BANDS];
	int32_t det_lo, det_w;           /* detect band (peak search); det_w=0: detector off */
	int32_t noise_lo, noise_w;       /* noise band (lower-quartile estimate)      */
	int32_t avg_bins;                /* bins averaged around the peak             */
	int32_t device;                  /* CUDA device ordinal                        */
	int32_t max_batch_frames;        /* 0 = default; frames per kernel launch in the streaming path */
	/* Extension stages, no reference counterpart: 0 or 1 = off, which keeps the
	 * reference's rows.  When on, an output row is the RMS amplitude of row_avg consecutive
	 * waterfall rows x bin_decim adjacent bins, sqrt(mean |X|^2); a group of rows left incomplete at
	 * ro_wf_flush / at the end of a batch call is dropped like the reference drops an incomplete
	 * frame.  Bands and detector bins then index the REDUCED row (bins / bin_decim wide). */
	int32_t row_avg;                 /* rows accumulated per output row, >= 1                      */
	int32_t bin_decim;               /* adjacent bins per output bin: power of two, 1 .. 128       */
	/* 0 = packed FP32 transform (the fast path; rows within 1e-3 dB of the reference's);
	 * 1 = FP64 transform and magnitude like the reference's FFTW path (src/FFTBackend.cpp:117-120,
	 * src/WaterfallBackend.cpp:492-505): rows within 1e-5 relative in linear power on every bin. */
	int32_t precision;
	int32_t tuning;                  /* RO_TUNE_* bits, 0 = the measured defaults                  */
	int32_t scratch_mb;              /* frames longer than 32768 points: MiB of device scratch per
	                                    batch of the split path; 0 = 256                            */
	int32_t reserved[1];
} ro_wf_config;

/* tuning bits: A/B switches for measurements and tests; results do not depend on them */
#define RO_TUNE_NO_FUSED   1   /* 32768- and 65536-point frames through the split path instead of the fused kernels */
#define RO_TUNE_NO_PAIR    2   /* split path: every residue class stores its own bins (no paired stores) */
#define RO_TUNE_NO_CHAIN   4   /* the warps of a scheduler do not take turns with their shared-memory load bursts */

/* Per-row detector statistics: what BolidRecorder::update computes before its
 * decision (reference src/BolidRecorder.cpp:121-132, 302-347). */
typedef struct ro_det_rec {
	float   noise;   /* 2 x the floor(w/4)-th smallest value of the noise band   */
	int32_t peak;    /* offset of the LAST maximum inside the detect band         */
	float   avg;     /* mean of avg_bins values starting at det_lo+peak-avg_bins/2 */
	int32_t pad;
} ro_det_rec;

/* ---- lifecycle --------------------------------------------------------- */

/* Replaces FFTBackend::FFTBackend + startStream (src/FFTBackend.cpp:103-127,
 * 144-189): validates N, builds the Nuttall table and twiddles, allocates the
 * device work space. */
int  ro_wf_create(const ro_wf_config *cfg, ro_wf **out);
void ro_wf_destroy(ro_wf *h);

/* floats per row handed to the HOST by poll/process_host: sum of band widths,
 * or bins when n_bands == 0. */
int  ro_wf_row_width(const ro_wf *h);
/* the clamped overlap actually in use */
int  ro_wf_overlap(const ro_wf *h);
/* floats per row on the DEVICE: bins / bin_decim */
int  ro_wf_out_bins(const ro_wf *h);

/* ---- streaming path (Backend::process, src/FFTBackend.cpp:192-279) ------ */

/* Borrow n_samples complex samples (encoding cfg.in_format) for the duration
 * of the call; they are copied to pinned staging and uploaded asynchronously.
 * Frames are formed exactly as the reference does: frame k = samples
 * [k*hop, k*hop+bins); a tail shorter than one frame is kept for the next call. */
int  ro_wf_push(ro_wf *h, const void *iq, size_t n_samples);

/* Harvest finished rows (in row order).  rows: max_rows x row_width floats
 * (may be NULL to skip row data), det: max_rows records (may be NULL).
 * *first_row is the global index (frame number) of the first returned row.
 * Blocks only for work already submitted. */
int  ro_wf_poll(ro_wf *h, float *rows, size_t max_rows, ro_det_rec *det,
                size_t *n_rows, uint64_t *first_row);

/* Backend::endStream (src/FFTBackend.cpp:282-286): submit every complete
 * frame still pending; the incomplete tail is dropped as in the reference.
 * Rows remain available to ro_wf_poll. */
int  ro_wf_flush(ro_wf *h);

/* Restart the stream (frame counter and tail reset), keeping the plan. */
int  ro_wf_reset(ro_wf *h);

/* ---- batch paths (archive reprocessing, benchmarks) ---------------------- */

/* Number of frames n_samples yield: floor((S-N)/hop)+1, or 0. */
size_t ro_wf_frame_count(const ro_wf *h, size_t n_samples);
/* Number of OUTPUT rows n_samples yield: frame_count / row_avg. */
size_t ro_wf_row_count(const ro_wf *h, size_t n_samples);

/* Input already in device memory (d_iq: cfg.in_format, int16 or float32).
 * Writes FULL rows (ro_wf_out_bins floats each, fftshifted) to d_rows and, if the
 * detector is configured and d_det != NULL, one record per row; max_frames and
 * *n_frames count output rows.  Asynchronous on the handle's stream unless
 * `sync` is non-zero.
 * Alignment: d_rows 16 bytes (what cudaMalloc returns), float32 d_iq 8 bytes; RO_E_INVALID
 * otherwise.  int16 d_iq may start anywhere on a sample; frames that start on 16-byte boundaries
 * (d_iq 16-byte aligned and hop a multiple of 4 samples) take the faster, TMA-staged kernels.
 * One stream at a time: a handle's work buffers (long-frame scratch, FP64 scratch, reduce scratch)
 * are shared by all of its calls, so asynchronous calls must all be issued on the SAME stream
 * (ro_wf_set_stream between two of them needs a ro_wf_sync first), and the batch and streaming
 * entry points of one handle must not be in flight together. */
int  ro_wf_process_resident(ro_wf *h, const void *d_iq, size_t n_samples,
                            float *d_rows, ro_det_rec *d_det,
                            size_t max_frames, size_t *n_frames, int sync);

/* Host buffers in, host buffers out: chunked, double-buffered H2D / kernels /
 * D2H on three streams.  rows receives row_width floats per frame (bands or
 * full rows), det one record per frame (either may be NULL). */
int  ro_wf_process_host(ro_wf *h, const void *iq, size_t n_samples,
                        float *rows, ro_det_rec *det,
                        size_t max_frames, size_t *n_frames);

/* ---- allocation-free variant --------------------------------------------- */

/* NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The long-frame transform of ro_wf_process_resident -- wf_split_kernel into a scratch, then
 * wf_fft2_kernel<14,32,16,MODE_SPLIT> out of it -- with that scratch and the window / twiddle
 * tables in file-scope __device__ arrays of the library (radio_observer_b200/csrc/ro_static.cu)
 * instead of a handle's cudaMalloc'ed buffers.  No handle, and no cudaMalloc, cudaFree or other
 * allocating routine on this path: the first call on a device looks the arrays' addresses up and
 * the first call for a frame size copies its tables (1.7 MB at most) into them; after that a call
 * is two kernel launches.
 *
 *   bins       a power of two, 32768 .. RO_STATIC_MAX_BINS
 *   overlap, in_format (RO_IN_INT16 | RO_IN_FLOAT32), iq_gain: as in ro_wf_config
 *   d_iq       n_samples complex samples in device memory (float32: 8-byte aligned)
 *   d_rows     receives min(frame count, max_frames) FULL rows of `bins` floats, fftshifted
 *              magnitudes, pitch `bins` (16-byte aligned); *n_frames = rows written
 *   cuda_stream  the caller's cudaStream_t (NULL = the default stream); sync != 0 waits for it
 *
 * Limit: frames * 8 * bins <= RO_STATIC_SCRATCH_BYTES per call (256 / 128 / 64 frames of 32k / 64k /
 * 128k); more is RO_E_INVALID with the numbers in ro_last_error(), and nothing is launched.  Every
 * GPU that runs this entry point holds the arrays (RO_STATIC_SCRATCH_BYTES + 1.7 MB) until the
 * process exits.  ALL calls on a device share them: calls that overlap on the GPU (two streams, two
 * threads) race; issue them on one stream or wait in between.  The scratch keeps the last call's
 * contents and needs no clearing (every element is written before it is read). */
#define RO_STATIC_SCRATCH_BYTES (64u << 20)
#define RO_STATIC_MAX_BINS      131072
int  ro_wf_long_rows_static(int32_t device, int32_t bins, int32_t overlap, int32_t in_format,
                            double iq_gain, const void *d_iq, size_t n_samples, float *d_rows,
                            size_t max_frames, size_t *n_frames, void *cuda_stream, int sync);

/* ---- stage variants (unit tests, ncu attribution) ------------------------ */

/* Only the reduction kernel (row_avg x bin_decim of the handle): n_in full rows of `bins` floats
 * -> n_in / row_avg rows of bins / bin_decim floats. */
int  ro_wf_reduce_resident(ro_wf *h, const float *d_rows, size_t n_in, float *d_out,
                           size_t *n_out, int sync);

/* Only the detector kernel over n_rows full rows already on the device. */
int  ro_wf_detect_resident(ro_wf *h, const float *d_rows, size_t n_rows,
                           ro_det_rec *d_det, int sync);

/* ---- stream / timing control --------------------------------------------- */

/* Run the batch paths on a caller-owned cudaStream_t (pass NULL to go back to
 * the handle's own stream). */
int  ro_wf_set_stream(ro_wf *h, void *cuda_stream);
int  ro_wf_sync(ro_wf *h);

/* When enabled every kernel launch is bracketed by CUDA events on its stream. */
int  ro_wf_set_profiling(ro_wf *h, int enable);
/* Sum of device time (ms) and launch counts since the last call, per kernel
 * family: [0] fused FFT kernel, [1] detector kernel.  Synchronises. */
int  ro_wf_kernel_times(ro_wf *h, double ms_sum[2], uint64_t launches[2]);
/* total kernel launches issued by this handle since creation */
uint64_t ro_wf_launch_count(const ro_wf *h);
/* which kernel the handle's last transform launch ran (a static string: diagnostics, bench labels) */
const char *ro_wf_kernel_name(const ro_wf *h);

/* ---- device memory helpers (so callers need no CUDA binding) -------------- */
int  ro_dev_count(int *n);
int  ro_dev_alloc(int device, size_t bytes, void **d_ptr);
int  ro_dev_free(int device, void *d_ptr);
int  ro_dev_upload(int device, void *d_dst, const void *src, size_t bytes);
int  ro_dev_download(int device, void *dst, const void *d_src, size_t bytes);
int  ro_dev_memset(int device, void *d_dst, int value, size_t bytes);
int  ro_host_alloc_pinned(size_t bytes, void **ptr);
int  ro_host_free_pinned(void *ptr);

/* ---- host-side tables (no GPU needed) ------------------------------------- */

/* The 4-term Nuttall table exactly as FFTBackend::startStream builds it
 * (src/FFTBackend.cpp:155-186). */
int  ro_window_table(int32_t bins, float *out);
/* exp(-2*pi*i*j/n), j in [0,n), rounded once from long double; interleaved re,im */
int  ro_twiddle_table(int32_t n, float *out);

const char *ro_last_error(void);
const char *ro_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RO_CUDA_H */
