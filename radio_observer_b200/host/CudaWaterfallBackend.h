/*
 * CudaWaterfallBackend.h -- drop-in Backend for radio-observer's Pipeline.
 *
 * Compiled AGAINST the reference headers (src/WaterfallBackend.h etc.), never
 * copied from them.  The class derives from the reference's WaterfallBackend so
 * that its row ring, handle table, recorder fan-out, FITS/CSV writers and every
 * configured Recorder ("snapshot", "bolid") keep running unmodified; only what
 * FFTBackend::process computes (reference src/FFTBackend.cpp:192-279: framing,
 * IQ gain, Nuttall window, N-point FFT) and the magnitude of
 * WaterfallBackend::processFFT (src/WaterfallBackend.cpp:492-505) move to the
 * GPU through the C ABI of include/ro_cuda.h.
 *
 * Selecting it is one string in the JSON config:  "factory": "waterfall_cuda"
 * (same keys as "waterfall", src/WaterfallBackend.cpp:620-644, plus the optional
 * keys cuda_device (0), cuda_batch_frames (1) and cuda_precision (0 = FP32, 1 = FP64)).
 */
#ifndef CUDA_WATERFALL_BACKEND_H
#define CUDA_WATERFALL_BACKEND_H

#include <deque>
#include <vector>

#include "WaterfallBackend.h"
#include "ro_cuda.h"

class CudaWaterfallBackend : public WaterfallBackend {
private:
	CudaWaterfallBackend(const CudaWaterfallBackend &other);

	struct PendingFrame {      // what FFTBackend::process hands to processFFT per frame
		DataInfo info;
		int      rawMark;
	};

	int     overlap_;          // clamped like FFTBackend::FFTBackend (src/FFTBackend.cpp:108-109)
	int     device_;
	int     batchFrames_;      // frames submitted per kernel launch; 1 = lowest latency
	int     precision_;        // 0 = packed FP32 transform, 1 = FP64 like the reference's FFTW path
	double  iqGain_;
	int     iqPhaseShift_;

	ro_wf  *wf_;
	bool    failed_;

	// host mirror of FFTBackend's private framing state (window_ fill level and
	// windowRaw_, src/FFTBackend.h:75-80): only the bookkeeping, no samples
	int                          fill_;
	std::vector<RawDataHandle>   windowRaw_;
	DataInfo                     info_;
	std::deque<PendingFrame>     pending_;

	std::vector<float>           rows_;      // harvested rows, bins_ floats each
	std::vector<double>          asComplex_; // a row re-expressed as (magnitude, 0) per FFT bin (fftw_complex layout)

	void pushSamples(const Complex *src, int count, WFTime timeOffset);
	void harvest(bool flush);
	void fail(const char *what);

public:
	CudaWaterfallBackend(int bins, int overlap, std::string origin, int device, int batchFrames, int precision = 0);
	virtual ~CudaWaterfallBackend();


	virtual void startStream(StreamInfo info);
	virtual void process(const std::vector<Complex> &data, DataInfo info);
	virtual void endStream();

	/// number of rows delivered to the recorders so far
	long deliveredRows() const { return delivered_; }

	static Ref<DIObject> make(Ref<DynObject> config, Ref<DIObject> parent);

private:
	long delivered_;
};

#endif /* CUDA_WATERFALL_BACKEND_H */
