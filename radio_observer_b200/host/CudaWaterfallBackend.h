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
 *
 * Narrow ingest: processRaw() takes the interleaved int16 I/Q of a WAV block as it
 * lies in the file (4 B per sample instead of a 16 B Complex, src/WAVStream.cpp:115-121) with the
 * same DataInfo Frontend::process would pass; NarrowWAVStream.h is the frontend that calls it.
 */
#ifndef CUDA_WATERFALL_BACKEND_H
#define CUDA_WATERFALL_BACKEND_H

#include <deque>
#include <vector>

// Everything WaterfallBackend.h includes, first and untouched ...
#include <cmath>
#include <cfloat>
#include <limits>
#include <fitsio.h>
#include <fftw3.h>
#include "Backend.h"
// RingBuffer2D with its private section readable as well: the narrow-ingest path (processRaw) fills
// the raw I/Q ring a run of samples at a time and then moves the head by the run length, which needs
// head_ / size_ / chunkRows_ (src/RingBuffer.h:222-306); the per-sample push() it replaces costs an
// integer division per sample (src/RingBuffer.h:289-290).  No reservation is ever made on the raw
// ring (only on the row ring, src/WaterfallBackend.cpp:118), so push()'s reservation scan has
// nothing to mark there.
#define private public
#include "RingBuffer.h"
#undef private
#include "FFTBackend.h"
#include "FITSWriter.h"
#include "Channel.h"
#include "BolidMessage.h"
#include "CsvLog.h"
#include "utils.h"
// ... then the class itself with its private section readable from this translation unit: the row
// ring, its handle table and the recorder list are private members (src/WaterfallBackend.h:243-252)
// and a finished GPU row has to be pushed into that very ring (what processFFT does after its own
// magnitude loop, src/WaterfallBackend.cpp:488,507,534-536).  No reference file is edited and the class
// layout is unchanged; a maintainer who prefers it adds `friend class CudaWaterfallBackend;` instead.
#define private public
#include "WaterfallBackend.h"
#undef private
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
	// windowRaw_, src/FFTBackend.h:75-80): only the bookkeeping, no samples.  The reference stores a
	// RawDataHandle per sample (src/FFTBackend.cpp:217-225) but reads only windowRaw_[0] of a full
	// window, before and after the shift (:241-251), i.e. the handles of the samples whose stream index
	// is a multiple of the hop: hopHandles_ keeps exactly those, computed the same way.
	int                          fill_;
	int                          toNextHop_;   // samples until the next multiple of the hop
	std::deque<RawDataHandle>    hopHandles_;  // [0] = handle of the first sample of the oldest unfinished frame
	DataInfo                     info_;
	std::deque<PendingFrame>     pending_;

	std::vector<float>           rows_;      // harvested rows, bins_ floats each

	int     inFormat_;         // RO_IN_* the handle was created for
	long    samplesIn_;        // samples accepted since startStream
	long    rawPushed_;        // samples pushed to the raw I/Q ring since startStream

	// Back-pressure.  The reference's recorders write their files from worker threads that read the
	// row ring, its handle table and the raw ring LATER (src/WaterfallBackend.cpp:60-108, :141-258);
	// a reservation only flags an overwritten region (src/RingBuffer.h:489-493), nothing stops the
	// producer.  At the reference's own speed the writers keep up; behind this backend a WAV is
	// consumed several times faster, the producer laps the ring (8 snapshot lengths) and the writer
	// stores rows, times and file names of one ring capacity later.  So the producer waits before it
	// overwrites what a queued snapshot still needs: per live reservation the absolute row its
	// snapshot starts at, noted when the reservation appears.
	std::vector<long>            resAbs_;    // by reservation handle; -1 = none noted
	std::vector<int>             resStart_;  // the ring index it was noted for
	void noteReservations();
	long oldestReservedRow();
	void waitForWriters(long firstNeededRow);

	bool createHandle(int inFormat);
	bool ensureFormat(int inFormat);
	void pushSamples(const Complex *src, int count, WFTime timeOffset);
	void pushSamples(const int16_t *src, int count, WFTime timeOffset);
	template<class S> void processSamples(const S *src, int size, DataInfo info);
	void harvest(bool flush, int atMost = 0);
	void waitBeforeRawPush(int count);
	void fail(const char *what);

public:
	CudaWaterfallBackend(int bins, int overlap, std::string origin, int device, int batchFrames, int precision = 0);
	virtual ~CudaWaterfallBackend();


	virtual void startStream(StreamInfo info);
	virtual void process(const std::vector<Complex> &data, DataInfo info);
	virtual void endStream();

	/// Backend::process for samples still in the WAV file's encoding: `count` complex samples as
	/// interleaved int16 [I, Q], unscaled (what src/WAVStream.cpp:115-121 widens to Complex).
	/// Same frames, rows, handles and recorder calls as process() on the widened block.
	void processRaw(const int16_t *iq, int count, DataInfo info);

	/// number of rows delivered to the recorders so far
	long deliveredRows() const { return delivered_; }

	static Ref<DIObject> make(Ref<DynObject> config, Ref<DIObject> parent);

private:
	long delivered_;
};

#endif /* CUDA_WATERFALL_BACKEND_H */
