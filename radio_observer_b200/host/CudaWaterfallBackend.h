/*
 * CudaWaterfallBackend.h -- drop-in Backend for radio-observer's Pipeline.
 *
 * Compiled AGAINST the reference headers (src/WaterfallBackend.h etc.), never copied from them, and
 * through their public and protected interfaces only: no reference file is edited, no `friend` is
 * added, no access specifier is redefined.
 *
 * The class derives from the reference's WaterfallBackend (so it IS the backend every Recorder's
 * Ref<WaterfallBackend> expects: getFFTSampleRate(), frequencyToBin(), getOrigin(), getMetadataFile()
 * ... keep working) but owns the data path below it:
 *
 *   FFTBackend::process              src/FFTBackend.cpp:192-279   framing, IQ gain, window, FFT
 *   WaterfallBackend::processFFT     src/WaterfallBackend.cpp:485-541  magnitude, fftshift, ring push,
 *                                                                 handle entry, recorder fan-out
 *   BolidRecorder::noise/peak/average src/BolidRecorder.cpp:302-347 per-row detector statistics
 *
 * run on the GPU behind the C ABI of include/ro_cuda.h; what comes back per frame is the union of the
 * recorders' bands (about 0.5 KB instead of the 64 KB row) and one (noise, peak, average) record.
 * WaterfallBackend keeps its row ring, handle table and recorder list private, so this class has its
 * own (same types, same index semantics: RingBuffer2D::mark() is the NEXT row, the handle of row r
 * sits in slot r + 1, src/WaterfallBackend.cpp:507) and hands THEM to the recorders with the public
 * Recorder::setBuffer() (src/WaterfallBackend.h:65-74).  The ring rows are bins_ floats wide like the
 * reference's, but only the band columns are ever written, so the untouched pages of its chunks are
 * never committed; the recorders index rows by absolute bin exactly as they do in the reference.
 *
 * The recorders the configuration injects ("snapshot", "bolid": src/WaterfallBackend.cpp:436-458,
 * src/BolidRecorder.cpp:360-414) are replaced, at injection, by subclasses of the same classes built
 * from the same parameters (CudaRecorders.h): same files, names, headers, CSV lines and messages --
 * the reference's code writes them -- plus what the reference lacks: a record of which snapshots
 * are queued (back-pressure, below), the detector statistics taken from the GPU record, and raw-I/Q
 * snapshots cut from this backend's own sample store instead of a per-sample ring push.
 *
 * Selecting it is one string in the JSON config:  "factory": "waterfall_cuda"
 * (same keys as "waterfall", src/WaterfallBackend.cpp:620-644, plus the optional keys
 *   cuda_device (0), cuda_batch_frames (64; 1 = every row is with the recorders before process()
 *   returns), cuda_precision (0 = packed FP32 transform, 1 = FP64 like the reference's FFTW path),
 *   cuda_rows ("bands" | "full": "full" brings whole rows back and lets the reference's own
 *   BolidRecorder::update() recompute the statistics on the host -- the bit-for-bit comparison mode),
 *   cuda_debug_write_delay_ms (0; a test hook that slows the FITS writers down)).
 *
 * Narrow ingest: processRaw() takes the interleaved int16 I/Q of a WAV block as it
 * lies in the file (4 B per sample instead of a 16 B Complex, src/WAVStream.cpp:115-121) with the
 * same DataInfo Frontend::process would pass; NarrowWAVStream.h is the frontend that calls it.
 */
#ifndef CUDA_WATERFALL_BACKEND_H
#define CUDA_WATERFALL_BACKEND_H

#include <stdint.h>
#include <deque>
#include <vector>

#include "WaterfallBackend.h"
#include "BolidRecorder.h"
#include "ro_cuda.h"

class CudaSnapshotRecorder;
class CudaBolidRecorder;

/// A Recorder that reads nothing but the row count and the handle table (a diagnostic tap): its
/// presence does not force whole rows back from the GPU.
class RowMetaTap : public Recorder {
public:
	RowMetaTap(Ref<WaterfallBackend> backend) : Recorder(backend) {}
};

class CudaWaterfallBackend : public WaterfallBackend {
public:
	/// Columns of the device row the host receives: [lo, hi), absolute bins.
	struct Band { int lo, hi; };

private:
	CudaWaterfallBackend(const CudaWaterfallBackend &other);

	struct PendingFrame {      // what FFTBackend::process hands to processFFT per frame
		DataInfo info;
		int      rawMark;
	};

	int     overlap_;          // clamped like FFTBackend::FFTBackend (src/FFTBackend.cpp:108-109)
	int     device_;
	int     batchFrames_;      // frames submitted per kernel launch; 1 = lowest latency
	int     precision_;
	bool    wantFullRows_;     // cuda_rows = "full"
	int     chunkSize_;        // buffer_chunk_size (the base class keeps its copy private)
	int     debugWriteDelayMs_; // cuda_debug_write_delay_ms: test hook, every FITS write sleeps first

	ro_wf  *wf_;
	bool    failed_;

	// ---- what WaterfallBackend keeps private, owned here (same types) ----
	FFTBuffer                    rows_;
	std::vector<RawDataHandle>   handles_;
	Mutex                        ringMutex_;
	std::vector<Ref<Recorder> >  recs_;
	std::vector<Ref<DIObject> >  retired_;    // the injected stock recorders (kept alive: they registered listeners)
	std::vector<CudaBolidRecorder *> bolids_;
	bool    fullRows_;         // whole rows come back (asked for, or a recorder of unknown type is attached)
	bool    gpuDetector_;      // the first bolid recorder takes (noise, peak, average) from the GPU record
	bool    refRawRing_;       // a recorder of unknown type may read FFTBackend::rawBuffer_: keep it filled
	Band    band_;

	// ---- host mirror of FFTBackend's private framing state (window_ fill level and windowRaw_,
	// src/FFTBackend.h:75-80): only the bookkeeping, no samples.  The reference stores a RawDataHandle
	// per sample (src/FFTBackend.cpp:217-225) but reads only windowRaw_[0] of a full window, before
	// and after the shift (:241-251), i.e. the handles of the samples whose stream index is a multiple
	// of the hop: hopHandles_ keeps exactly those, computed the same way.
	int                          fill_;
	int                          toNextHop_;   // samples until the next multiple of the hop
	std::deque<RawDataHandle>    hopHandles_;  // [0] = handle of the first sample of the oldest unfinished frame
	DataInfo                     info_;
	std::deque<PendingFrame>     pending_;

	std::vector<float>           harvestRows_; // harvested rows, rowWidth_ floats each
	std::vector<ro_det_rec>      harvestDet_;
	int                          rowWidth_;
	ro_det_rec                   currentDet_;  // record of the row the recorders are being updated for

	int     inFormat_;         // RO_IN_* the handle was created for
	long    samplesIn_;        // samples accepted since startStream
	long    delivered_;        // rows handed to the recorders
	long    rowsAtLastLog_;

	// ---- raw I/Q store: the samples as they arrived, ring of rawCapacity_ samples with
	// the index semantics of FFTBackend::rawBuffer_ (sample g sits at g % capacity; the mark after its
	// push is (g + 1) % capacity, src/RingBuffer.h:482-509) ----
	unsigned char *rawStore_;
	long    rawCapacity_;      // samples
	int     rawBytes_;         // 4: int16 pairs, 8: float pairs (what floatToInt(Complex, float*) keeps)
	long    rawPushed_;

	// ---- back-pressure.  The reference's recorders write their files from worker threads that read
	// the row ring, its handle table and the raw samples LATER (src/WaterfallBackend.cpp:60-108,
	// :141-258); a reservation only flags an overwritten region (src/RingBuffer.h:489-493), nothing
	// stops the producer.  At the reference's speed the writers keep up; this backend consumes a WAV
	// orders of magnitude faster and would lap the ring (8 snapshot lengths).  So every queued
	// snapshot is noted with the absolute row it starts at (by the recorder subclasses) until its
	// files are written, and the producer waits before it overwrites what the oldest one needs. ----
	struct Live { const void *who; int start; long absRow; };
	Mutex               liveMutex_;
	std::vector<Live>   live_;
	long oldestLiveRow();
	void waitForWriters(long firstNeededRow);
	void waitBeforeRawPush(int count);

	// ---- the reference's three timing counters (src/FFTBackend.h:86-92,208-235), kept alive ----
	RunningAverage2<double> procTime_, fftTime_, analTime_;
	Stopwatch               procWatch_;

	bool createHandle(int inFormat);
	bool ensureFormat(int inFormat);
	void storeRaw(const Complex *src, int count);
	void storeRaw(const int16_t *src, int count);
	template<class S> void pushSamples(const S *src, int count, WFTime timeOffset);
	template<class S> void processSamples(const S *src, int size, DataInfo info);
	void harvest(bool flush, int atMost = 0);
	void fail(const char *what);
	void attach(Ref<Recorder> recorder);

public:
	CudaWaterfallBackend(int bins, int overlap, std::string origin, int device, int batchFrames,
	                     int precision, bool fullRows, int chunkSize);
	virtual ~CudaWaterfallBackend();

	virtual void startStream(StreamInfo info);
	virtual void process(const std::vector<Complex> &data, DataInfo info);
	virtual void endStream();
	virtual bool injectDependency(Ref<DIObject> obj, std::string key);

	/// Backend::process for samples still in the WAV file's encoding: `count` complex samples as
	/// interleaved int16 [I, Q], unscaled (what src/WAVStream.cpp:115-121 widens to Complex).
	/// Same frames, rows, handles and recorder calls as process() on the widened block.
	void processRaw(const int16_t *iq, int count, DataInfo info);

	/// Attaches a recorder that did not come through the Injector (what WaterfallBackend::addRecorder
	/// does for the base class, src/WaterfallBackend.cpp:563-567).  Anything but a RowMetaTap makes
	/// whole rows come back from the GPU.
	void addTap(Ref<Recorder> recorder);

	/// number of rows delivered to the recorders so far
	long deliveredRows() const { return delivered_; }
	bool deliversFullRows() const { return fullRows_; }
	Band deliveredBand() const { return band_; }

	// ---- for the recorder subclasses ----
	bool gpuDetector() const { return gpuDetector_; }
	const ro_det_rec &currentRecord() const { return currentDet_; }
	/// absolute row of ring index `mark`, seen from the row that has just been pushed
	long absoluteRow(int mark);
	void noteQueued(const void *who, int start);
	void noteWritten(const void *who, int start);
	/// `length` samples from raw-ring index `start` (wrapped) as float pairs, the way writeRaw reads them
	void copyRaw(int start, int length, float *dst);
	/// capacity of this backend's own sample store; 0 while the reference's raw ring is in use instead
	long rawCapacity() const { return refRawRing_ ? 0 : rawCapacity_; }
	int  debugWriteDelayMs() const { return debugWriteDelayMs_; }
	void setDebugWriteDelayMs(int ms) { debugWriteDelayMs_ = ms > 0 ? ms : 0; }
	/// the DEBUG line of FFTBackend::logProcessingTimes / clearProcessingTime with this backend's counters
	void logProcessingTimes();
	void clearProcessingTime();

	static Ref<DIObject> make(Ref<DynObject> config, Ref<DIObject> parent);
};

#endif /* CUDA_WATERFALL_BACKEND_H */
