/*
 * CudaWaterfallBackend.cpp -- see CudaWaterfallBackend.h.
 *
 * What stays on the host, restated from FFTBackend::process (reference src/FFTBackend.cpp:192-279)
 * because the recorders depend on it:
 *   - every incoming sample has a slot in the raw I/Q ring and a RawDataHandle {ring mark after its
 *     push, stream time}                                            (:217-223, :264-270)
 *   - a frame is complete when the window holds bins_ samples; its row time is the handle time of
 *     the window's first sample                                      (:225)
 *   - the handle table is shifted by the hop BEFORE processFFT is called, so the rawMark handed
 *     over is the one of the NEXT frame's first sample whenever overlap > 0   (:241-251)
 * and from WaterfallBackend::processFFT (src/WaterfallBackend.cpp:485-541): push a row, store the
 * handle at the ring's mark (the slot AFTER the row), call every recorder's update().
 * What moves to the GPU: IQ gain, window, FFT, magnitude, fftshift, the bolid statistics.
 */
#include "CudaWaterfallBackend.h"
#include "CudaRecorders.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <climits>
#include <typeinfo>
#include <unistd.h>


CudaWaterfallBackend::CudaWaterfallBackend(int bins, int overlap, std::string origin, int device,
                                           int batchFrames, int precision, bool fullRows, int chunkSize) :
	WaterfallBackend(bins, overlap, origin),
	overlap_(overlap),
	device_(device),
	batchFrames_(batchFrames < 1 ? 1 : batchFrames),
	precision_(precision),
	wantFullRows_(fullRows),
	chunkSize_(chunkSize),
	debugWriteDelayMs_(0),
	wf_(NULL), failed_(false),
	fullRows_(fullRows), gpuDetector_(false), refRawRing_(false),
	fill_(0), toNextHop_(0),
	rowWidth_(bins),
	inFormat_(RO_IN_COMPLEX64), samplesIn_(0), delivered_(0), rowsAtLastLog_(0),
	rawStore_(NULL), rawCapacity_(0), rawBytes_(8), rawPushed_(0)
{
	if (overlap_ < 0) overlap_ = 0;
	if (overlap_ >= bins_) overlap_ = bins_ - 1;
	band_.lo = 0;
	band_.hi = bins_;
	memset(&currentDet_, 0, sizeof currentDet_);
}


CudaWaterfallBackend::~CudaWaterfallBackend()
{
	if (wf_) ro_wf_destroy(wf_);
	wf_ = NULL;
	free(rawStore_);
	rawStore_ = NULL;
}


void CudaWaterfallBackend::fail(const char *what)
{
	// the reference's convention on the data path: log and stop producing
	LOG_ERROR("CudaWaterfallBackend: " << what << ": " << ro_last_error());
	failed_ = true;
}


////////////////////////////////////////////////////////////////////////////////
// recorders
////////////////////////////////////////////////////////////////////////////////

void CudaWaterfallBackend::attach(Ref<Recorder> recorder)
{
	// WaterfallBackend::addRecorder (src/WaterfallBackend.cpp:563-567) on this class's own ring
	recs_.push_back(recorder);
	recorder->setBuffer(&rows_, &rawBuffer_, &ringMutex_, &handles_);
}


bool CudaWaterfallBackend::injectDependency(Ref<DIObject> obj, std::string key)
{
	if (key.compare("recorder") != 0) return WaterfallBackend::injectDependency(obj, key);

	// WaterfallBackend::injectDependency (src/WaterfallBackend.cpp:610-617) would hand the recorder its
	// private ring.  The two recorder classes the reference ships are rebuilt as their GPU-aware
	// subclasses from the same parameters; the stock objects stay alive (a snapshot recorder that
	// listens to NoiseMessage registered `this` with the dispatcher and never unregisters).
	DIObject *raw = obj.getPtr();
	if (raw && typeid(*raw) == typeid(BolidRecorder)) {
		Ref<CudaBolidRecorder> r = CudaBolidRecorder::like(*static_cast<BolidRecorder *>(raw), this);
		bolids_.push_back(r.getPtr());
		retired_.push_back(obj);
		attach(r);
	} else if (raw && typeid(*raw) == typeid(SnapshotRecorder)) {
		Ref<CudaSnapshotRecorder> r = CudaSnapshotRecorder::like(*static_cast<SnapshotRecorder *>(raw), this);
		retired_.push_back(obj);
		attach(r);
	} else {
		addTap(obj.as<Recorder>());
	}
	return true;
}


void CudaWaterfallBackend::addTap(Ref<Recorder> recorder)
{
	if (recorder.isNull()) return;
	if (dynamic_cast<RowMetaTap *>(recorder.getPtr()) == NULL) {
		// a recorder this backend knows nothing about: it may read any bin of a row and the reference's
		// raw ring, so both are kept the way the reference keeps them
		fullRows_ = true;
		refRawRing_ = true;
	}
	attach(recorder);
}


////////////////////////////////////////////////////////////////////////////////
// stream start / end
////////////////////////////////////////////////////////////////////////////////

void CudaWaterfallBackend::startStream(StreamInfo info)
{
	// WaterfallBackend::startStream (src/WaterfallBackend.cpp:573-594) on this class's own ring:
	// FFTBackend::startStream sets streamInfo_ and fftSampleRate_ (src/FFTBackend.cpp:144-151)
	FFTBackend::startStream(info);

	int bufferSize = 1;
	for (size_t i = 0; i < recs_.size(); i++) {
		int requested = recs_[i]->requestBufferSize();
		if (requested > bufferSize) bufferSize = requested;
	}
	// rows are bins_ floats wide as in the reference; only the band columns are ever written, the rest
	// of a chunk stays untouched virtual memory
	rows_.resize(getBins(), chunkSize_, bufferSize);
	handles_.assign(rows_.getCapacity(), RawDataHandle());

	// the raw I/Q ring of the reference holds fftSamplesToRaw(bufferSize) samples rounded up to whole
	// 1 MiB chunks of float pairs (src/FFTBackend.h:129-132, src/RingBuffer.h:428-448)
	{
		const long want = fftSamplesToRaw(bufferSize);
		const long chunkRows = (1024 * 1024) / (2 * sizeof(float));
		long chunks = want / chunkRows + ((want % chunkRows) > 0 ? 1 : 0);
		rawCapacity_ = chunks * chunkRows;
	}
	if (refRawRing_) resizeRawBuffer(fftSamplesToRaw(bufferSize));

	for (size_t i = 0; i < recs_.size(); i++) recs_[i]->start();

	fill_ = 0;
	info_ = DataInfo();
	pending_.clear();
	toNextHop_ = 0;
	hopHandles_.clear();
	delivered_ = 0;
	rowsAtLastLog_ = 0;
	failed_ = false;
	samplesIn_ = 0;
	rawPushed_ = 0;
	{
		MutexLock lock(&liveMutex_);
		live_.clear();
	}
	clearProcessingTime();

	// the columns the recorders read: every snapshot band, and for a bolid recorder its noise band, its
	// detect band and the mean's window, which may reach avg/2 bins below and avg bins above the detect band
	fullRows_ = wantFullRows_ || refRawRing_;
	gpuDetector_ = false;
	int lo = bins_, hi = 0;
	for (size_t i = 0; i < recs_.size() && !fullRows_; i++) {
		Recorder *r = recs_[i].getPtr();
		if (CudaBolidRecorder *b = dynamic_cast<CudaBolidRecorder *>(r)) {
			lo = std::min(lo, std::min(b->leftBin(), std::min(b->lowNoiseBin(), b->lowDetectBin() - b->averageBinRange() / 2)));
			hi = std::max(hi, std::max(b->rightBin(), std::max(b->lowNoiseBin() + b->noiseWidth(),
			                                                   b->lowDetectBin() + b->detectWidth() + b->averageBinRange())));
		} else if (CudaSnapshotRecorder *s = dynamic_cast<CudaSnapshotRecorder *>(r)) {
			lo = std::min(lo, s->leftBin());
			hi = std::max(hi, s->rightBin());
		} else if (dynamic_cast<RowMetaTap *>(r) == NULL) {
			fullRows_ = true;
		}
	}
	if (lo < 0) lo = 0;
	if (hi > bins_) hi = bins_;
	if (fullRows_ || lo >= hi) {
		band_.lo = 0;
		band_.hi = bins_;
	} else {
		band_.lo = lo;
		band_.hi = hi;
	}
	// the first bolid recorder's statistics come from the GPU (one detector per handle); a second one
	// and the "full" mode run the reference's own update() on the host row
	for (size_t i = 0; i < bolids_.size(); i++) bolids_[i]->useGpuRecords(false);
	if (!fullRows_ && !bolids_.empty() && bolids_[0]->detectWidth() > 0 && bolids_[0]->noiseWidth() > 0) {
		gpuDetector_ = true;
		bolids_[0]->useGpuRecords(true);
	}

	// the handle is created by the first process() / processRaw(): the Pipeline's frontends hand over
	// vector<Complex>, a narrow frontend int16, and the device work space depends on the encoding
	if (wf_) { ro_wf_destroy(wf_); wf_ = NULL; }
	inFormat_ = -1;
}


bool CudaWaterfallBackend::createHandle(int inFormat)
{
	if (wf_) { ro_wf_destroy(wf_); wf_ = NULL; }
	ro_wf_config cfg;
	memset(&cfg, 0, sizeof cfg);
	cfg.bins = bins_;
	cfg.overlap = overlap_;
	cfg.sample_rate = streamInfo_.sampleRate;
	cfg.in_format = inFormat;
	cfg.iq_gain = getGain();
	cfg.iq_phase_shift = getPhaseShift();
	if (band_.lo > 0 || band_.hi < bins_) {
		cfg.n_bands = 1;
		cfg.band_lo[0] = band_.lo;
		cfg.band_hi[0] = band_.hi;
	}
	if (gpuDetector_) {
		const CudaBolidRecorder *b = bolids_[0];
		cfg.det_lo = b->lowDetectBin();
		cfg.det_w = b->detectWidth();
		cfg.noise_lo = b->lowNoiseBin();
		cfg.noise_w = b->noiseWidth();
		cfg.avg_bins = b->averageBinRange();
	}
	cfg.device = device_;
	cfg.max_batch_frames = batchFrames_;
	cfg.precision = precision_;
	inFormat_ = inFormat;
	rawBytes_ = inFormat == RO_IN_INT16 ? 4 : 8;
	if (!refRawRing_) {
		// zero pages, committed only as samples arrive (the reference's ring is uninitialised memory)
		free(rawStore_);
		rawStore_ = (unsigned char *)calloc((size_t)rawCapacity_, (size_t)rawBytes_);
		if (!rawStore_) { LOG_ERROR("CudaWaterfallBackend: out of memory for the raw I/Q store"); failed_ = true; return false; }
	}
	if (ro_wf_create(&cfg, &wf_) != RO_OK) {
		wf_ = NULL;
		fail("ro_wf_create");
		return false;
	}
	ro_wf_set_profiling(wf_, 1);      // kernel times for the reference's "fft time" / "anal. time" line
	rowWidth_ = ro_wf_row_width(wf_);
	{
		size_t cap = ((size_t)1 << 22) / (size_t)rowWidth_;          // about 16 MiB of harvested rows at a time
		if (cap < 1) cap = 1;
		if (cap > 4096) cap = 4096;
		harvestRows_.resize((size_t)rowWidth_ * cap);
		harvestDet_.resize(cap);
	}
	return true;
}


bool CudaWaterfallBackend::ensureFormat(int inFormat)
{
	if (inFormat_ == inFormat && wf_) return true;
	if (samplesIn_ > 0) {
		// one stream, one encoding: the device-side sample ring holds what was pushed so far
		LOG_ERROR("CudaWaterfallBackend: process() and processRaw() mixed within one stream");
		failed_ = true;
		return false;
	}
	return createHandle(inFormat);
}


void CudaWaterfallBackend::endStream()
{
	// complete frames still in flight are delivered; the incomplete tail is dropped exactly as the
	// reference never flushes window_ (src/FFTBackend.cpp:282-286).  Then WaterfallBackend::endStream
	// (src/WaterfallBackend.cpp:600-607) on this class's recorder list.
	if (wf_ && !failed_) harvest(true);
	FFTBackend::endStream();
	for (size_t i = 0; i < recs_.size(); i++) recs_[i]->stop();
	if (wf_) { ro_wf_destroy(wf_); wf_ = NULL; }
}


////////////////////////////////////////////////////////////////////////////////
// samples in
////////////////////////////////////////////////////////////////////////////////

void CudaWaterfallBackend::storeRaw(const Complex *src, int count)
{
	if (refRawRing_) {
		for (int i = 0; i < count; i++) floatToInt(src[i], rawBuffer_.push());     // src/FFTBackend.cpp:218
		return;
	}
	// (float)re, (float)im: what floatToInt(Complex, float*) stores (src/FFTBackend.h:250-254)
	float *ring = reinterpret_cast<float *>(rawStore_);
	long at = rawPushed_ % rawCapacity_;
	for (int i = 0; i < count; i++) {
		ring[2 * at] = (float)src[i].real;
		ring[2 * at + 1] = (float)src[i].imag;
		if (++at == rawCapacity_) at = 0;
	}
}


void CudaWaterfallBackend::storeRaw(const int16_t *src, int count)
{
	if (refRawRing_) {
		for (int i = 0; i < count; i++) {
			float *dst = rawBuffer_.push();
			dst[0] = (float)src[2 * i];
			dst[1] = (float)src[2 * i + 1];
		}
		return;
	}
	// the int16 pairs as they are; (float)(double)int16 == (float)int16 is formed when a snapshot is cut
	int16_t *ring = reinterpret_cast<int16_t *>(rawStore_);
	long at = rawPushed_ % rawCapacity_;
	int done = 0;
	while (done < count) {
		long run = rawCapacity_ - at;
		if (run > count - done) run = count - done;
		memcpy(ring + 2 * at, src + 2 * (size_t)done, (size_t)run * 4);
		done += (int)run;
		at = (at + run) % rawCapacity_;
	}
}


void CudaWaterfallBackend::copyRaw(int start, int length, float *dst)
{
	if (rawCapacity_ <= 0 || refRawRing_) return;
	long at = ((long)start % rawCapacity_ + rawCapacity_) % rawCapacity_;
	for (int y = 0; y < length; y++) {
		if (rawBytes_ == 4) {
			const int16_t *s = reinterpret_cast<const int16_t *>(rawStore_) + 2 * at;
			dst[2 * (size_t)y] = (float)s[0];
			dst[2 * (size_t)y + 1] = (float)s[1];
		} else {
			const float *s = reinterpret_cast<const float *>(rawStore_) + 2 * at;
			dst[2 * (size_t)y] = s[0];
			dst[2 * (size_t)y + 1] = s[1];
		}
		if (++at == rawCapacity_) at = 0;
	}
}


template<class S>
void CudaWaterfallBackend::pushSamples(const S *src, int count, WFTime timeOffset)
{
	// Handles: sample i of this run starts a frame when i = toNextHop_ (mod hop); the ring mark after
	// its push is (samples pushed so far + i + 1) mod capacity (src/RingBuffer.h:482-509).
	const int hop = bins_ - overlap_;
	const long cap = refRawRing_ ? rawBuffer_.getCapacity() : rawCapacity_;
	waitBeforeRawPush(count);
	{
		int i = toNextHop_;
		for (; i < count; i += hop) {
			hopHandles_.push_back(RawDataHandle(
				(int)((rawPushed_ + i + 1) % cap),
				timeOffset.addSamples(i, streamInfo_.sampleRate)
			));
		}
		toNextHop_ = i - count;
	}
	storeRaw(src, count);
	rawPushed_ += count;
	if (ro_wf_push(wf_, src, (size_t)count) != RO_OK) fail("ro_wf_push");
}


void CudaWaterfallBackend::process(const std::vector<Complex> &data, DataInfo info)
{
	if (failed_ || data.empty()) return;
	if (!ensureFormat(RO_IN_COMPLEX64)) return;
	procWatch_.start();
	processSamples(&(data[0]), (int)data.size(), info);
	procWatch_.end();
	procTime_.add(procWatch_.getMilliseconds());
}


void CudaWaterfallBackend::processRaw(const int16_t *iq, int count, DataInfo info)
{
	if (failed_ || count <= 0) return;
	if (!ensureFormat(RO_IN_INT16)) return;
	procWatch_.start();
	processSamples(iq, count, info);
	procWatch_.end();
	procTime_.add(procWatch_.getMilliseconds());
}


template<class S> static inline const S *advanceSamples(const S *p, int n);
template<> inline const Complex *advanceSamples<Complex>(const Complex *p, int n) { return p + n; }
template<> inline const int16_t *advanceSamples<int16_t>(const int16_t *p, int n) { return p + 2 * (size_t)n; }


template<class S>
void CudaWaterfallBackend::processSamples(const S *src, int size, DataInfo info)
{
	WFTime timeOffset = info.timeOffset;
	samplesIn_ += size;

	while (size >= (bins_ - fill_)) {
		int count = bins_ - fill_;
		pushSamples(src, count, timeOffset);
		if (failed_) return;

		// windowRaw_[0] of the full window, then of the shifted one (src/FFTBackend.cpp:241-251):
		// with overlap 0 nothing is shifted and both are the same handle
		info_.timeOffset = hopHandles_[0].time;
		PendingFrame p;
		p.info = info_;
		p.rawMark = overlap_ > 0 ? hopHandles_[1].mark : hopHandles_[0].mark;
		pending_.push_back(p);
		hopHandles_.pop_front();

		fill_ = overlap_;
		size -= count;
		src = advanceSamples<S>(src, count);

		timeOffset = timeOffset.addSamples(count, streamInfo_.sampleRate);
		info_.offset++;

		// realtime (batch 1): the row is with the recorders before process() returns.  Otherwise rows
		// are taken over one batch late, while the GPU works on the next one.
		if (batchFrames_ <= 1) harvest(true);
		else if ((int)pending_.size() >= 2 * batchFrames_) harvest(false, batchFrames_);
		if (failed_) return;
	}

	if (size > 0) {
		pushSamples(src, size, timeOffset);
		fill_ += size;
	}
}


////////////////////////////////////////////////////////////////////////////////
// rows out
////////////////////////////////////////////////////////////////////////////////

void CudaWaterfallBackend::harvest(bool flush, int atMost)
{
	if (flush && ro_wf_flush(wf_) != RO_OK) { fail("ro_wf_flush"); return; }
	const size_t cap = harvestDet_.size();
	const int bandWidth = band_.hi - band_.lo;
	size_t want = atMost > 0 ? (size_t)atMost : pending_.size();
	Stopwatch analysis;
	while (want > 0 && !pending_.empty()) {
		size_t n = 0;
		uint64_t first = 0;
		if (ro_wf_poll(wf_, &harvestRows_[0], cap < want ? cap : want, gpuDetector_ ? &harvestDet_[0] : NULL,
		               &n, &first) != RO_OK) { fail("ro_wf_poll"); return; }
		if (n == 0) break;
		analysis.start();
		for (size_t r = 0; r < n && !pending_.empty(); r++) {
			// Pushing absolute row d overwrites ring row d - capacity and the handle slot (d + 1) % capacity,
			// which is where a snapshot starting at row d + 1 - capacity reads its time and raw mark
			// (fftMarkToTime / fftMarkToRaw of its start, at write time): wait for starts <= d + 1 - capacity
			waitForWriters(delivered_ - rows_.getCapacity() + 2);
			float *row = rows_.push();
			memcpy(row + band_.lo, &harvestRows_[r * (size_t)rowWidth_], sizeof(float) * (size_t)bandWidth);
			PendingFrame p = pending_.front();
			pending_.pop_front();
			handles_[rows_.mark()] = RawDataHandle(p.rawMark, p.info.timeOffset);     // src/WaterfallBackend.cpp:507
			if (gpuDetector_) currentDet_ = harvestDet_[r];
			for (size_t i = 0; i < recs_.size(); i++) recs_[i]->update();            // :534-536
			delivered_++;
		}
		analysis.end();
		analTime_.add(analysis.getMilliseconds() / (double)n);
		want -= n < want ? n : want;
	}
}


long CudaWaterfallBackend::absoluteRow(int mark)
{
	// called from a recorder's update(): the row being announced is already in the ring, delivered_ not
	// yet counted.  rows pushed = delivered_ + 1, the ring's mark is their count mod capacity.
	const int cap = rows_.getCapacity();
	const int head = rows_.mark();
	const int start = ((mark % cap) + cap) % cap;
	const int back = ((head - start) % cap + cap) % cap;     // rows between the snapshot's first row and the head
	return delivered_ + 1 - back;                            // negative: a bolid right after the start of the stream
}


void CudaWaterfallBackend::noteQueued(const void *who, int start)
{
	Live l;
	l.who = who;
	l.start = start;
	l.absRow = absoluteRow(start);
	MutexLock lock(&liveMutex_);
	live_.push_back(l);
}


void CudaWaterfallBackend::noteWritten(const void *who, int start)
{
	MutexLock lock(&liveMutex_);
	for (size_t i = 0; i < live_.size(); i++) {
		if (live_[i].who == who && live_[i].start == start) {
			live_.erase(live_.begin() + i);
			return;
		}
	}
}


long CudaWaterfallBackend::oldestLiveRow()
{
	MutexLock lock(&liveMutex_);
	long oldest = LONG_MAX;
	for (size_t i = 0; i < live_.size(); i++)
		if (live_[i].absRow < oldest) oldest = live_[i].absRow;
	return oldest;
}


void CudaWaterfallBackend::waitForWriters(long firstNeededRow)
{
	// rows below firstNeededRow are about to be lost: wait until no queued snapshot starts there.
	// A snapshot that old is complete, so its writer does not depend on rows still to come; the
	// time limit only guards against a writer that died.
	if (firstNeededRow <= 0) {
		MutexLock lock(&liveMutex_);
		if (live_.empty()) return;
	}
	long waited = 0;
	while (oldestLiveRow() < firstNeededRow) {
		usleep(100);
		waited += 100;
		if (waited == 2L * 1000 * 1000)
			LOG_WARNING("CudaWaterfallBackend: waiting for a recorder to write the snapshot that starts at row "
			            << oldestLiveRow() << " (row " << delivered_ << " is next; rows below " << firstNeededRow
			            << " are about to be overwritten)");
		if (waited > 60L * 1000 * 1000) {
			LOG_WARNING("CudaWaterfallBackend: a recorder has not written a snapshot for 60 s; overwriting its rows");
			break;
		}
	}
}


void CudaWaterfallBackend::waitBeforeRawPush(int count)
{
	// the raw I/Q of a queued bolid snapshot starts at (or one hop after) the first sample of its
	// first row: sample row * hop
	const long hop = bins_ - overlap_;
	const long cap = refRawRing_ ? rawBuffer_.getCapacity() : rawCapacity_;
	const long firstNeededSample = rawPushed_ + count - cap;
	if (firstNeededSample <= 0) return;
	waitForWriters((firstNeededSample + hop - 1) / hop);
}


////////////////////////////////////////////////////////////////////////////////
// timing counters
////////////////////////////////////////////////////////////////////////////////

void CudaWaterfallBackend::logProcessingTimes()
{
	// FFTBackend::logProcessingTimes (src/FFTBackend.h:208-229) prints its private counters, which only
	// FFTBackend::process feeds; the same line from this backend's own: wall time per process() call,
	// CUDA-event time of the transform kernels per frame, and per row the detector kernel plus the
	// recorders' update() on the host
	double ms[2] = {0, 0};
	uint64_t launches[2] = {0, 0};
	const long rows = delivered_ - rowsAtLastLog_;
	if (wf_ && rows > 0 && ro_wf_kernel_times(wf_, ms, launches) == RO_OK) {
		fftTime_.add(ms[0] / (double)rows);
		if (launches[1]) analTime_.add(ms[1] / (double)rows);
		rowsAtLastLog_ = delivered_;
	}
	LOG_DEBUG("FFTBackend: avg. proc. time (ms) = "
	          << procTime_.getValue()
	          << ", max. proc. time (ms) = "
	          << procTime_.max
	          << ", call count = "
	          << procTime_.count
	          << ", avg. fft time (ms) = "
	          << fftTime_.getValue()
	          << ", max. fft time (ms) = "
	          << fftTime_.max
	          << ", fft count = "
	          << fftTime_.count
	          << ", avg. anal. time (ms) = "
	          << analTime_.getValue()
	          << ", max. anal. time (ms) = "
	          << analTime_.max
	);
}


void CudaWaterfallBackend::clearProcessingTime()
{
	procTime_.clear();
	fftTime_.clear();
	analTime_.clear();
}


////////////////////////////////////////////////////////////////////////////////
// factory
////////////////////////////////////////////////////////////////////////////////

Ref<DIObject> CudaWaterfallBackend::make(Ref<DynObject> config, Ref<DIObject> parent)
{
	// the keys of WaterfallBackend::make (src/WaterfallBackend.cpp:620-644) with the same defaults,
	// plus cuda_device / cuda_batch_frames / cuda_precision / cuda_rows
	int bins      = config->getStrInt("bins",        32768);
	int overlap   = config->getStrInt("overlap",         0);
	string origin = config->getStrString("origin", "debug");
	string rows   = config->getStrString("cuda_rows", "bands");
	int chunkSize = config->getStrInt("buffer_chunk_size", WATERFALL_BACKEND_CHUNK_SIZE);

	if (rows != "bands" && rows != "full")
		LOG_WARNING("CudaWaterfallBackend: cuda_rows = \"" << rows << "\" (expected \"bands\" or \"full\"), using \"bands\"");

	Ref<CudaWaterfallBackend> backend = new CudaWaterfallBackend(
		bins, overlap, origin,
		config->getStrInt("cuda_device", 0),
		config->getStrInt("cuda_batch_frames", 64),
		config->getStrInt("cuda_precision", 0),
		rows == "full",
		chunkSize
	);

	backend->setDebugWriteDelayMs(config->getStrInt("cuda_debug_write_delay_ms", 0));
	backend->setMetadataPath(
		config->getStrString("metadata_path", "."));
	backend->setBufferChunkSize(chunkSize);
	backend->setGain(
		config->getStrDouble("iq_gain", 0));
	backend->setPhaseShift(
		config->getStrInt("iq_phase_shift", 0));

	return backend;
}

CPPAPP_DI_METHOD("waterfall_cuda", CudaWaterfallBackend, make);
