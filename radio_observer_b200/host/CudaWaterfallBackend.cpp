/*
 * CudaWaterfallBackend.cpp -- see CudaWaterfallBackend.h.
 *
 * What stays on the host, restated from FFTBackend::process
 * (reference src/FFTBackend.cpp:192-279) because the recorders depend on it:
 *   - every incoming sample is pushed to the raw I/Q ring (rawBuffer_) and gets a
 *     RawDataHandle {ring mark, stream time}            (:217-223, :264-270)
 *   - a frame is complete when the window holds bins_ samples; its row time is the
 *     handle time of the window's first sample            (:225)
 *   - the handle table is shifted by the hop BEFORE processFFT is called, so the
 *     rawMark handed over is the one of the NEXT frame's first sample whenever
 *     overlap > 0                                           (:241-251)
 * What moves to the GPU: IQ gain, window, FFT, magnitude.  A finished row comes
 * back as bins_ fftshifted floats; it is handed to the unmodified
 * WaterfallBackend::processFFT as the complex vector (row[(i+N/2)%N], 0), whose
 * magnitude sqrt(x*x + 0) is exactly x in double, so the ring receives the GPU
 * row bit for bit and every recorder runs as in the reference.
 */
#include "CudaWaterfallBackend.h"

#include <cstring>
#include <climits>
#include <unistd.h>

CudaWaterfallBackend::CudaWaterfallBackend(int bins, int overlap, std::string origin,
                                           int device, int batchFrames, int precision) :
	WaterfallBackend(bins, overlap, origin),
	overlap_(overlap),
	device_(device),
	batchFrames_(batchFrames < 1 ? 1 : batchFrames),
	precision_(precision),
	iqGain_(0), iqPhaseShift_(0),
	wf_(NULL), failed_(false),
	fill_(0),
	inFormat_(RO_IN_COMPLEX64), samplesIn_(0), rawPushed_(0),
	delivered_(0)
{
	if (overlap_ < 0) overlap_ = 0;
	if (overlap_ >= bins_) overlap_ = bins_ - 1;
}


CudaWaterfallBackend::~CudaWaterfallBackend()
{
	if (wf_) ro_wf_destroy(wf_);
	wf_ = NULL;
}


void CudaWaterfallBackend::fail(const char *what)
{
	// the reference's convention on the data path: log and stop producing
	LOG_ERROR("CudaWaterfallBackend: " << what << ": " << ro_last_error());
	failed_ = true;
}


void CudaWaterfallBackend::startStream(StreamInfo info)
{
	// sizes the row ring and the raw ring, starts the recorders, sets
	// fftSampleRate_ (src/WaterfallBackend.cpp:573-594, src/FFTBackend.cpp:144-189)
	WaterfallBackend::startStream(info);

	fill_ = 0;
	info_ = DataInfo();
	pending_.clear();
	toNextHop_ = 0;
	hopHandles_.clear();
	{
		int cap = (1 << 22) / bins_;          // about 16 MiB of harvested rows at a time
		if (cap < 1) cap = 1;
		if (cap > 64) cap = 64;
		rows_.resize((size_t)bins_ * cap);
	}
	delivered_ = 0;
	failed_ = false;

	samplesIn_ = 0;
	rawPushed_ = 0;
	resAbs_.clear();
	resStart_.clear();
	// the Pipeline's frontends hand over vector<Complex>; a narrow frontend switches the format with
	// its first processRaw()
	createHandle(RO_IN_COMPLEX64);
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
	cfg.n_bands = 0;                          // recorders slice the full row themselves
	cfg.device = device_;
	cfg.max_batch_frames = batchFrames_;
	cfg.precision = precision_;
	inFormat_ = inFormat;
	if (ro_wf_create(&cfg, &wf_) != RO_OK) {
		wf_ = NULL;
		fail("ro_wf_create");
		return false;
	}
	return true;
}


bool CudaWaterfallBackend::ensureFormat(int inFormat)
{
	if (inFormat_ == inFormat) return true;
	if (samplesIn_ > 0) {
		// one stream, one encoding: the device-side sample ring holds what was pushed so far
		LOG_ERROR("CudaWaterfallBackend: process() and processRaw() mixed within one stream");
		failed_ = true;
		return false;
	}
	return createHandle(inFormat);
}


void CudaWaterfallBackend::pushSamples(const Complex *src, int count, WFTime timeOffset)
{
	// raw I/Q ring: one push per sample like src/FFTBackend.cpp:218; the handle (ring mark after the
	// push, stream time of the sample) only where a frame can start
	waitBeforeRawPush(count);
	rawPushed_ += count;
	for (int i = 0; i < count; i++) {
		floatToInt(src[i], rawBuffer_.push());
		if (toNextHop_ == 0) {
			hopHandles_.push_back(RawDataHandle(
				rawBuffer_.mark(),
				timeOffset.addSamples(i, streamInfo_.sampleRate)
			));
			toNextHop_ = bins_ - overlap_;
		}
		toNextHop_--;
	}
	if (ro_wf_push(wf_, src, (size_t)count) != RO_OK) fail("ro_wf_push");
}


void CudaWaterfallBackend::pushSamples(const int16_t *src, int count, WFTime timeOffset)
{
	// Same effects as the Complex overload on the widened block, without a per-sample step.
	// Handles: sample i of this run starts a frame when i = toNextHop_ (mod hop); the ring mark after
	// its push is the mark before the run + i + 1, wrapped (src/RingBuffer.h:289-290).
	const int hop = bins_ - overlap_;
	const int cap = rawBuffer_.capacity_;
	waitBeforeRawPush(count);
	rawPushed_ += count;
	{
		const long mark0 = rawBuffer_.mark();
		int i = toNextHop_;
		for (; i < count; i += hop) {
			hopHandles_.push_back(RawDataHandle(
				(int)((mark0 + i + 1) % cap),
				timeOffset.addSamples(i, streamInfo_.sampleRate)
			));
		}
		toNextHop_ = i - count;
	}

	// Raw I/Q ring: (float)(double)int16 = (float)int16, exactly what floatToInt(Complex, float*)
	// stores (src/FFTBackend.h:250-254); rows of one chunk are contiguous (src/RingBuffer.h:257-259)
	int done = 0;
	while (done < count) {
		const int index = rawBuffer_.head_.index;
		int run = rawBuffer_.chunkRows_ - index % rawBuffer_.chunkRows_;
		if (run > count - done) run = count - done;
		float *dst = rawBuffer_.head_.item;
		const int16_t *s = src + 2 * (size_t)done;
		for (int j = 0; j < 2 * run; j++) dst[j] = (float)s[j];
		rawBuffer_.head_ = IQBuffer::ChunkItem(rawBuffer_, (index + run) % cap,
		                                       rawBuffer_.head_.generation + run);
		rawBuffer_.size_ = rawBuffer_.size_ + run < cap ? rawBuffer_.size_ + run : cap;
		done += run;
	}

	if (ro_wf_push(wf_, src, (size_t)count) != RO_OK) fail("ro_wf_push");
}


void CudaWaterfallBackend::process(const std::vector<Complex> &data, DataInfo info)
{
	if (failed_ || !wf_ || data.empty()) return;
	if (!ensureFormat(RO_IN_COMPLEX64)) return;
	processSamples(&(data[0]), (int)data.size(), info);
}


void CudaWaterfallBackend::processRaw(const int16_t *iq, int count, DataInfo info)
{
	if (failed_ || !wf_ || count <= 0) return;
	if (!ensureFormat(RO_IN_INT16)) return;
	processSamples(iq, count, info);
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


void CudaWaterfallBackend::harvest(bool flush, int atMost)
{
	if (flush && ro_wf_flush(wf_) != RO_OK) { fail("ro_wf_flush"); return; }
	const size_t cap = rows_.size() / (size_t)bins_;
	size_t want = atMost > 0 ? (size_t)atMost : pending_.size();
	while (want > 0 && !pending_.empty()) {
		size_t n = 0;
		uint64_t first = 0;
		if (ro_wf_poll(wf_, &rows_[0], cap < want ? cap : want, NULL, &n, &first) != RO_OK) { fail("ro_wf_poll"); return; }
		if (n == 0) break;
		for (size_t r = 0; r < n && !pending_.empty(); r++) {
			// what processFFT does after its magnitude loop (src/WaterfallBackend.cpp:488,507,534-536):
			// the GPU row is already magnitude, fftshifted, float
			// this push overwrites absolute row delivered_ - capacity
			waitForWriters(delivered_ - buffer_.capacity_ + 1);
			float *row = buffer_.push();
			memcpy(row, &rows_[r * (size_t)bins_], sizeof(float) * (size_t)bins_);
			PendingFrame p = pending_.front();
			pending_.pop_front();
			rawHandles_[buffer_.mark()] = RawDataHandle(p.rawMark, p.info.timeOffset);
			FOR_EACH(recorders_, it) {
				(*it)->update();
			}
			delivered_++;
			noteReservations();
		}
		want -= n < want ? n : want;
	}
}


void CudaWaterfallBackend::noteReservations()
{
	// called on the producer thread right after the recorders' update(): reserve() only ever runs
	// there (SnapshotRecorder::startWriting, src/WaterfallBackend.cpp:108-121), the writers only free
	MutexLock lock(&bufferMutex_);
	const size_t n = buffer_.reservations_.size();
	if (resAbs_.size() < n) { resAbs_.resize(n, -1); resStart_.resize(n, -1); }
	const int cap = buffer_.capacity_;
	const int head = buffer_.mark();             // index of the next row = delivered_ mod capacity
	for (size_t i = 0; i < n; i++) {
		if (!buffer_.reservations_[i].alive) { resAbs_[i] = -1; continue; }
		const int start = buffer_.reservations_[i].start;
		if (resAbs_[i] != -1 && resStart_[i] == start) continue;
		const int back = ((head - start) % cap + cap) % cap;   // rows between the snapshot's first row and the head
		resAbs_[i] = delivered_ - back;          // may be negative: a bolid right after the start of the stream
		resStart_[i] = start;
		if (resAbs_[i] == -1) resAbs_[i] = -2;   // keep -1 for "none"
	}
}


long CudaWaterfallBackend::oldestReservedRow()
{
	MutexLock lock(&bufferMutex_);
	long oldest = LONG_MAX;
	const size_t n = buffer_.reservations_.size() < resAbs_.size() ? buffer_.reservations_.size() : resAbs_.size();
	for (size_t i = 0; i < n; i++) {
		if (resAbs_[i] == -1) continue;
		if (!buffer_.reservations_[i].alive || buffer_.reservations_[i].start != resStart_[i]) continue;
		if (resAbs_[i] < oldest) oldest = resAbs_[i];
	}
	return oldest;
}


void CudaWaterfallBackend::waitForWriters(long firstNeededRow)
{
	// rows below firstNeededRow are about to be lost: wait until no queued snapshot starts there.
	// A snapshot that old is complete, so its writer does not depend on rows still to come; the
	// time limit only guards against a writer that died.
	if (firstNeededRow <= 0 && resAbs_.empty()) return;
	long waited = 0;
	while (oldestReservedRow() < firstNeededRow) {
		usleep(200);
		waited += 200;
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
	const long firstNeededSample = rawPushed_ + count - rawBuffer_.capacity_;
	if (firstNeededSample <= 0) return;
	waitForWriters((firstNeededSample + hop - 1) / hop);
}


void CudaWaterfallBackend::endStream()
{
	// complete frames still in flight are delivered; the incomplete tail is dropped
	// exactly as the reference never flushes window_ (src/FFTBackend.cpp:282-286)
	if (wf_ && !failed_) harvest(true);
	WaterfallBackend::endStream();
	if (wf_) { ro_wf_destroy(wf_); wf_ = NULL; }
}


Ref<DIObject> CudaWaterfallBackend::make(Ref<DynObject> config, Ref<DIObject> parent)
{
	// the keys of WaterfallBackend::make (src/WaterfallBackend.cpp:620-644) with
	// the same defaults, plus cuda_device / cuda_batch_frames / cuda_precision
	int bins      = config->getStrInt("bins",        32768);
	int overlap   = config->getStrInt("overlap",         0);
	string origin = config->getStrString("origin", "debug");

	Ref<CudaWaterfallBackend> backend = new CudaWaterfallBackend(
		bins, overlap, origin,
		config->getStrInt("cuda_device", 0),
		config->getStrInt("cuda_batch_frames", 1),
		config->getStrInt("cuda_precision", 0)
	);

	backend->setMetadataPath(
		config->getStrString("metadata_path", "."));
	backend->setBufferChunkSize(
		config->getStrInt("buffer_chunk_size", WATERFALL_BACKEND_CHUNK_SIZE));
	backend->setGain(
		config->getStrDouble("iq_gain", 0));
	backend->setPhaseShift(
		config->getStrInt("iq_phase_shift", 0));

	return backend;
}

CPPAPP_DI_METHOD("waterfall_cuda", CudaWaterfallBackend, make);
