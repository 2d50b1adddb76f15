/*
 * CudaRecorders.cpp -- see CudaRecorders.h.
 */
#include "CudaRecorders.h"

#include "BolidMessage.h"
#include "MessageDispatch.h"
#include "CsvLog.h"
#include "utils.h"

#include <unistd.h>

// WFTime::format goes through the non-reentrant gmtime() (src/WFTime.cpp:25-30), and the reference
// formats times on the producer thread (file names, CSV entries) and on every recorder's writer
// thread (DATE-OBS, "Local time", CSV entries) with nothing in between.  At the reference's speed the
// calls rarely meet; behind this backend the producer queues snapshots while earlier ones are being
// written, and a name now and then came out with another thread's date.  Everything these subclasses
// do that formats a time happens under one lock.
static Mutex g_timeFormat;


////////////////////////////////////////////////////////////////////////////////
// SNAPSHOT RECORDER
////////////////////////////////////////////////////////////////////////////////

CudaSnapshotRecorder::CudaSnapshotRecorder(CudaWaterfallBackend *backend, int snapshotLength,
                                           float leftFrequency, float rightFrequency, string outputDir,
                                           string outputType, bool compressOutput, bool listenToNoise) :
	SnapshotRecorder(backend, snapshotLength, leftFrequency, rightFrequency, outputDir, outputType,
	                 compressOutput, listenToNoise),
	cuda_(backend),
	writerRunning_(false)
{
}


// SnapshotRecorder::start (src/WaterfallBackend.cpp:364-397) up to the thread: the recorded columns
// and the first snapshot.  A macro because both subclasses reach the same protected members through
// different bases.
#define RO_BEGIN_SNAPSHOTS()                                                                  \
	do {                                                                                      \
		LOG_INFO("Snapshot recording starting...");                                           \
		if (leftFrequency_ == rightFrequency_) {                                              \
			StreamInfo info = backend_->getStreamInfo();                                      \
			leftFrequency_ = -(float)info.sampleRate / 2.0;                                   \
			rightFrequency_ = (float)info.sampleRate / 2.0;                                   \
			leftBin_  = 0;                                                                    \
			rightBin_ = backend_->getBins();                                                  \
		} else {                                                                              \
			leftBin_  = backend_->frequencyToBin(leftFrequency_);                             \
			rightBin_ = backend_->frequencyToBin(rightFrequency_);                            \
		}                                                                                     \
		nextSnapshot_ = Snapshot(0);                                                          \
		nextSnapshot_.fileName = getFileName(nextSnapshot_.start);                            \
		workerThread_ = NULL;                                                                 \
		writerRunning_ = pthread_create(&writer_, NULL, writerEntry, this) == 0;              \
		if (!writerRunning_) LOG_ERROR("CudaRecorders: could not start the writer thread");   \
	} while (0)

// SnapshotRecorder::stop (src/WaterfallBackend.cpp:400-412)
#define RO_END_SNAPSHOTS()                                                                    \
	do {                                                                                      \
		if ((buffer_->size(nextSnapshot_.start) >= 0) && writeUnfinished_) {                  \
			MutexLock lock(&g_timeFormat);                                                    \
			startWriting();                                                                   \
		}                                                                                     \
		snapshots_.close();                                                                   \
		if (writerRunning_) pthread_join(writer_, NULL);                                      \
		writerRunning_ = false;                                                               \
	} while (0)


void *CudaSnapshotRecorder::writerEntry(void *self)
{
	return static_cast<CudaSnapshotRecorder *>(self)->threadMethod();
}


void CudaSnapshotRecorder::start()
{
	RO_BEGIN_SNAPSHOTS();
}


void CudaSnapshotRecorder::stop()
{
	RO_END_SNAPSHOTS();
}


Ref<CudaSnapshotRecorder> CudaSnapshotRecorder::like(const SnapshotRecorder &stock, CudaWaterfallBackend *backend)
{
	// a derived class may form pointers to the protected members of its base; applied to `stock` they
	// read back what SnapshotRecorder::make handed to the constructor (src/WaterfallBackend.cpp:436-456)
	typedef CudaSnapshotRecorder Me;
	Ref<CudaSnapshotRecorder> r = new CudaSnapshotRecorder(
		backend,
		stock.*(&Me::snapshotLength_),
		stock.*(&Me::leftFrequency_),
		stock.*(&Me::rightFrequency_),
		stock.*(&Me::outputDir_),
		stock.*(&Me::outputType_),
		stock.*(&Me::compressOutput_),
		stock.*(&Me::listenToNoise_));
	r->writeUnfinished_ = stock.*(&Me::writeUnfinished_);
	return r;
}


void CudaSnapshotRecorder::update()
{
	// src/WaterfallBackend.cpp:415-427, with the snapshot noted before it is queued
	if (buffer_->size(nextSnapshot_.start) >= snapshotRows_ + 2) {
		LOG_DEBUG("SnapshotRecorder: Snapshot full [start_: " << nextSnapshot_.start <<
		          ", snapshotRows_: " << snapshotRows_ <<
		          ", buffer_->size(start_): " << buffer_->size(nextSnapshot_.start) << "].");
		cuda_->logProcessingTimes();
		cuda_->clearProcessingTime();
		cuda_->noteQueued(this, nextSnapshot_.start);
		MutexLock lock(&g_timeFormat);
		startWriting();
	}
}


void CudaSnapshotRecorder::write(Snapshot snapshot)
{
	if (cuda_->debugWriteDelayMs()) usleep(1000 * cuda_->debugWriteDelayMs());
	{
		MutexLock lock(&g_timeFormat);
		SnapshotRecorder::write(snapshot);
	}
	cuda_->noteWritten(this, snapshot.start);
}


////////////////////////////////////////////////////////////////////////////////
// BOLID RECORDER
////////////////////////////////////////////////////////////////////////////////

CudaBolidRecorder::CudaBolidRecorder(CudaWaterfallBackend *backend, const BolidRecorder &stock) :
	BolidRecorder(backend,
	              stock.*(&CudaBolidRecorder::snapshotLength_),
	              stock.*(&CudaBolidRecorder::leftFrequency_),
	              stock.*(&CudaBolidRecorder::rightFrequency_),
	              stock.*(&CudaBolidRecorder::outputDir_),
	              stock.*(&CudaBolidRecorder::outputType_),
	              stock.*(&CudaBolidRecorder::compressOutput_),
	              stock.*(&CudaBolidRecorder::minDetectFq_),
	              stock.*(&CudaBolidRecorder::maxDetectFq_),
	              stock.*(&CudaBolidRecorder::minNoiseFq_),
	              stock.*(&CudaBolidRecorder::maxNoiseFq_),
	              stock.*(&CudaBolidRecorder::advanceTime_),
	              stock.*(&CudaBolidRecorder::jitterTime_),
	              stock.*(&CudaBolidRecorder::averageFrequencyRange_),
	              stock.*(&CudaBolidRecorder::thresholdRatio_),
	              stock.*(&CudaBolidRecorder::noiseMetadataTime_)),
	cuda_(backend),
	fromGpu_(false),
	writerRunning_(false)
{
}


void *CudaBolidRecorder::writerEntry(void *self)
{
	return static_cast<CudaBolidRecorder *>(self)->threadMethod();
}


void CudaBolidRecorder::start()
{
	// BolidRecorder::start (src/BolidRecorder.cpp:80-116): bins from frequencies, rows from times
	int minFqBin = backend_->frequencyToBin(minDetectFq_);
	int maxFqBin = backend_->frequencyToBin(maxDetectFq_);
	ORDER_PAIR(minFqBin, maxFqBin);
	lowDetectBin_ = minFqBin;
	detectWidth_  = maxFqBin - minFqBin;

	minFqBin = backend_->frequencyToBin(minNoiseFq_);
	maxFqBin = backend_->frequencyToBin(maxNoiseFq_);
	lowNoiseBin_ = min(minFqBin, maxFqBin);
	noiseWidth_  = max(minFqBin, maxFqBin) - lowNoiseBin_;
	noiseBuffer_.resize(noiseWidth_);

	CPPAPP_ASSERT(advanceTime_ >= 0.0);
	CPPAPP_ASSERT(jitterTime_ >= 0.0);
	CPPAPP_ASSERT(averageFrequencyRange_ > 0.0);
	advance_           = backend_->timeToFFTSamples(advanceTime_);
	jitter_            = backend_->timeToFFTSamples(jitterTime_);
	averageBinRange_   = backend_->frequencyToBin(averageFrequencyRange_) - backend_->frequencyToBin(0);
	noiseMetadataRows_ = backend_->timeToFFTSamples(noiseMetadataTime_);
	CPPAPP_ASSERT(averageBinRange_ > 0);

	lastNoiseMetadataEntry_ = 0;
	state_ = STATE_INIT;

	LOG_INFO("Bolid detector starting...");
	LOG_INFO("Freq.: " << leftFrequency_ << "-" << rightFrequency_ <<
	         ", detect. freq.: " << minDetectFq_ << "-" << maxDetectFq_ <<
	         ", noise freq.: " << minNoiseFq_ << "-" << maxNoiseFq_);
	RO_BEGIN_SNAPSHOTS();
}


void CudaBolidRecorder::stop()
{
	RO_END_SNAPSHOTS();
}


Ref<CudaBolidRecorder> CudaBolidRecorder::like(const BolidRecorder &stock, CudaWaterfallBackend *backend)
{
	return new CudaBolidRecorder(backend, stock);
}


void CudaBolidRecorder::queued(int start)
{
	cuda_->noteQueued(this, start);
}


void CudaBolidRecorder::update()
{
	if (!fromGpu_) {
		// The reference's update() as it stands (whole rows are on the host).  It queues a snapshot only
		// on the transition BOLID_ENDED -> INIT; the writer thread may finish before update() returns,
		// so the note is made beforehand and withdrawn when nothing was queued.
		const bool mayQueue = state_ == STATE_BOLID_ENDED;
		const int  start = nextSnapshot_.start;
		if (mayQueue) queued(start);
		{
			MutexLock lock(&g_timeFormat);
			BolidRecorder::update();
		}
		if (mayQueue && state_ != STATE_INIT) cuda_->noteWritten(this, start);
		return;
	}

	// src/BolidRecorder.cpp:119-138 with the three statistics taken from the GPU's record of this row
	// (same definitions: ro_cuda.h, ro_det_rec), then the decision and the state machine of :135-273.
	const ro_det_rec &rec = cuda_->currentRecord();
	const float n = rec.noise;
	const int   p = rec.peak;
	const float a = rec.avg;
	const float peakFq = backend_->binToFrequency(lowDetectBin_ + p);

	const bool detect = (a > (n * 2.0));

	NoiseMessage msg(WFTime::now(), n, peakFq, a);
	sendMessage(msg);

	switch (state_) {
	case STATE_INIT:
		if (detect) {
			peakFreq_ = peakFq;
			noise_ = n;
			magnitude_ = a;
			duration_ = 1;
			nextSnapshot_.start = buffer_->mark() - advance_;
			nextSnapshot_.length = 2 * advance_;
			{
				MutexLock lock(&g_timeFormat);
				nextSnapshot_.fileName = getFileName(nextSnapshot_.start);
			}
			state_ = STATE_BOLID;
		}
		break;

	case STATE_BOLID:
		if (detect) {
			duration_ += 1;
		} else {
			nextSnapshot_.length += duration_;
			duration_ = 1;
			state_ = STATE_BOLID_ENDED;
		}
		break;

	case STATE_BOLID_ENDED:
		duration_ += 1;
		if (detect) {
			state_ = STATE_BOLID;
		} else if (duration_ >= jitter_) {
			const float duration = (float)(nextSnapshot_.length - 2 * advance_) / (float)backend_->getFFTSampleRate();
			const float quarter = (maxDetectFq_ - minDetectFq_) / 4;
			WFTime t = WFTime::now();
			MutexLock lock(&g_timeFormat);

			CSV_LOG_ENTRY(
				backend_->getMetadataFile(),
				t,
				Path::basename(nextSnapshot_.fileName) << ";"
				<< noise_ << ";"
				<< peakFreq_ << ";"
				<< magnitude_ << ";"
				<< duration
			);

			sendMessage(BolidMessage(t, noise_, peakFreq_, magnitude_,
			                         peakFreq_ - quarter, peakFreq_ + quarter,
			                         0, fftSamplesToRaw(nextSnapshot_.length)));

			cout << "met;" << t << ";"
			     << noise_ << ";"
			     << peakFreq_ << ";"
			     << magnitude_ << ";"
			     << peakFreq_ - quarter << ";"
			     << peakFreq_ + quarter << ";"
			     << duration << ";"
			     << fftSamplesToRaw(nextSnapshot_.length) << "#" << endl;

			LOG_WARNING("************** METEOR DETECTED **************");
			LOG_INFO("Duration: " << duration << "s" << "  |  Frequency: " << peakFreq_ << "Hz");
			nextSnapshot_.includeRawData = true;
			queued(nextSnapshot_.start);
			startWriting();
			state_ = STATE_INIT;
		}
		break;

	default:
		LOG_ERROR("Internal error: unknown state " << state_ << " in bolid detector!");
		state_ = STATE_INIT;
		break;
	}
}


void CudaBolidRecorder::write(Snapshot snapshot)
{
	if (cuda_->debugWriteDelayMs()) usleep(1000 * cuda_->debugWriteDelayMs());
	{
		MutexLock lock(&g_timeFormat);
		SnapshotRecorder::write(snapshot);
	}
	// a bolid's raw-I/Q file follows (threadMethod, src/WaterfallBackend.cpp:78-80): done after that one
	if (!snapshot.includeRawData) cuda_->noteWritten(this, snapshot.start);
}


void CudaBolidRecorder::writeRaw(Snapshot snapshot)
{
	// SnapshotRecorder::writeRaw (src/WaterfallBackend.cpp:214-267) card for card, with the pixels handed
	// to the FITS writer in ONE call.  The reference fetches and writes a raw-I/Q image one sample (two
	// floats) at a time -- rawBuffer_->at(start + y), then fits_write_pix -- a million calls per bolid; a
	// 30 min station recording has 300 of them and 3.2 GB of raw-I/Q files, and behind a backend that
	// needs 2 s for its DSP the writer thread was the whole run time (12 s).  Same file, byte for byte:
	// start = fftMarkToRaw(snapshot.start), length = fftSamplesToRaw(snapshot.length) with Recorder's
	// int-truncated row rate (src/WaterfallBackend.h:84-87), samples from the backend's own store.
	if (cuda_->rawCapacity() <= 0) {            // the reference's raw ring is in use (a recorder of unknown type is attached)
		MutexLock lock(&g_timeFormat);
		SnapshotRecorder::writeRaw(snapshot);
		cuda_->noteWritten(this, snapshot.start);
		return;
	}
	// (Recorder::fftMarkToRaw / fftMarkToTime are declared inline but defined in the reference's own translation unit)
	const RawDataHandle &handle = rawHandles_->at(wrap(snapshot.start, rawHandles_->size()));
	const int start  = handle.mark;
	const int length = fftSamplesToRaw(snapshot.length);
	WFTime time = handle.time;
	std::vector<float> pixels(2 * (size_t)(length > 0 ? length : 0));
	if (length > 0) cuda_->copyRaw(start, length, &pixels[0]);
	{
		MutexLock lock(&g_timeFormat);
		string origin = backend_->getOrigin();
		float sampleRate = (float)backend_->getStreamInfo().sampleRate;

		string fileName = "!";
		fileName += Path::join(outputDir_, getFileName("raws", "fits", time));
		LOG_INFO("Writing raw snapshot \"" << (fileName.c_str() + 1) << "\"...");

		FITSWriter w;
		if (w.open(fileName.c_str())) {
			w.createImage(2, length, FLOAT_IMG);
			writeHeader(&w);
			w.writeHeader("ORIGIN", origin.c_str(), "");
			w.date();
			w.comment(WFTime::now().format("Local time: %Y-%m-%d %H:%M:%S %Z", true).c_str());
			w.writeHeader("DATE-OBS", time.format("%Y-%m-%dT%H:%M:%S").c_str(), "observation date (UTC)");

			w.writeHeader("CTYPE2", "TIME", "in seconds");
			w.writeHeader("CRPIX2", 1, "");
			w.writeHeader("CRVAL2", (long long)time.toMilliseconds(), "unix time of the first IQ sample in this file in ms");
			w.writeHeader("CDELT2", ((float)MS_IN_SECOND) / (float)sampleRate, "time difference between two IQ samples in ms");

			w.writeHeader("CTYPE1", "CHAN", "in Hz");
			w.writeHeader("CRPIX1", 1.f, "");
			w.writeHeader("CRVAL1", 0, "");
			w.writeHeader("CDELT1", 1, "");
			w.checkStatus("Error occured while writing FITS file header.");

			if (length > 0) w.write(0, length, &pixels[0]);
			w.checkStatus("Error occured while writing data to a FITS file.");
			w.close();
			LOG_DEBUG("Finished writing raw snapshot.");
		}
	}
	cuda_->noteWritten(this, snapshot.start);
}
