/*
 * CudaRecorders.h -- the reference's recorders, subclassed for CudaWaterfallBackend.
 *
 * SnapshotRecorder and BolidRecorder (src/WaterfallBackend.h:112-225, src/BolidRecorder.h:37-210)
 * keep doing what they do: slicing snapshots out of the row ring, naming files, writing FITS headers
 * and pixels, CSV lines, the `met;` line, Noise/BolidMessage.  Their data members are `protected`,
 * so a subclass can add the three things the GPU backend needs without touching a reference file:
 *
 *   1. every queued snapshot is reported to the backend with the absolute row it starts at, and
 *      reported again when its files are on disk (the producer's back-pressure, see
 *      CudaWaterfallBackend.h);
 *   2. the bolid detector's per-row statistics -- noise, peak bin, local mean
 *      (BolidRecorder::noise/peak/average, src/BolidRecorder.cpp:302-347) -- come from the GPU's
 *      record of the row instead of a qsort on the host; the decision and the state machine
 *      (src/BolidRecorder.cpp:135-273) run here on those three numbers, in row order;
 *   3. raw-I/Q snapshots (SnapshotRecorder::writeRaw, src/WaterfallBackend.cpp:214-267) are cut from
 *      the backend's own sample store; the reference's writeRaw still writes the file.
 *
 *   4. the writer thread is an ordinary pthread on threadMethod().  The reference creates it as a
 *      cppapp MethodThread, whose base-class constructor starts the thread before the derived object
 *      (vtable, instance and method pointers) exists (cppapp/cppapp/Thread.cpp:47-56, Thread.h:88-95):
 *      when the new thread is scheduled first, the process dies at start-up with "pure virtual method
 *      called" or a segmentation fault.  start() / stop() are therefore restated here
 *      (src/WaterfallBackend.cpp:364-412, src/BolidRecorder.cpp:80-116) around a thread that starts
 *      after its arguments are complete.
 *
 * An instance is built from the stock recorder the Injector created, with the same constructor
 * arguments (read back through pointers to the protected members, which a derived class may form).
 */
#ifndef CUDA_RECORDERS_H
#define CUDA_RECORDERS_H

#include <pthread.h>

#include "CudaWaterfallBackend.h"

class CudaSnapshotRecorder : public SnapshotRecorder {
	CudaWaterfallBackend *cuda_;

	CudaSnapshotRecorder(CudaWaterfallBackend *backend, int snapshotLength, float leftFrequency,
	                     float rightFrequency, string outputDir, string outputType, bool compressOutput,
	                     bool listenToNoise);

	pthread_t writer_;
	bool      writerRunning_;
	static void *writerEntry(void *self);

protected:
	virtual void write(Snapshot snapshot);

public:
	/// the same recorder as `stock`, attached to `backend`
	static Ref<CudaSnapshotRecorder> like(const SnapshotRecorder &stock, CudaWaterfallBackend *backend);

	virtual void start();
	virtual void stop();
	virtual void update();

	int leftBin() const { return leftBin_; }
	int rightBin() const { return rightBin_; }
};


class CudaBolidRecorder : public BolidRecorder {
	CudaWaterfallBackend *cuda_;
	bool                  fromGpu_;   // statistics from the GPU record (else: the reference's update())

	CudaBolidRecorder(CudaWaterfallBackend *backend, const BolidRecorder &stock);

	void queued(int start);

	pthread_t writer_;
	bool      writerRunning_;
	static void *writerEntry(void *self);

protected:
	virtual void write(Snapshot snapshot);
	virtual void writeRaw(Snapshot snapshot);

public:
	static Ref<CudaBolidRecorder> like(const BolidRecorder &stock, CudaWaterfallBackend *backend);

	virtual void start();
	virtual void stop();
	virtual void update();

	void useGpuRecords(bool on) { fromGpu_ = on; }

	int leftBin() const { return leftBin_; }
	int rightBin() const { return rightBin_; }
	int lowDetectBin() const { return lowDetectBin_; }
	int detectWidth() const { return detectWidth_; }
	int lowNoiseBin() const { return lowNoiseBin_; }
	int noiseWidth() const { return noiseWidth_; }
	int averageBinRange() const { return averageBinRange_; }
};

#endif /* CUDA_RECORDERS_H */
