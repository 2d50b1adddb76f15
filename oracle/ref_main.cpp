/*
 * oracle/ref_main.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Driver for the UNMODIFIED reference classes, compiled in place from
 * /root/reference by oracle/Makefile into oracle/_ref/ref_oracle.  It does
 * what src/App.cpp:172-248 does (parse the JSON config, build the Pipeline
 * through the Injector, attach a WAVStream frontend, run) minus JACK and
 * signal handling, and adds observation taps that use only the reference's
 * public/protected extension points:
 *   - a Recorder subclass (src/WaterfallBackend.h:46-103) that dumps every
 *     row the backend pushes, with the ring mark and the row's handle;
 *   - NoiseMessage / BolidMessage listeners (src/MessageDispatch.h:196-207)
 *     that capture the detector's per-row (noise, peakFreq, magnitude) and
 *     the emitted bolid events.
 *
 * usage: ref_oracle -c CONFIG.json [-n NAME] [-d DUMPDIR] [-r] [-q] [-N] WAV_FILE
 *   -d DUMPDIR  write rowinfo.bin / noise.bin / bolid.txt / timing.json there
 *   -r          also write every full row to DUMPDIR/rows.f32
 *   -q          quiet (log warnings and errors only, to stderr)
 *   -N          (adapter build only) read the WAV with NarrowWAVStream instead of WAVStream
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/time.h>

#include <cppapp/cppapp.h>
using namespace cppapp;

#include "Pipeline.h"
#include "WAVStream.h"
#include "WaterfallBackend.h"
#include "BolidRecorder.h"
#include "BolidMessage.h"
#include "MessageDispatch.h"
#ifdef RO_NARROW_FRONTEND
/* only in the adapter build (radio_observer_b200/host/Makefile): -N selects the narrow-ingest
 * frontend instead of the reference's WAVStream */
#include "NarrowWAVStream.h"
#endif

struct RowInfo {
	int32_t  mark;      /* buffer_->mark() after the push = index of the NEXT row */
	int32_t  rawMark;
	int64_t  sec;
	int64_t  usec;
};

struct NoiseRec {
	float noise;
	float peakFreq;
	float magnitude;
};

class DumpRecorder : public Recorder {
	FILE *rows_;
	FILE *info_;
	long  count_;
public:
	DumpRecorder(Ref<WaterfallBackend> backend, const std::string &dir, bool dumpRows) :
		Recorder(backend), rows_(NULL), info_(NULL), count_(0)
	{
		if (!dir.empty()) {
			info_ = fopen((dir + "/rowinfo.bin").c_str(), "wb");
			if (dumpRows) rows_ = fopen((dir + "/rows.f32").c_str(), "wb");
		}
	}
	virtual ~DumpRecorder() { close(); }
	void close()
	{
		if (rows_) { fclose(rows_); rows_ = NULL; }
		if (info_) { fclose(info_); info_ = NULL; }
	}
	long count() const { return count_; }
	virtual void update()
	{
		count_++;
		int m = buffer_->mark();
		if (rows_) {
			float *row = buffer_->at(m - 1);
			fwrite(row, sizeof(float), buffer_->getWidth(), rows_);
		}
		if (info_) {
			RawDataHandle h = rawHandles_->at(m);
			RowInfo ri;
			ri.mark = m;
			ri.rawMark = h.mark;
			ri.sec = h.time.seconds();
			ri.usec = h.time.microseconds();
			fwrite(&ri, sizeof ri, 1, info_);
		}
	}
};

static FILE *g_noise = NULL;
static FILE *g_bolid = NULL;

static void onNoise(const NoiseMessage &msg, void *)
{
	if (!g_noise) return;
	NoiseRec r = { msg.noise, msg.peakFrequency, msg.magnitude };
	fwrite(&r, sizeof r, 1, g_noise);
}

static void onBolid(const BolidMessage &msg, void *)
{
	if (!g_bolid) return;
	fprintf(g_bolid, "%.9g;%.9g;%.9g;%.9g;%.9g;%d;%d\n",
	        msg.noise, msg.peakFrequency, msg.magnitude,
	        msg.minFreq, msg.maxFreq, msg.startSample, msg.endSample);
	fflush(g_bolid);
}

static double nowSeconds()
{
	struct timeval tv;
	gettimeofday(&tv, NULL);
	return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

int main(int argc, char **argv)
{
	std::string configFile, cfgName, dumpDir, wavFile;
	bool dumpRows = false, quiet = false, narrow = false;
	for (int i = 1; i < argc; i++) {
		std::string a = argv[i];
		if (a == "-c" && i + 1 < argc) configFile = argv[++i];
		else if (a == "-n" && i + 1 < argc) cfgName = argv[++i];
		else if (a == "-d" && i + 1 < argc) dumpDir = argv[++i];
		else if (a == "-r") dumpRows = true;
		else if (a == "-q") quiet = true;
		else if (a == "-N") narrow = true;
		else wavFile = a;
	}
	if (configFile.empty() || wavFile.empty()) {
		fprintf(stderr, "usage: %s -c CONFIG.json [-n NAME] [-d DUMPDIR] [-r] [-q] WAV_FILE\n", argv[0]);
		return 2;
	}

	Logger::clearConfig();
	Logger::addOutput(quiet ? LOG_LVL_WARNING : LOG_LVL_INFO, "-");

	Ref<FileInput> configInput = new FileInput(configFile);
	if (!configInput->exists()) {
		fprintf(stderr, "config file %s does not exist\n", configFile.c_str());
		return 2;
	}
	JSONParser parser;
	Ref<DynObject> config = parser.parse(configInput);
	configInput->close();
	if (config->isError()) {
		Ref<DynError> err = config.as<DynError>();
		std::cerr << "Syntax error in config file at " << err->getErrorLoc()
		          << ": " << err->getString() << std::endl;
		return 1;
	}

	if (cfgName.empty()) cfgName = config->getStrString("configuration", "default");
	Injector::getInstance().makePlans(config->getStrItem("configurations"));
	Ref<Pipeline> pipeline = Injector::getInstance().instantiateAs<Pipeline>(cfgName);
	if (pipeline.isNull()) {
		fprintf(stderr, "initialization failed\n");
		return 3;
	}

	Ref<WaterfallBackend> backend = pipeline->getBackend().as<WaterfallBackend>();
	if (backend.isNull()) {
		fprintf(stderr, "backend is not a WaterfallBackend\n");
		return 3;
	}
	Ref<DumpRecorder> dump = new DumpRecorder(backend, dumpDir, dumpRows);
	backend->addRecorder(dump);

	if (!dumpDir.empty()) {
		g_noise = fopen((dumpDir + "/noise.bin").c_str(), "wb");
		g_bolid = fopen((dumpDir + "/bolid.txt").c_str(), "w");
	}
	addListener<NoiseMessage>(onNoise, NULL);
	addListener<BolidMessage>(onBolid, NULL);

	if (narrow) {
#ifdef RO_NARROW_FRONTEND
		pipeline->setFrontend(new NarrowWAVStream(new FileInput(wavFile)));
#else
		fprintf(stderr, "-N: this build has no narrow-ingest frontend\n");
		return 2;
#endif
	} else {
		pipeline->setFrontend(new WAVStream(new FileInput(wavFile)));
	}

	double t0 = nowSeconds();
	pipeline->run();
	double t1 = nowSeconds();

	dump->close();
	if (g_noise) fclose(g_noise);
	if (g_bolid) fclose(g_bolid);

	StreamInfo si = backend->getStreamInfo();
	char line[512];
	snprintf(line, sizeof line,
	         "{\"rows\": %ld, \"bins\": %d, \"fft_sample_rate\": %.9g, \"sample_rate\": %d, "
	         "\"run_seconds\": %.6f}",
	         dump->count(), backend->getBins(), (double)backend->getFFTSampleRate(),
	         si.sampleRate, t1 - t0);
	fprintf(stderr, "%s\n", line);
	if (!dumpDir.empty()) {
		FILE *f = fopen((dumpDir + "/timing.json").c_str(), "w");
		if (f) { fprintf(f, "%s\n", line); fclose(f); }
	}
	return 0;
}
