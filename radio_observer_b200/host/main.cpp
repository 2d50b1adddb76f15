/*
 * host/main.cpp -- radio-observer-cuda: the reference's Pipeline with the CUDA backend linked in.
 *
 * The sequence of App::onRun (reference src/App.cpp:172-248): read the JSON configuration, build the
 * Pipeline through the Injector (every factory of the reference is linked in, plus "waterfall_cuda"),
 * choose a frontend (App::createFrontend, src/App.cpp:57-102), run until the stream ends or SIGINT
 * asks the frontend to stop.  The shipped src/App.cpp cannot be compiled here (it includes the JACK
 * headers); the JACK frontend is therefore absent from this binary, everything else is the reference's
 * own code compiled in place.
 *
 * usage: radio-observer-cuda -c CONFIG.json [-n NAME] [-q] [-N] [-d DIR [-r]] [WAV_FILE]
 *   -c CONFIG   configuration file (as the reference's -c; default ~/.radio-observer.json)
 *   -n NAME     configuration to instantiate (default: the file's "configuration" key)
 *   -q          log warnings and errors only, to stderr (default: the file's "logging" list)
 *   -N          read WAV_FILE with the narrow-ingest frontend (int16 blocks straight to the GPU backend)
 *   -d DIR      diagnostics: per-row handles (rowinfo.bin), per-row detector messages (noise.bin),
 *               bolid events (bolid.txt), run time (timing.json); with -r also every row (rows.f32),
 *               which makes the CUDA backend bring whole rows back from the GPU
 *   -v          print the version and exit
 */
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/time.h>

#include <cppapp/cppapp.h>
using namespace cppapp;

#include "Pipeline.h"
#include "WAVStream.h"
#include "RawStream.h"
#include "BolidMessage.h"
#include "MessageDispatch.h"
#include "config.h"
#include "CudaWaterfallBackend.h"
#include "NarrowWAVStream.h"

namespace {

// ---- diagnostics (-d): what a test or an operator compares between two backends --------------
struct RowInfoRecord { int32_t mark, rawMark; int64_t sec, usec; };
struct NoiseRecord { float noise, peakFrequency, magnitude; };

/// per row: the ring mark after the push (= index of the NEXT row) and the handle stored there
class RowInfoTap : public RowMetaTap {
protected:
	FILE *info_;
	long  count_;
public:
	RowInfoTap(Ref<WaterfallBackend> backend, const std::string &dir) :
		RowMetaTap(backend), info_(dir.empty() ? NULL : fopen((dir + "/rowinfo.bin").c_str(), "wb")), count_(0) {}
	virtual ~RowInfoTap() { finish(); }
	virtual void finish() { if (info_) fclose(info_); info_ = NULL; }
	long count() const { return count_; }
	virtual void update()
	{
		count_++;
		if (!info_) return;
		const int m = buffer_->mark();
		const RawDataHandle &h = rawHandles_->at(m);
		RowInfoRecord r = { m, h.mark, (int64_t)h.time.seconds(), (int64_t)h.time.microseconds() };
		fwrite(&r, sizeof r, 1, info_);
	}
};

/// the same plus every row: an ordinary Recorder (not a RowMetaTap), so whole rows must reach the host
class RowDumpTap : public Recorder {
	RowInfoTap meta_;
	FILE *rows_;
public:
	RowDumpTap(Ref<WaterfallBackend> backend, const std::string &dir) :
		Recorder(backend), meta_(backend, dir), rows_(fopen((dir + "/rows.f32").c_str(), "wb")) {}
	virtual ~RowDumpTap() { finish(); }
	void finish() { meta_.finish(); if (rows_) fclose(rows_); rows_ = NULL; }
	long count() const { return meta_.count(); }
	virtual void update()
	{
		meta_.setBuffer(buffer_, rawBuffer_, bufferMutex_, rawHandles_);
		meta_.update();
		if (rows_) fwrite(buffer_->at(buffer_->mark() - 1), sizeof(float), buffer_->getWidth(), rows_);
	}
};

FILE *g_noise = NULL, *g_bolid = NULL;

void onNoise(const NoiseMessage &msg, void *)
{
	if (!g_noise) return;
	NoiseRecord r = { msg.noise, msg.peakFrequency, msg.magnitude };
	fwrite(&r, sizeof r, 1, g_noise);
}

void onBolid(const BolidMessage &msg, void *)
{
	if (!g_bolid) return;
	fprintf(g_bolid, "%.9g;%.9g;%.9g;%.9g;%.9g;%d;%d\n", msg.noise, msg.peakFrequency, msg.magnitude,
	        msg.minFreq, msg.maxFreq, msg.startSample, msg.endSample);
	fflush(g_bolid);
}

Pipeline *g_pipeline = NULL;

void onInterrupt(int)
{
	// App::interruptHandler (src/App.cpp:251-256)
	if (g_pipeline) g_pipeline->stop();
}

void onTerminate(int)
{
	// App::termHandler (src/App.cpp:259-263): EXIT_TERM_RECEIVED
	_exit(1);
}

double wallSeconds()
{
	struct timeval tv;
	gettimeofday(&tv, NULL);
	return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

int usage(const char *argv0)
{
	fprintf(stderr, "usage: %s -c CONFIG.json [-n NAME] [-q] [-N] [-d DIR [-r]] [-v] [WAV_FILE]\n", argv0);
	return 2;
}

}  // namespace


int main(int argc, char **argv)
{
	std::string configFile, cfgName, dumpDir, wavFile;
	bool dumpRows = false, quiet = false, narrow = false;
	for (int i = 1; i < argc; i++) {
		const std::string a = argv[i];
		if (a == "-c" && i + 1 < argc) configFile = argv[++i];
		else if (a == "-n" && i + 1 < argc) cfgName = argv[++i];
		else if (a == "-d" && i + 1 < argc) dumpDir = argv[++i];
		else if (a == "-r") dumpRows = true;
		else if (a == "-q") quiet = true;
		else if (a == "-N") narrow = true;
		else if (a == "-v") { printf("%s (radio-observer-cuda, %s)\n", PACKAGE_STRING, ro_version()); return 0; }
		else if (a == "-h" || (a.size() > 1 && a[0] == '-')) return usage(argv[0]);
		else wavFile = a;
	}
	if (configFile.empty()) {
		const char *home = getenv("HOME");
		configFile = std::string(home ? home : ".") + "/.radio-observer.json";     // App::getDefaultConfigFile
	}

	Logger::clearConfig();
	Logger::addOutput(quiet ? LOG_LVL_WARNING : LOG_LVL_INFO, "-");

	Ref<FileInput> configInput = new FileInput(configFile);
	if (!configInput->exists()) {
		LOG_ERROR("Configuration file " << configFile << " does not exist.");
		return 2;                                   // EXIT_NO_CONFIG
	}
	JSONParser parser;
	Ref<DynObject> config = parser.parse(configInput);
	configInput->close();
	if (config->isError()) {
		Ref<DynError> err = config.as<DynError>();
		LOG_ERROR("Syntax error in config file at " << err->getErrorLoc() << ": " << err->getString());
		return 1;
	}

	Ref<DynObject> loggingConfig = config->getStrItem("logging");
	if (!quiet && !loggingConfig.isNull()) {
		Logger::clearConfig();
		Logger::readConfig(loggingConfig);
	}
	LOG_INFO("***** Starting Radio Observer v" PACKAGE_VERSION " with " << ro_version() << " *****");

	if (cfgName.empty()) cfgName = config->getStrString("configuration", "default");
	Injector::getInstance().makePlans(config->getStrItem("configurations"));
	Ref<Pipeline> pipeline = Injector::getInstance().instantiateAs<Pipeline>(cfgName);
	if (pipeline.isNull()) {
		LOG_ERROR("Initialization failed.");
		return 3;                                   // EXIT_INIT_FAILED
	}

	// ---- frontend: App::createFrontend ----
	if (pipeline->getFrontend().isNull()) {
		if (!wavFile.empty()) {
			LOG_INFO("Using WAV frontend, reading " << wavFile << "...");
			if (narrow) pipeline->setFrontend(new NarrowWAVStream(new FileInput(wavFile)));
			else pipeline->setFrontend(new WAVStream(new FileInput(wavFile)));
		} else {
			const std::string name = config->getStrString("frontend", "jack");
			if (name == "tcp_raw") {
				LOG_INFO("Using raw TCP frontend.");
				pipeline->setFrontend(new RawTCPStream(config->getStrString("tcp_host", "localhost"),
				                                       config->getStrInt("tcp_port", 4000),
				                                       config->getStrInt("raw_sample_rate", 96000)));
			} else {
				LOG_ERROR("No frontend to use: \"" << name << "\" is not built into radio-observer-cuda "
				          "(give a WAV file or set \"frontend\": \"tcp_raw\").");
				return 1;
			}
		}
	}

	// ---- diagnostics ----
	Ref<WaterfallBackend> backend = pipeline->getBackend().as<WaterfallBackend>();
	CudaWaterfallBackend *cuda = dynamic_cast<CudaWaterfallBackend *>(backend.getPtr());
	Ref<RowInfoTap> infoTap;
	Ref<RowDumpTap> dumpTap;
	if (!dumpDir.empty() && !backend.isNull()) {
		Ref<Recorder> tap;
		if (dumpRows) { dumpTap = new RowDumpTap(backend, dumpDir); tap = dumpTap; }
		else { infoTap = new RowInfoTap(backend, dumpDir); tap = infoTap; }
		if (cuda) cuda->addTap(tap);
		else backend->addRecorder(tap);
		g_noise = fopen((dumpDir + "/noise.bin").c_str(), "wb");
		g_bolid = fopen((dumpDir + "/bolid.txt").c_str(), "w");
		addListener<NoiseMessage>(onNoise, NULL);
		addListener<BolidMessage>(onBolid, NULL);
	}

	g_pipeline = pipeline.getPtr();
	signal(SIGINT, onInterrupt);
	signal(SIGTERM, onTerminate);
	const double t0 = wallSeconds();
	pipeline->run();
	const double t1 = wallSeconds();
	signal(SIGINT, SIG_DFL);
	signal(SIGTERM, SIG_DFL);
	g_pipeline = NULL;
	LOG_INFO("Exiting.");

	long rows = 0;
	if (!dumpTap.isNull()) { rows = dumpTap->count(); dumpTap->finish(); }
	if (!infoTap.isNull()) { rows = infoTap->count(); infoTap->finish(); }
	if (g_noise) fclose(g_noise);
	if (g_bolid) fclose(g_bolid);

	if (!dumpDir.empty() && !backend.isNull()) {
		char line[640];
		snprintf(line, sizeof line,
		         "{\"rows\": %ld, \"bins\": %d, \"fft_sample_rate\": %.9g, \"sample_rate\": %d, \"run_seconds\": %.6f, "
		         "\"backend\": \"%s\", \"full_rows\": %s, \"band\": [%d, %d]}",
		         rows, backend->getBins(), (double)backend->getFFTSampleRate(), backend->getStreamInfo().sampleRate,
		         t1 - t0, cuda ? "waterfall_cuda" : "waterfall",
		         (!cuda || cuda->deliversFullRows()) ? "true" : "false",
		         cuda ? cuda->deliveredBand().lo : 0, cuda ? cuda->deliveredBand().hi : backend->getBins());
		fprintf(stderr, "%s\n", line);
		if (FILE *f = fopen((dumpDir + "/timing.json").c_str(), "w")) { fprintf(f, "%s\n", line); fclose(f); }
	}
	return 0;
}
