/*
 * NarrowWAVStream.cpp -- see NarrowWAVStream.h.
 */
#include "NarrowWAVStream.h"
#include "CudaWaterfallBackend.h"

#include <cstring>
#include <istream>


NarrowWAVStream::NarrowWAVStream(Ref<Input> input, int readBlocks) :
	input_(input),
	readBlocks_(readBlocks < 1 ? 1 : readBlocks),
	channelCount_(0), blockAlign_(0), bitsPerSample_(0),
	started_(false),
	narrow_(NULL),
	lastBlock_(2 * BLOCK_SAMPLES, 0)
{
}


bool NarrowWAVStream::readBytes(void *dst, size_t n)
{
	std::istream *in = input_->getStream();
	in->read((char *)dst, (std::streamsize)n);
	return (size_t)in->gcount() == n;
}


bool NarrowWAVStream::readFormat(int64_t size)
{
	// the 16 bytes WAVStream::readFormatSubchunk takes (src/WAVStream.cpp:62-83); like the reference
	// the rest of a longer fmt chunk is not skipped
	unsigned char f[16];
	if (!readBytes(f, sizeof f)) return false;
	channelCount_  = (int16_t)(f[2] | (f[3] << 8));
	int sampleRate = (int32_t)((uint32_t)f[4] | ((uint32_t)f[5] << 8) | ((uint32_t)f[6] << 16) | ((uint32_t)f[7] << 24));
	blockAlign_    = (int16_t)(f[12] | (f[13] << 8));
	bitsPerSample_ = (int16_t)(f[14] | (f[15] << 8));
	streamInfo_.sampleRate = sampleRate;
	LOG_INFO("NarrowWAVStream: channelCount=" << channelCount_ << ", sampleRate=" << sampleRate <<
	         ", blockAlign=" << blockAlign_ << ", bitsPerSample=" << bitsPerSample_);
	return true;
}


void NarrowWAVStream::emit(const int16_t *iq, int count)
{
	// Frontend::process (src/Frontend.cpp:41-52) for a block that is still int16
	if (narrow_) {
		narrow_->processRaw(iq, count, dataInfo_);
	} else if (backend_.isNotNull()) {
		wide_.resize(count);
		for (int i = 0; i < count; i++) {
			wide_[i].real = (double)iq[2 * i];
			wide_[i].imag = (double)iq[2 * i + 1];
		}
		backend_->process(wide_, dataInfo_);
	}
	dataInfo_.offset += count;
	dataInfo_.timeOffset = streamInfo_.timeOffset.addSamples(dataInfo_.offset, streamInfo_.sampleRate);
}


bool NarrowWAVStream::readData(int64_t size)
{
	if (bitsPerSample_ != 16) {
		LOG_ERROR("Can only read 16 bits per sample! Stopping now.");
		return false;
	}
	if (channelCount_ != 2 || blockAlign_ != 4) {
		// the reference indexes dataBuffer_[2*sample], [2*sample+1] whatever the header says
		// (src/WAVStream.cpp:118-121): only stereo I/Q is meaningful
		LOG_ERROR("NarrowWAVStream: only 2-channel (I/Q) 16-bit data is supported. Stopping now.");
		return false;
	}

	const int64_t blockBytes = (int64_t)BLOCK_SAMPLES * blockAlign_;
	int64_t       fullBlocks = size / blockBytes;
	const int     remainder  = (int)(size % blockBytes);

	chunk_.resize((size_t)readBlocks_ * 2 * BLOCK_SAMPLES);
	while (fullBlocks > 0) {
		int     want = fullBlocks < readBlocks_ ? (int)fullBlocks : readBlocks_;
		std::istream *in = input_->getStream();
		in->read((char *)&chunk_[0], (std::streamsize)(want * blockBytes));
		int got = (int)(in->gcount() / blockBytes);
		for (int b = 0; b < got; b++)
			emit(&chunk_[(size_t)b * 2 * BLOCK_SAMPLES], BLOCK_SAMPLES);
		if (got > 0)
			memcpy(&lastBlock_[0], &chunk_[(size_t)(got - 1) * 2 * BLOCK_SAMPLES], (size_t)blockBytes);
		if (got < want) return false;      // file shorter than its header says: stop at EOF
		fullBlocks -= got;
	}

	if (remainder > 0) {
		int tail = remainder / blockAlign_;
		if (!readBytes(&lastBlock_[0], (size_t)remainder)) return false;
		int count = tail * channelCount_;                  // the reference's sampleCount (:127-128)
		if (count > BLOCK_SAMPLES) count = BLOCK_SAMPLES;  // what the block buffer holds
		emit(&lastBlock_[0], count);
	}
	return true;
}


void NarrowWAVStream::run()
{
	streamInfo_ = StreamInfo();
	started_ = false;
	narrow_ = dynamic_cast<CudaWaterfallBackend *>(backend_.getPtr());

	char id[5] = { 0, 0, 0, 0, 0 };
	uint32_t size32 = 0;
	if (!readBytes(id, 4) || strcmp(id, "RIFF") != 0) {
		LOG_ERROR("Invalid chunk ID. Stream may not be in WAV format.");
		return;
	}
	if (!readBytes(&size32, 4)) return;
	int64_t left = (int64_t)size32;
	if (!readBytes(id, 4) || strcmp(id, "WAVE") != 0) {
		LOG_ERROR("Invalid chunk format. Stream may not be in WAV format.");
		return;
	}
	left -= 4;
	dataInfo_ = DataInfo();

	while (left > 0) {
		if (!readBytes(id, 4) || !readBytes(&size32, 4)) break;     // EOF
		int64_t size = (int64_t)size32;
		bool ok = true;
		if (strcmp(id, "fmt ") == 0) {
			ok = readFormat(size);
		} else if (strcmp(id, "data") == 0) {
			if (!started_) {
				startStream();
				started_ = true;
			}
			ok = readData(size);
		} else if (strcmp(id, "inf1") == 0) {
			std::string inf1((size_t)size, '\0');
			ok = readBytes(&inf1[0], (size_t)size);
			if (ok) LOG_INFO("INF1 subchunk found: " << inf1.c_str());
		} else {
			input_->getStream()->ignore(size);
		}
		if (!ok) break;
		left -= size + 4;
	}

	if (started_) endStream();
}
