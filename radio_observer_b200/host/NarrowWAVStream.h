/*
 * NarrowWAVStream.h -- narrow-ingest WAV frontend.
 *
 * A Frontend (reference src/Frontend.h:20-52) that reads the same 16-bit stereo RIFF/WAVE files as
 * the reference's WAVStream (src/WAVStream.cpp:62-83, 101-142, 231-276) and hands the samples to a
 * CudaWaterfallBackend in the encoding they have in the file -- interleaved int16 [I, Q], 4 B per
 * sample -- through processRaw(), instead of widening each block to a vector<Complex> (16 B per
 * sample) that the backend narrows again.  Any other Backend still gets the widened block, so the
 * class is a stand-in for WAVStream with every backend.
 *
 * Kept from the reference because results depend on it:
 *   - the block cadence: one backend call per 1024 samples, DataInfo advanced by
 *     Frontend::process's rule (src/Frontend.cpp:41-52), which is what the frame-time chain of
 *     FFTBackend::process starts from;
 *   - the tail block: a data chunk that is not a multiple of 4096 bytes ends with a block of
 *     r < 1024 samples for which WAVStream passes on 2r samples, the second half left over from the
 *     previous block (src/WAVStream.cpp:126-139); reproduced for r <= 512 (beyond that the reference
 *     reads past its buffer);
 *   - the chunk walk's accounting (size + 4 per subchunk, src/WAVStream.cpp:178, 234-238).
 * Not kept: reading on after the end of the file (with a standard RIFF size field the reference's
 * accounting runs past EOF and re-processes stale buffers; this reader stops at EOF).
 *
 * Compiled against the reference headers; nothing copied from them.
 */
#ifndef NARROW_WAV_STREAM_H
#define NARROW_WAV_STREAM_H

#include <stdint.h>
#include <string>
#include <vector>

#include <cppapp/Object.h>
#include <cppapp/Input.h>
#include <cppapp/Logger.h>

using namespace cppapp;

#include "Frontend.h"
#include "Backend.h"

class CudaWaterfallBackend;

class NarrowWAVStream : public Frontend {
public:
	enum { BLOCK_SAMPLES = 1024 };   // WAVStream's dataBufferSize_ (src/WAVStream.cpp:189)

private:
	NarrowWAVStream(const NarrowWAVStream &other);

	Ref<Input>            input_;
	int                   readBlocks_;    // blocks fetched per read() call
	int                   channelCount_, blockAlign_, bitsPerSample_;
	bool                  started_;
	CudaWaterfallBackend *narrow_;        // backend_ when it takes int16, else NULL
	std::vector<int16_t>  chunk_;         // readBlocks_ blocks of interleaved int16
	std::vector<int16_t>  lastBlock_;     // WAVStream's dataBuffer_: contents of the last block read
	std::vector<Complex>  wide_;          // only for backends without processRaw

	bool     readBytes(void *dst, size_t n);
	bool     readFormat(int64_t size);
	bool     readData(int64_t size);
	void     emit(const int16_t *iq, int count);

public:
	NarrowWAVStream(Ref<Input> input, int readBlocks = 256);

	virtual void run();
};

#endif /* NARROW_WAV_STREAM_H */
