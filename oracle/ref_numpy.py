"""oracle/ref_numpy.py -- TEST INFRASTRUCTURE, not product code.

NumPy restatement (float64 FFT, explicit float32 casts where the reference
rounds) of the reference's spectral hot path.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.

Parity status: PINNED against the reference itself -- oracle/_ref/ref_oracle is
the reference's own C++ (compiled in place from /root/reference) and
tests/test_oracle_vs_ref.py + tests/golden/ check this restatement against its
outputs.  The reference holds no golden vectors of its own for this path
(SURVEY.md 4), and its FFT lives in FFTW3 (unpinned version, not vendored):
here numpy.fft (pocketfft, double) stands in, which differs from any
double-precision FFT by O(1e-15 log N) relative.

All `file:line` citations are relative to /root/reference/.
"""
from __future__ import annotations

import math
import time as _time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

f32 = np.float32
f64 = np.float64


# --------------------------------------------------------------------------
# A4  window table                                   src/FFTBackend.cpp:155-186
# --------------------------------------------------------------------------
def nuttall_window(bins: int) -> np.ndarray:
    """4-term Nuttall table exactly as FFTBackend::startStream builds it:
    float coefficients promoted to double, angle = 2.0*PI*(float)i/(float)(N-1)
    evaluated left to right in double, result rounded once to float."""
    a0, a1, a2, a3 = (f64(f32(0.355768)), f64(f32(0.487396)),
                      f64(f32(0.144232)), f64(f32(0.012604)))
    pi = 4.0 * math.atan(1.0)
    i = np.arange(bins, dtype=np.float64)          # (float)i is exact for i < 2^24
    d = f64(f32(bins - 1))
    w = (a0
         - a1 * np.cos(2.0 * pi * i / d)
         + a2 * np.cos(4.0 * pi * i / d)
         - a3 * np.cos(6.0 * pi * i / d))
    return w.astype(np.float32)


# --------------------------------------------------------------------------
# A9  Hz <-> bin, rows <-> seconds                   src/FFTBackend.h:134-200
# --------------------------------------------------------------------------
def fft_sample_rate(sample_rate: int, bins: int, overlap: int) -> np.float32:
    """src/FFTBackend.cpp:150-151."""
    return f32(f32(sample_rate) / f32(bins - overlap))


def clamp_overlap(bins: int, overlap: int) -> int:
    """src/FFTBackend.cpp:108-109."""
    if overlap < 0:
        overlap = 0
    if overlap >= bins:
        overlap = bins - 1
    return overlap


def bin_to_frequency(b: int, sample_rate: int, bins: int) -> np.float32:
    """src/FFTBackend.h:134-147: `sr * (-0.5 + b / n)` with float b, sr, n; the
    division is float, the sum and product are double, the return is float."""
    bf, sr, n = f32(b), f32(sample_rate), f32(bins)
    q = f32(bf / n)
    return f32(f64(sr) * (-0.5 + f64(q)))


def bin_width(sample_rate: int, bins: int) -> np.float32:
    """src/FFTBackend.h:149-157: binToFrequency(1) - binToFrequency(0), in float."""
    return f32(bin_to_frequency(1, sample_rate, bins) - bin_to_frequency(0, sample_rate, bins))


def frequency_to_bin(freq, sample_rate: int, bins: int) -> int:
    """src/FFTBackend.h:159-178: `int bin = n * ((frequency / sr) + 0.5)`;
    frequency/sr is a float division, the rest double, truncated toward zero,
    clamped to [0, bins-1]."""
    fq, sr, n = f32(freq), f32(sample_rate), f32(bins)
    q = f32(fq / sr)
    v = f64(n) * (f64(q) + 0.5)
    b = int(v)  # C truncation toward zero
    if b < 0:
        return 0
    if b >= bins:
        return bins - 1
    return b


def time_to_fft_samples(t: float, fsr: np.float32) -> int:
    """src/FFTBackend.h:197-200: `return time * fftSampleRate_` (double*float -> int)."""
    return int(f64(t) * f64(fsr))


# --------------------------------------------------------------------------
# WFTime                                              src/WFTime.h:92-119
# --------------------------------------------------------------------------
US = 1000000


@dataclass(frozen=True)
class WFTime:
    sec: int = 0
    usec: int = 0

    def add(self, seconds: int, microseconds: int) -> "WFTime":
        """src/WFTime.h:92-103 (non-negative operands on this path)."""
        usec = self.usec + microseconds % US
        sec = self.sec + seconds
        sec += microseconds // US
        sec += usec // US
        usec %= US
        return WFTime(sec, usec)

    def add_samples(self, count: int, sample_rate: int) -> "WFTime":
        """src/WFTime.h:105-114: whole seconds exactly, remainder truncated to us."""
        seconds = count // sample_rate
        remainder = count % sample_rate
        microseconds = int((f64(remainder) / f64(sample_rate)) * f64(US))
        return self.add(seconds, microseconds)

    def to_milliseconds(self) -> float:
        """src/WFTime.h:56-60."""
        return float(f64(self.sec) * 1000.0 + f64(self.usec) / 1000.0)

    def format(self, fmt: str) -> str:
        """src/WFTime.cpp:19-35 (UTC)."""
        return _time.strftime(fmt, _time.gmtime(self.sec))


def frame_times(n_samples: int, bins: int, overlap: int, sample_rate: int,
                      block: int, t0: WFTime = WFTime()) -> List[WFTime]:
    """A7: stream time of every frame's first sample, reproducing the chain of
    truncating addSamples() calls in Frontend::process (src/Frontend.cpp:47-52)
    and FFTBackend::process (src/FFTBackend.cpp:208-256, 261-270), which depends
    on the frontend's block size (1024 WAV, 4096 raw).  Only the handles of
    samples k*hop (future windowRaw_[0] entries) are kept."""
    hop = bins - overlap
    handles: Dict[int, WFTime] = {}
    out: List[WFTime] = []
    fill = 0
    frame = 0
    pos = 0
    while pos < n_samples:
        size = min(block, n_samples - pos)
        t = t0.add_samples(pos, sample_rate)
        gpos = pos
        while size >= bins - fill:
            count = bins - fill
            s = -(-gpos // hop) * hop             # first multiple of hop >= gpos
            while s < gpos + count:
                handles[s] = t.add_samples(s - gpos, sample_rate)
                s += hop
            out.append(handles.pop(frame * hop))
            frame += 1
            t = t.add_samples(count, sample_rate)
            gpos += count
            size -= count
            fill = overlap
        if size > 0:
            s = -(-gpos // hop) * hop
            while s < gpos + size:
                handles[s] = t.add_samples(s - gpos, sample_rate)
                s += hop
            fill += size
        pos += min(block, n_samples - pos)
    return out


# --------------------------------------------------------------------------
# A1-A6  samples -> rows
# --------------------------------------------------------------------------
def to_complex(iq: np.ndarray) -> np.ndarray:
    """A1: interleaved int16 (src/WAVStream.cpp:118-121, unscaled) or float32
    (src/RawStream.cpp:59-63) pairs -> complex128."""
    iq = np.asarray(iq)
    if iq.dtype == np.complex128:
        return iq
    v = iq.astype(np.float64)
    return v[0::2] + 1j * v[1::2]


def frame_count(n_samples: int, bins: int, overlap: int) -> int:
    """A3: F = floor((S - N)/h) + 1 (src/FFTBackend.cpp:211-257; the tail is
    dropped at endStream, :282-286)."""
    if n_samples < bins:
        return 0
    return (n_samples - bins) // (bins - overlap) + 1


def waterfall_rows(iq: np.ndarray, bins: int, overlap: int, iq_gain: float = 0.0,
                   first_frame: int = 0, n_frames: Optional[int] = None,
                   window: Optional[np.ndarray] = None) -> np.ndarray:
    """A2-A6: rows[k, j] for frames first_frame .. first_frame+n_frames-1.

    frame k = samples [k*h, k*h+N)                    (src/FFTBackend.cpp:211-247)
    Q += iq_gain (additive, phase shift 0)            (src/FFTBackend.cpp:71-79)
    in[i] = x[i] * (double)w[i]                       (src/FFTBackend.cpp:229-232)
    X = forward DFT, unnormalised, double             (src/FFTBackend.cpp:120,236)
    row[(i + N/2) mod N] = (float)sqrt(re^2 + im^2)   (src/WaterfallBackend.cpp:492-505)
    """
    x = to_complex(iq)
    overlap = clamp_overlap(bins, overlap)
    hop = bins - overlap
    total = frame_count(len(x), bins, overlap)
    if n_frames is None:
        n_frames = total - first_frame
    n_frames = max(0, min(n_frames, total - first_frame))
    w = (nuttall_window(bins) if window is None else window).astype(np.float64)
    rows = np.empty((n_frames, bins), dtype=np.float32)
    half = bins // 2
    batch = max(1, (1 << 22) // bins)
    for b0 in range(0, n_frames, batch):
        b1 = min(n_frames, b0 + batch)
        idx = (np.arange(first_frame + b0, first_frame + b1)[:, None] * hop
               + np.arange(bins)[None, :])
        fr = x[idx]
        if iq_gain != 0.0:
            fr = fr.real + 1j * (fr.imag + iq_gain)
        X = np.fft.fft(fr * w[None, :], axis=1)
        mag = np.sqrt(X.real * X.real + X.imag * X.imag).astype(np.float32)
        rows[b0:b1, half:] = mag[:, :half]
        rows[b0:b1, :half] = mag[:, half:]
    return rows


# --------------------------------------------------------------------------
# A11-A13  detector statistics                       src/BolidRecorder.cpp:302-347
# --------------------------------------------------------------------------
def noise(buf: np.ndarray) -> np.float32:
    """Ascending sort, element [len/4], times 2.0 in double, returned as float
    (src/BolidRecorder.cpp:313-320)."""
    s = np.sort(np.asarray(buf, dtype=np.float32), kind="stable")
    return f32(f64(s[len(s) // 4]) * 2.0)


def peak(buf: np.ndarray) -> int:
    """Index of the maximum, LAST index on ties (`>=`, src/BolidRecorder.cpp:323-335)."""
    buf = np.asarray(buf, dtype=np.float32)
    m = buf.max()
    return int(np.flatnonzero(buf == m)[-1])


def average(buf: np.ndarray) -> np.float32:
    """Sequential double accumulation, double division, float result
    (src/BolidRecorder.cpp:338-347)."""
    acc = 0.0
    for v in np.asarray(buf, dtype=np.float32):
        acc += float(v)
    return f32(acc / float(len(buf)))


@dataclass
class DetectorBands:
    """Bins derived in BolidRecorder::start (src/BolidRecorder.cpp:80-103)."""
    low_detect_bin: int
    detect_width: int
    low_noise_bin: int
    noise_width: int
    avg_bins: int
    advance: int
    jitter: int


def detector_bands(sample_rate: int, bins: int, overlap: int, *,
                   low_detect_freq=10000.0, hi_detect_freq=10900.0,
                   low_noise_freq=9000.0, hi_noise_freq=10000.0,
                   advance_time=1.0, jitter_time=1.0, avg_freq_range=40.0) -> DetectorBands:
    lo_f, hi_f = f32(low_detect_freq), f32(hi_detect_freq)
    if lo_f > hi_f:                                   # ORDER_PAIR, src/BolidRecorder.h:161
        lo_f, hi_f = hi_f, lo_f
    a = frequency_to_bin(lo_f, sample_rate, bins)
    b = frequency_to_bin(hi_f, sample_rate, bins)
    if a > b:
        a, b = b, a
    na = frequency_to_bin(f32(low_noise_freq), sample_rate, bins)
    nb = frequency_to_bin(f32(hi_noise_freq), sample_rate, bins)
    fsr = fft_sample_rate(sample_rate, bins, overlap)
    avg = (frequency_to_bin(f32(avg_freq_range), sample_rate, bins)
           - frequency_to_bin(0, sample_rate, bins))
    return DetectorBands(
        low_detect_bin=a, detect_width=b - a,
        low_noise_bin=min(na, nb), noise_width=max(na, nb) - min(na, nb),
        avg_bins=avg,
        advance=time_to_fft_samples(advance_time, fsr),
        jitter=time_to_fft_samples(jitter_time, fsr))


def detector_row(row: np.ndarray, d: DetectorBands) -> Tuple[np.float32, int, np.float32]:
    """(n, p, a) of BolidRecorder::update (src/BolidRecorder.cpp:121-132)."""
    n = noise(row[d.low_noise_bin:d.low_noise_bin + d.noise_width].copy())
    p = peak(row[d.low_detect_bin:d.low_detect_bin + d.detect_width])
    s = d.low_detect_bin + p - d.avg_bins // 2
    a = average(row[s:s + d.avg_bins])
    return n, p, a


def detect(n: np.float32, a: np.float32) -> bool:
    """`a > (n * 2.0)` in double (src/BolidRecorder.cpp:135); `threshold` is inert."""
    return bool(f64(a) > f64(n) * 2.0)


def detector_rows(rows: np.ndarray, d: DetectorBands):
    """Vectorised (n, p, a) for many rows; same results as detector_row."""
    nb = np.sort(rows[:, d.low_noise_bin:d.low_noise_bin + d.noise_width], axis=1)
    n = (nb[:, d.noise_width // 4].astype(np.float64) * 2.0).astype(np.float32)
    db = rows[:, d.low_detect_bin:d.low_detect_bin + d.detect_width]
    p = (d.detect_width - 1 - np.argmax(db[:, ::-1], axis=1)).astype(np.int32)
    a = np.empty(len(rows), dtype=np.float32)
    for r in range(len(rows)):
        s = d.low_detect_bin + int(p[r]) - d.avg_bins // 2
        acc = 0.0
        for v in rows[r, s:s + d.avg_bins]:
            acc += float(v)
        a[r] = f32(acc / float(d.avg_bins))
    return n, p, a


# --------------------------------------------------------------------------
# Extension stages row_avg / bin_decim (NO reference counterpart: SURVEY.md D2;
# the reference's row rate is set by `overlap` alone, src/FFTBackend.cpp:150-151,
# 241-245, and its only bin reduction is the band crop, src/WaterfallBackend.cpp:
# 368-380).  Definition restated here so the CUDA stage has a checker:
#   out[j][i] = sqrt(mean over r < R, d < D of rows[j*R + r][i*D + d]^2)
# computed in float64, rounded once to float32; an incomplete last group is dropped.
# --------------------------------------------------------------------------
def reduce_rows(rows: np.ndarray, row_avg: int = 1, bin_decim: int = 1) -> np.ndarray:
    R_, D_ = max(1, int(row_avg)), max(1, int(bin_decim))
    F, N = rows.shape
    G = F // R_
    x = rows[:G * R_].astype(np.float64)
    p = (x * x).reshape(G, R_, N // D_, D_).sum(axis=(1, 3)) / float(R_ * D_)
    return np.sqrt(p).astype(np.float32)


# --------------------------------------------------------------------------
# Row ring index arithmetic                           src/RingBuffer.h:428-571
# --------------------------------------------------------------------------
def ring_capacity(bins: int, chunk_bytes: int, min_rows: int) -> int:
    """RingBuffer2D<float>::resize (src/RingBuffer.h:428-457)."""
    row_size = 4 * bins
    chunk_rows = chunk_bytes // row_size
    if chunk_bytes % row_size:
        chunk_rows += 1
    chunk_count = min_rows // chunk_rows
    if min_rows % chunk_rows > 0:
        chunk_count += 1
    return chunk_count * chunk_rows


class RowRing:
    """Index-only model of RingBuffer2D: mark() is the index of the NEXT row
    (src/RingBuffer.h:482-509, pinned by tests/RingBufferTest.h:229-246)."""

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.head = 0
        self.pushed = 0              # global count of rows pushed

    def normalize(self, v: int) -> int:
        """src/RingBuffer.h:361-370."""
        while v < 0:
            v += self.capacity
        return v % self.capacity

    def push(self) -> int:
        r = self.pushed
        self.pushed += 1
        self.head = (self.head + 1) % self.capacity
        return r

    def mark(self) -> int:
        return self.head

    def size(self, start: int, end: Optional[int] = None) -> int:
        """src/RingBuffer.h:543-563; returns capacity when start == end (Q4)."""
        start = self.normalize(start)
        end = self.mark() if end is None else self.normalize(end)
        if end > start:
            return end - start
        return (self.capacity - start) + end


# --------------------------------------------------------------------------
# A10, A14  recorders (host-side sequential logic)
# --------------------------------------------------------------------------
@dataclass
class SnapshotDesc:
    kind: str                 # output_type
    start: int                # ring index as the reference holds it (may be <0 or >= capacity)
    length: int
    file_name: str
    time: WFTime              # fftMarkToTime(start) -> CRVAL2 / DATE-OBS
    global_start: int         # global row number of the first row (None-like -1 if unknown)
    raw: bool = False
    written: bool = True      # set by run_recorders: did the worker thread get to write it (see there)


def file_basename(t: WFTime, origin: str, typ: str, ext: str = "fits") -> str:
    """src/WaterfallBackend.cpp:320-336."""
    return "%s%03d_%s_%s.%s" % (t.format("%Y%m%d%H%M%S"), t.usec // 1000, origin, typ, ext)


class SnapshotModel:
    """SnapshotRecorder (src/WaterfallBackend.cpp:107-127, 339-427)."""

    def __init__(self, *, sample_rate, bins, overlap, origin="debug", output_type="snap",
                 snapshot_length=60, low_freq=0.0, hi_freq=0.0, write_unfinished=True):
        self.sr, self.bins, self.overlap = sample_rate, bins, clamp_overlap(bins, overlap)
        self.origin, self.output_type = origin, output_type
        self.fsr = fft_sample_rate(sample_rate, bins, self.overlap)
        self.snapshot_length = int(snapshot_length)
        lo, hi = f32(low_freq), f32(hi_freq)
        if lo > hi:
            lo, hi = hi, lo
        self.left_freq, self.right_freq = lo, hi
        self.write_unfinished = write_unfinished
        # requestBufferSize (:339-347): float product, ceil, at least 1
        self.snapshot_rows = max(1, int(math.ceil(f32(f32(self.snapshot_length) * self.fsr))))
        self.out: List[SnapshotDesc] = []

    def request_buffer_size(self) -> int:
        return self.snapshot_rows * 8

    def start(self, ring: RowRing, handles: List[WFTime]):
        self.ring, self.handles = ring, handles
        if self.left_freq == self.right_freq:          # :368-374
            self.left_freq = f32(-f64(f32(self.sr)) / 2.0)
            self.right_freq = f32(f64(f32(self.sr)) / 2.0)
            self.left_bin, self.right_bin = 0, self.bins
        else:
            self.left_bin = frequency_to_bin(self.left_freq, self.sr, self.bins)
            self.right_bin = frequency_to_bin(self.right_freq, self.sr, self.bins)
        self.next_start, self.next_length, self.next_raw = 0, 0, False
        self.next_file = self.file_name(0)

    def mark_to_time(self, mark: int) -> WFTime:
        """Recorder::fftMarkToTime (src/WaterfallBackend.cpp:42-46) incl. the
        mark()-is-next off-by-one (Q3): handles[r+1] holds row r's time."""
        return self.handles[self.ring.normalize(mark)]

    def file_name(self, mark: int) -> str:
        return file_basename(self.mark_to_time(mark), self.origin, self.output_type)

    def global_row(self, start: int) -> int:
        """Global number of the row stored at ring index `start` most recently
        (or about to be), given ring.pushed rows so far."""
        idx = self.ring.normalize(start)
        head_global = self.ring.pushed           # row number the head slot will get
        back = (self.ring.head - idx) % self.ring.capacity
        if back == 0:
            back = self.ring.capacity
        return head_global - back

    def start_writing(self):
        """:107-127."""
        if self.next_length == 0:
            self.next_length = self.ring.size(self.next_start)
        if self.snapshot_rows < self.next_length:
            self.next_length = self.snapshot_rows
        self.out.append(SnapshotDesc(
            kind=self.output_type, start=self.next_start, length=self.next_length,
            file_name=self.next_file, time=self.mark_to_time(self.next_start),
            global_start=self.global_row(self.next_start), raw=self.next_raw))
        self.next_start = self.next_start + self.next_length
        self.next_length, self.next_raw = 0, False
        self.next_file = self.file_name(self.next_start)

    def update(self, row: Optional[np.ndarray] = None):
        """:415-427."""
        if self.ring.size(self.next_start) >= self.snapshot_rows + 2:
            self.start_writing()

    def stop(self):
        """:400-412."""
        if self.ring.size(self.next_start) >= 0 and self.write_unfinished:
            self.start_writing()

    def header(self, snap: SnapshotDesc) -> Dict[str, object]:
        """Header values written by SnapshotRecorder::write (:186-198)."""
        return {
            "NAXIS1": self.right_bin - self.left_bin,
            "NAXIS2": snap.length,
            "CRVAL2": int(snap.time.to_milliseconds()),
            "CDELT2": float(f64(1000.0) / f64(self.fsr)),
            "CRVAL1": float(f32(self.left_freq)),
            "CDELT1": float(bin_width(self.sr, self.bins)),
            "DATE-OBS": snap.time.format("%Y-%m-%dT%H:%M:%S"),
            "ORIGIN": self.origin,
        }


@dataclass
class BolidEvent:
    row: int                  # global row at which the event was emitted
    start: int                # nextSnapshot_.start (ring index, may be negative)
    length: int               # rows, before the snapshotRows_ cap
    file_name: str
    noise: np.float32
    peak_freq: np.float32
    magnitude: np.float32
    duration: np.float32
    end_sample: int


class BolidModel(SnapshotModel):
    """BolidRecorder (src/BolidRecorder.cpp:80-299, src/BolidRecorder.h:126-184)."""
    INIT, BOLID, ENDED = 0, 1, 2

    def __init__(self, *, sample_rate, bins, overlap, origin="debug", output_type="blid",
                 snapshot_length=60, low_freq=9000.0, hi_freq=12000.0,
                 low_detect_freq=10000.0, hi_detect_freq=10900.0,
                 low_noise_freq=9000.0, hi_noise_freq=10000.0,
                 advance_time=1.0, jitter_time=1.0, avg_freq_range=40.0):
        super().__init__(sample_rate=sample_rate, bins=bins, overlap=overlap, origin=origin,
                         output_type=output_type, snapshot_length=snapshot_length,
                         low_freq=low_freq, hi_freq=hi_freq, write_unfinished=False)
        self.min_detect_fq, self.max_detect_fq = sorted((f32(low_detect_freq), f32(hi_detect_freq)))
        self.bands = detector_bands(
            sample_rate, bins, self.overlap,
            low_detect_freq=low_detect_freq, hi_detect_freq=hi_detect_freq,
            low_noise_freq=low_noise_freq, hi_noise_freq=hi_noise_freq,
            advance_time=advance_time, jitter_time=jitter_time, avg_freq_range=avg_freq_range)
        self.events: List[BolidEvent] = []
        self.records: List[Tuple[np.float32, int, np.float32]] = []

    def start(self, ring, handles):
        self.state = self.INIT
        self.duration = 0
        super().start(ring, handles)

    def fft_samples_to_raw(self, count: int) -> int:
        """Recorder::fftSamplesToRaw with getFFTSampleRate() truncated to int (Q6,
        src/WaterfallBackend.cpp:29-32, src/WaterfallBackend.h:84-87)."""
        rate = int(self.fsr)
        if rate == 0:
            return 0   # the reference divides by zero here (inf -> int is UB)
        return int((f64(count) / f64(rate)) * f64(self.sr))

    def update(self, row: np.ndarray = None, npa=None):
        d = self.bands
        n, p, a = detector_row(row, d) if npa is None else npa
        self.records.append((n, p, a))
        det = detect(n, a)
        if self.state == self.INIT:
            if det:
                self.peak_freq = bin_to_frequency(d.low_detect_bin + int(p), self.sr, self.bins)
                self.noise_, self.magnitude = n, a
                self.duration = 1
                self.next_start = self.ring.mark() - d.advance
                self.next_length = 2 * d.advance
                self.next_file = self.file_name(self.next_start)
                self.state = self.BOLID
        elif self.state == self.BOLID:
            if det:
                self.duration += 1
            else:
                self.next_length += self.duration
                self.duration = 1
                self.state = self.ENDED
        else:
            self.duration += 1
            if det:
                self.state = self.BOLID
            elif self.duration >= d.jitter:
                dur = f32(f32(self.next_length - 2 * d.advance) / f32(self.fsr))
                self.events.append(BolidEvent(
                    row=self.ring.pushed - 1, start=self.next_start, length=self.next_length,
                    file_name=self.next_file, noise=self.noise_, peak_freq=self.peak_freq,
                    magnitude=self.magnitude, duration=dur,
                    end_sample=self.fft_samples_to_raw(self.next_length)))
                self.next_raw = True
                self.start_writing()
                self.state = self.INIT


def run_recorders(rows_or_none, n_rows: int, times: Sequence[WFTime], recorders: Sequence[SnapshotModel],
                  bins: int, chunk_bytes: int = 1024 * 1024, npa=None):
    """WaterfallBackend::startStream / processFFT / endStream sequencing
    (src/WaterfallBackend.cpp:573-607, 485-541): size the ring, then per row
    push, store the handle at the NEXT index, update every recorder in order."""
    buffer_size = 1
    for r in recorders:
        buffer_size = max(buffer_size, r.request_buffer_size())
    ring = RowRing(ring_capacity(bins, chunk_bytes, buffer_size))
    handles = [WFTime() for _ in range(ring.capacity)]
    for r in recorders:
        r.start(ring, handles)
    for k in range(n_rows):
        ring.push()
        handles[ring.mark()] = times[k]
        row = None if rows_or_none is None else rows_or_none[k]
        for r in recorders:
            if isinstance(r, BolidModel):
                r.update(row, None if npa is None else (npa[0][k], int(npa[1][k]), npa[2][k]))
            else:
                r.update(row)
    for r in recorders:
        r.stop()
    # SnapshotRecorder::threadMethod (src/WaterfallBackend.cpp:60-104): a queued snapshot is written once
    # buffer_->size(start) >= length and put back otherwise; after Channel::close the queue is drained one
    # last time and what is still incomplete is dropped.  A bolid snapshot reaches `advance` rows past
    # the last detection but is queued only `jitter` rows after it, so an event near the end of the
    # stream is reported and never gets its file.
    for r in recorders:
        for d in r.out:
            d.written = n_rows - d.global_start >= d.length
    return ring
