"""recorders.py -- the sequential, O(1)-per-row remainder of the recorders for the ARCHIVE path
(8e "sequential remainder"): which rows go into which FITS file, under
which name and with which time stamp.  The streaming path does not use this: there the reference's own
SnapshotRecorder / BolidRecorder run unmodified behind the adapter.

Everything is expressed in GLOBAL row numbers (row k = frame k of the archive).  The reference works
on a ring of rows and a table of per-row handles; as long as the ring is not overrun between the moment
a snapshot is planned and the moment it is written -- it is sized 8 x the longest snapshot for that
reason (reference src/WaterfallBackend.cpp:339-347) -- ring indices and global rows are the same thing,
with two quirks kept:
  * the handle of row r is stored at ring index r+1 (RingBuffer2D::push post-increments, reference
    src/RingBuffer.h:482-509, src/WaterfallBackend.cpp:507), so the time a snapshot starting at row s
    is named and stamped with is the time of row s-1 -- and WFTime(0, 0) when s <= 0, the slot never
    having been written (src/WaterfallBackend.cpp:42-46);
  * a bolid snapshot may start before row 0 (start = first detecting row + 1 - advance).
"""
import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import archive as A
from . import fits as F

f32 = np.float32
US = 1000000


# ---------------------------------------------------------------------------
# A7  stream time of a sample: WFTime::addSamples / add (reference src/WFTime.h:85-119)
# ---------------------------------------------------------------------------
def _add_samples(t: Tuple[int, int], count: int, sample_rate: int) -> Tuple[int, int]:
    sec, rem = divmod(int(count), int(sample_rate))
    micro = int((float(rem) / float(sample_rate)) * float(US))          # truncation toward zero
    usec = t[1] + micro % US
    s = t[0] + sec + micro // US + usec // US
    return (s, usec % US)


def frame_times(n_frames: int, bins: int, overlap: int, sample_rate: int, block: int = 1024,
                t0: Tuple[int, int] = (0, 0)) -> List[Tuple[int, int]]:
    """Time of the first sample of every frame as the reference derives it: the frontend stamps each
    block of `block` samples with t0.addSamples(block offset) (src/Frontend.cpp:47-52; 1024 for WAV,
    4096 for raw TCP), FFTBackend::process re-bases that stamp with addSamples(consumed) after every
    window it completes inside the block and adds the sample's index within the current stretch
    (src/FFTBackend.cpp:208-256); every addSamples truncates to whole microseconds."""
    ov = A.clamp_overlap(bins, overlap)
    hop = bins - ov
    out = []
    for k in range(n_frames):
        s = k * hop                                   # global index of the frame's first sample
        b0 = (s // block) * block
        t = _add_samples(t0, b0, sample_rate)
        pos = b0
        # windows that complete inside this block before sample s: frame j is complete once sample
        # j*hop + bins - 1 is in, i.e. the stretch ends at e_j = j*hop + bins
        j = max(0, -(-(b0 + 1 - bins) // hop))        # first j with e_j > b0
        while True:
            e = j * hop + bins
            if e > s or e > b0 + block:
                break
            t = _add_samples(t, e - pos, sample_rate)
            pos = e
            j += 1
        out.append(_add_samples(t, s - pos, sample_rate))
    return out


# ---------------------------------------------------------------------------
# A10  snapshot slicing (reference src/WaterfallBackend.cpp:107-127, 339-347, 364-427)
# ---------------------------------------------------------------------------
@dataclass
class PlannedFile:
    kind: str                    # output_type ("snap", "blid")
    start_row: int               # global row of the first pixel row (may be < 0 for a bolid)
    length: int                  # rows
    time: Tuple[int, int]        # (sec, usec) used for the name, DATE-OBS and CRVAL2
    file_name: str
    event: Optional[A.BolidEvent] = None


def snapshot_rows(snapshot_length: float, fsr: np.float32) -> int:
    """max(1, ceil(snapshot_length * fftSampleRate)) in float arithmetic (:339-347)."""
    return max(1, int(math.ceil(f32(f32(int(snapshot_length)) * f32(fsr)))))


def _stamp(times: Sequence[Tuple[int, int]], start_row: int) -> Tuple[int, int]:
    return times[start_row - 1] if start_row >= 1 else (0, 0)


def plan_snapshots(n_rows: int, times: Sequence[Tuple[int, int]], *, sample_rate: int, bins: int, overlap: int,
                   snapshot_length: float, origin: str, output_type: str = "snap") -> List[PlannedFile]:
    """SnapshotRecorder: a file of snapshotRows rows whenever that many + 2 rows are waiting, and the
    unfinished rest at endStream."""
    fsr = A.fft_sample_rate(sample_rate, bins, A.clamp_overlap(bins, overlap))
    S = snapshot_rows(snapshot_length, fsr)
    out, start = [], 0
    for pushed in range(1, n_rows + 1):                # update() after every row
        if pushed - start >= S + 2:
            t = _stamp(times, start)
            out.append(PlannedFile(output_type, start, S, t, F.file_basename(t[0], t[1], origin, output_type)))
            start += S
    if n_rows > start:                                 # stop(): write_unfinished
        t = _stamp(times, start)
        out.append(PlannedFile(output_type, start, min(S, n_rows - start), t,
                               F.file_basename(t[0], t[1], origin, output_type)))
    return out


def plan_bolids(events: Sequence[A.BolidEvent], times: Sequence[Tuple[int, int]], *, sample_rate: int, bins: int,
                overlap: int, snapshot_length: float, origin: str, output_type: str = "blid",
                n_rows: Optional[int] = None) -> List[PlannedFile]:
    """BolidRecorder: one file per event, named at the first detection with the stamp of its start row,
    length capped to snapshotRows when it is written (src/BolidRecorder.cpp:171-180, :112-116).

    An event is emitted `jitter` rows after its last detection, but its snapshot runs `advance` rows
    past it: the worker thread keeps the snapshot back until those rows have arrived
    (`buffer_->size(snapshot->start) >= snapshot->length`, src/WaterfallBackend.cpp:77) and drops it if
    the stream ends first (the last pass after Channel::close, :66-104).  With `n_rows` (rows of the
    whole stream) such events get no file, as in the reference; the event itself was still reported."""
    fsr = A.fft_sample_rate(sample_rate, bins, A.clamp_overlap(bins, overlap))
    S = snapshot_rows(snapshot_length, fsr)
    out = []
    for ev in events:
        if n_rows is not None and n_rows - ev.start_row < min(S, ev.length):
            continue
        t = _stamp(times, ev.start_row)
        out.append(PlannedFile(output_type, ev.start_row, min(S, ev.length), t,
                               F.file_basename(t[0], t[1], origin, output_type), ev))
    return out


def write_planned(path: str, plan: PlannedFile, band_rows: np.ndarray, *, left_bin: int, sample_rate: int, bins: int,
                  overlap: int, low_freq: float, origin: str, git_version: str = "", local_time: str = "") -> None:
    """One FITS file: band_rows[r] is global row r cropped to [left_bin, left_bin + width)."""
    if plan.start_row < 0:
        raise ValueError("rows before the start of the archive are not available")
    px = band_rows[plan.start_row:plan.start_row + plan.length]
    cards = F.snapshot_cards(px.shape[1], px.shape[0], origin=origin, time_sec=plan.time[0], time_usec=plan.time[1],
                             fft_sample_rate=A.fft_sample_rate(sample_rate, bins, A.clamp_overlap(bins, overlap)),
                             left_frequency=f32(low_freq),
                             bin_width=A.bin_to_frequency(1, sample_rate, bins) - A.bin_to_frequency(0, sample_rate, bins),
                             git_version=git_version, local_time=local_time)
    F.write_image(path, cards, px)


# ---------------------------------------------------------------------------
# 8f N3  raw-I/Q file of a bolid (reference src/WaterfallBackend.cpp:35-46, 214-267,
#        src/WaterfallBackend.h:84-87, src/FFTBackend.cpp:217-251, src/FFTBackend.h:129-132)
# ---------------------------------------------------------------------------
RAW_CHUNK_ROWS = (1024 * 1024) // 8          # samples per 1 MiB chunk of the raw ring (float pairs)


def raw_ring_capacity(buffer_rows: int, *, sample_rate: int, bins: int, overlap: int) -> int:
    """Samples the reference's raw I/Q ring holds: WaterfallBackend::fftSamplesToRaw(bufferSize) rounded up
    to whole chunks (RingBuffer2D::resize)."""
    fsr = A.fft_sample_rate(sample_rate, bins, A.clamp_overlap(bins, overlap))
    want = int((float(buffer_rows) / float(fsr)) * float(sample_rate))
    return -(-want // RAW_CHUNK_ROWS) * RAW_CHUNK_ROWS


def raw_extent(plan: PlannedFile, *, sample_rate: int, bins: int, overlap: int, ring_capacity: int) -> Tuple[int, int]:
    """(first sample, sample count) of the raw-I/Q file that goes with bolid file `plan`.

    start = fftMarkToRaw(snapshot.start) is the raw mark stored in handle slot `start`, i.e. the one of
    row start - 1 (the slot quirk above): the ring mark after the push of the first sample of the frame
    that FOLLOWS that row when overlap > 0 (FFTBackend::process shifts its handle table before it calls
    processFFT), of that row's own first sample when overlap = 0.  A mark points at the NEXT slot, so the
    file begins one sample later.  A slot never written (start <= 0) holds mark 0.
    count = Recorder::fftSamplesToRaw(length): the row rate truncated to an integer (Q6).
    Valid while the ring has not been lapped between detection and writing (it holds 8 snapshot lengths)."""
    ov = A.clamp_overlap(bins, overlap)
    hop = bins - ov
    fsr = A.fft_sample_rate(sample_rate, bins, ov)
    count = A.fft_samples_to_raw(plan.length, fsr, sample_rate)
    if plan.start_row <= 0:
        return 0, count
    r = plan.start_row - 1
    s = (r + 1) * hop if ov > 0 else r * hop
    # the mark is (s + 1) % ring_capacity and that slot receives sample s + 1
    return s + 1, count


def raw_pixels(iq_int16: np.ndarray, first: int, count: int) -> np.ndarray:
    """[count][2] float32 (I, Q) from interleaved int16 I/Q; samples past the end of the archive read as
    zeros (the reference reads whatever the ring slot held)."""
    n = iq_int16.shape[0] // 2
    out = np.zeros((count, 2), dtype=np.float32)
    a, b = max(0, first), min(n, first + count)
    if b > a:
        out[a - first:b - first] = np.asarray(iq_int16[2 * a:2 * b]).reshape(-1, 2).astype(np.float32)
    return out


def write_raw(path: str, plan: PlannedFile, iq_int16: np.ndarray, *, sample_rate: int, bins: int, overlap: int,
              ring_capacity: int, origin: str, git_version: str = "", local_time: str = "") -> None:
    """The `raws` FITS file of a bolid: named like the bolid file with type "raws" (getFileName("raws", "fits", time))."""
    first, count = raw_extent(plan, sample_rate=sample_rate, bins=bins, overlap=overlap, ring_capacity=ring_capacity)
    cards = F.raw_cards(count, origin=origin, time_sec=plan.time[0], time_usec=plan.time[1], sample_rate=sample_rate,
                        git_version=git_version, local_time=local_time)
    F.write_image(path, cards, raw_pixels(iq_int16, first, count))


def raw_file_name(plan: PlannedFile, origin: str) -> str:
    return F.file_basename(plan.time[0], plan.time[1], origin, "raws")
