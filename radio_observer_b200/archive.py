"""archive.py -- time-sharded archive reprocessing (BASELINE config 4).

The waterfall path shards by TIME with no exchange between GPUs: frames are independent
(reference src/FFTBackend.cpp:192-279 forms frame k from samples [k*hop, k*hop + bins) and nothing
else), so rank g of G takes frames [floor(g*F/G), floor((g+1)*F/G)) and reads its samples -- plus a
halo of `overlap` samples that the next shard also reads -- straight from the input.  Rows and
detector records are concatenated in frame order on the host.  No data-path collective, no NCCL;
`torch.distributed` (any backend, gloo on the host) is used only to hand the per-rank results to
rank 0.

The order-dependent remainder is O(1) per row and runs once over the merged records:
`BolidTrigger` is the decision and state machine of BolidRecorder::update
(reference src/BolidRecorder.cpp:119-299) in GLOBAL row numbers.

Host-side Hz<->bin arithmetic follows reference src/FFTBackend.h:134-200 with its float/double
mixture (the band edges decide which bins the detector sees, so they must be bit-exact).

Nothing here imports oracle/: the compute callable defaults to the CUDA library.
"""
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import os

import numpy as np

f32 = np.float32


# ---------------------------------------------------------------------------
# A9  Hz <-> bin (reference src/FFTBackend.h:134-200, src/FFTBackend.cpp:150-151)
# ---------------------------------------------------------------------------
def frequency_to_bin(freq: float, sample_rate: int, bins: int) -> int:
    """clamp((int)(N_f * ((double)(f / sr_f) + 0.5)), 0, N-1); f/sr is a FLOAT division."""
    q = f32(f32(freq) / f32(sample_rate))
    v = float(f32(bins)) * (float(q) + 0.5)
    return min(max(int(v), 0), bins - 1)


def bin_to_frequency(b: int, sample_rate: int, bins: int) -> np.float32:
    """(float)(sr_f * (-0.5 + (double)(b_f / N_f)))."""
    return f32(float(f32(sample_rate)) * (-0.5 + float(f32(f32(b) / f32(bins)))))


def fft_sample_rate(sample_rate: int, bins: int, overlap: int) -> np.float32:
    """rows per second: (float)sr / (float)(bins - overlap)."""
    return f32(f32(sample_rate) / f32(bins - overlap))


def time_to_fft_samples(t: float, fsr: np.float32) -> int:
    """reference src/FFTBackend.h:197-200: `return time * fftSampleRate_` (double * float -> int)."""
    return int(float(t) * float(fsr))


def clamp_overlap(bins: int, overlap: int) -> int:
    """reference src/FFTBackend.cpp:108-109."""
    return min(max(int(overlap), 0), bins - 1)


def frame_count(n_samples: int, bins: int, overlap: int) -> int:
    hop = bins - clamp_overlap(bins, overlap)
    return 0 if n_samples < bins else (n_samples - bins) // hop + 1


def bolid_bands(sample_rate: int, bins: int, overlap: int, *, low_detect_freq: float, hi_detect_freq: float,
                low_noise_freq: float, hi_noise_freq: float, avg_freq_range: float,
                advance_time: float = 1.0, jitter_time: float = 1.0) -> dict:
    """The bin values BolidRecorder::start derives from its JSON keys
    (reference src/BolidRecorder.cpp:80-103): detector config for `Waterfall(detector=...)`
    plus the advance / jitter row counts of the state machine."""
    dlo, dhi = sorted((frequency_to_bin(low_detect_freq, sample_rate, bins),
                       frequency_to_bin(hi_detect_freq, sample_rate, bins)))
    nlo, nhi = sorted((frequency_to_bin(low_noise_freq, sample_rate, bins),
                       frequency_to_bin(hi_noise_freq, sample_rate, bins)))
    avg = frequency_to_bin(avg_freq_range, sample_rate, bins) - frequency_to_bin(0.0, sample_rate, bins)
    fsr = fft_sample_rate(sample_rate, bins, clamp_overlap(bins, overlap))
    return dict(det_lo=dlo, det_w=dhi - dlo, noise_lo=nlo, noise_w=nhi - nlo, avg_bins=avg,
                advance=time_to_fft_samples(advance_time, fsr), jitter=time_to_fft_samples(jitter_time, fsr))


# ---------------------------------------------------------------------------
# 8e  shard plan
# ---------------------------------------------------------------------------
@dataclass(frozen=True)
class Shard:
    rank: int
    frame0: int        # first global frame (row) of the shard
    frame1: int        # one past the last
    sample0: int       # first input sample the shard reads
    sample1: int       # one past the last (includes the overlap halo)

    @property
    def n_frames(self) -> int:
        return self.frame1 - self.frame0

    @property
    def n_samples(self) -> int:
        return self.sample1 - self.sample0


def plan_shards(n_samples: int, bins: int, overlap: int, world: int) -> List[Shard]:
    """Contiguous frame ranges, one per rank; an empty shard has sample0 == sample1."""
    if world < 1:
        raise ValueError("world must be >= 1")
    ov = clamp_overlap(bins, overlap)
    hop = bins - ov
    F = frame_count(n_samples, bins, ov)
    out = []
    for g in range(world):
        f0, f1 = (g * F) // world, ((g + 1) * F) // world
        if f1 > f0:
            out.append(Shard(g, f0, f1, f0 * hop, (f1 - 1) * hop + bins))
        else:
            out.append(Shard(g, f0, f0, f0 * hop, f0 * hop))
    return out


# ---------------------------------------------------------------------------
# A14  detector decision + state machine over merged records
# ---------------------------------------------------------------------------
@dataclass
class BolidEvent:
    emit_row: int       # global row at which the event was emitted (jitter expired)
    start_row: int      # first row of the snapshot: (row of first detection + 1) - advance; may be < 0
    length: int         # rows: 2*advance + rows between first detection and the last gap
    peak_bin: int       # row index of the peak of the FIRST detecting row (det_lo + peak)
    noise: np.float32   # of the first detecting row
    magnitude: np.float32
    duration_s: np.float32


@dataclass
class BolidTrigger:
    """BolidRecorder::update (reference src/BolidRecorder.cpp:119-299) on (noise, peak, avg)
    records; `row` numbers are global, the ring index the reference uses is row mod capacity.
    The threshold is the reference's hard-coded 2.0 (`threshold` JSON key is inert, :135)."""
    det_lo: int
    advance: int
    jitter: int
    fft_sample_rate: np.float32
    events: List[BolidEvent] = field(default_factory=list)
    _state: int = 0     # 0 INIT, 1 BOLID, 2 ENDED
    _duration: int = 0
    _row: int = 0
    _start: int = 0
    _length: int = 0
    _first: Optional[Tuple[int, np.float32, np.float32]] = None

    def feed(self, det: np.ndarray) -> None:
        """det: structured records (noise, peak, avg) of consecutive rows, in row order."""
        noise = np.asarray(det["noise"], dtype=np.float32)
        avg = np.asarray(det["avg"], dtype=np.float32)
        peak = np.asarray(det["peak"])
        hit = avg.astype(np.float64) > noise.astype(np.float64) * 2.0
        for i in range(len(hit)):
            r = self._row
            self._row += 1
            if self._state == 0:
                if hit[i]:
                    self._first = (self.det_lo + int(peak[i]), noise[i], avg[i])
                    self._duration = 1
                    self._start = (r + 1) - self.advance       # mark() is the index of the NEXT row
                    self._length = 2 * self.advance
                    self._state = 1
            elif self._state == 1:
                if hit[i]:
                    self._duration += 1
                else:
                    self._length += self._duration
                    self._duration = 1
                    self._state = 2
            else:
                self._duration += 1
                if hit[i]:
                    self._state = 1
                elif self._duration >= self.jitter:
                    pb, n0, a0 = self._first
                    dur = f32(f32(self._length - 2 * self.advance) / f32(self.fft_sample_rate))
                    self.events.append(BolidEvent(r, self._start, self._length, pb, n0, a0, dur))
                    self._state = 0


# ---------------------------------------------------------------------------
# driver: one rank's shard, then the merge on rank 0
# ---------------------------------------------------------------------------
def default_compute(bins: int, overlap: int, sample_rate: int, bands: Sequence[Tuple[int, int]],
                    detector: Optional[dict], device: int) -> Callable:
    """The product compute: ro_wf_process_host on `device` (fails loudly without the CUDA library)."""
    from .capi import Waterfall, IN_INT16
    det = None if detector is None else {k: detector[k] for k in ("det_lo", "det_w", "noise_lo", "noise_w", "avg_bins")}
    wf = Waterfall(bins, overlap, sample_rate=sample_rate, in_format=IN_INT16, bands=list(bands), detector=det,
                   device=device)

    def run(iq_int16: np.ndarray):
        rows, recs, n = wf.process_host(iq_int16)
        return rows[:n], (None if recs is None else recs[:n])
    return run


def staged_compute(bins: int, overlap: int, sample_rate: int, bands: Sequence[Tuple[int, int]],
                   detector: Optional[dict], device: int, *, chunk_frames: int = 16384, copy_threads: int = 4) -> Callable:
    """The product compute for an archive that is a memory map of a file (page cache, tmpfs): the shard is
    walked in chunks of `chunk_frames` frames (plus the overlap halo); `copy_threads` threads copy chunk
    c + 1 from the map into one of two PINNED staging buffers while the GPU works on chunk c out of the
    other (ro_wf_process_host: async H2D on a side stream, kernels, D2H of the band rows and records).
    Handing the map itself to the library makes every H2D copy go through the driver's pageable bounce
    buffer on one thread: 1.9 Gsample/s for a 24 h archive on one GPU against the 13.6 of pinned input.

    run(iq_map, sample0, sample1, rows_out, det_out) fills rows_out / det_out (this shard's slices)."""
    import threading
    from .capi import Waterfall, IN_INT16, PinnedArray
    det = None if detector is None else {k: detector[k] for k in ("det_lo", "det_w", "noise_lo", "noise_w", "avg_bins")}
    wf = Waterfall(bins, overlap, sample_rate=sample_rate, in_format=IN_INT16, bands=list(bands), detector=det,
                   device=device)
    hop = bins - clamp_overlap(bins, overlap)
    chunk_samples = (chunk_frames - 1) * hop + bins
    stage = [PinnedArray((2 * chunk_samples,), np.int16) for _ in range(2)]

    def fill(buf: np.ndarray, iq_map: np.ndarray, e0: int, e1: int):
        """buf[:e1 - e0] = iq_map[e0:e1] (int16 elements).  A numpy memmap of a file is read with
        os.preadv straight into the pinned buffer, one slice per thread: copying out of the mapping
        instead takes a minor page fault per 4 KB under the process-wide mm lock and stays at the speed
        of one thread (23 GB/s measured on the B200 box, whatever the thread count)."""
        n = e1 - e0
        fn = getattr(iq_map, "filename", None)
        if isinstance(iq_map, np.memmap) and fn is not None and iq_map.dtype == np.int16 and iq_map.ndim == 1:
            base = int(iq_map.offset) + 2 * e0
            nbytes = 2 * n
            raw = memoryview(buf).cast("B")
            parts = max(1, copy_threads)
            step = -(-nbytes // parts)
            step += (-step) % 4096

            def rd(a):
                fd = os.open(fn, os.O_RDONLY)
                try:
                    at, end = a, min(nbytes, a + step)
                    while at < end:
                        got = os.preadv(fd, [raw[at:end]], base + at)
                        if got <= 0:
                            raise IOError("short read from %s" % fn)
                        at += got
                finally:
                    os.close(fd)
            ths = [threading.Thread(target=rd, args=(a,)) for a in range(0, nbytes, step)]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            return
        src = iq_map[e0:e1]
        if copy_threads <= 1 or n < (1 << 22):
            buf[:n] = src
            return
        step = -(-n // copy_threads)
        step += step & 1
        ths = [threading.Thread(target=lambda a=a: buf.__setitem__(slice(a, min(n, a + step)), src[a:min(n, a + step)]))
               for a in range(0, n, step)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()

    def run(iq_map: np.ndarray, sample0: int, sample1: int, rows_out: np.ndarray, det_out: Optional[np.ndarray]) -> int:
        n_frames = frame_count(sample1 - sample0, bins, overlap)
        spans = []
        for f0 in range(0, n_frames, chunk_frames):
            f1 = min(n_frames, f0 + chunk_frames)
            spans.append((f0, f1, sample0 + f0 * hop, sample0 + (f1 - 1) * hop + bins))
        if not spans:
            return 0
        pending = None

        def load(i):
            f0, f1, s0, s1 = spans[i]
            fill(stage[i & 1].array, iq_map, 2 * s0, 2 * s1)
        load(0)
        for i, (f0, f1, s0, s1) in enumerate(spans):
            if i + 1 < len(spans):
                pending = threading.Thread(target=load, args=(i + 1,))
                pending.start()
            wf.process_host(stage[i & 1].array[:2 * (s1 - s0)], rows_out[f0:f1], None if det_out is None else det_out[f0:f1])
            if pending is not None:
                pending.join()
                pending = None
        return n_frames
    run.close = wf.close
    run.row_width = wf.row_width
    return run


def reprocess_archive(iq: np.ndarray, bins: int, overlap: int, *, sample_rate: int = 96000,
                      bands: Sequence[Tuple[int, int]] = (), detector: Optional[dict] = None,
                      rank: int = 0, world: int = 1, dist=None, device: Optional[int] = None,
                      compute: Optional[Callable] = None, shared_dir: Optional[str] = None, staged: Optional[Callable] = None):
    """Process this rank's time shard of `iq` (interleaved int16 [I,Q], e.g. a numpy memmap of the
    archive: only the shard's slice is read) and merge on rank 0.

    Returns (rows, det, shards) on rank 0 -- rows of all shards concatenated in frame order -- and
    (None, None, shards) elsewhere.

    Merge, two ways.  With `shared_dir` (a directory every rank of the box sees, e.g. /dev/shm) rank 0
    creates the merged arrays as files there, every rank maps them and writes its own frame range in
    place: nothing is gathered, `dist` is used for two barriers only (`staged` = a staged_compute()
    callable then does the work).  Without it the finished rows travel to rank 0 as host objects over
    `dist` (gather_object: pickled; fine for tests and small archives).  Either way the data path itself
    has no collective."""
    n_samples = iq.shape[0] // 2
    shards = plan_shards(n_samples, bins, overlap, world)
    me = shards[rank]
    if shared_dir is not None:
        if staged is None:
            staged = staged_compute(bins, overlap, sample_rate, bands, detector, rank if device is None else device)
        from .capi import DET_DTYPE
        F = frame_count(n_samples, bins, overlap)
        W = staged.row_width
        rows_path, det_path = os.path.join(shared_dir, "ro_rows_%d.f32" % F), os.path.join(shared_dir, "ro_det_%d.rec" % F)
        if rank == 0:
            np.lib.format.open_memmap(rows_path, mode="w+", dtype=np.float32, shape=(F, W)).flush()
            if detector is not None:
                np.lib.format.open_memmap(det_path, mode="w+", dtype=DET_DTYPE, shape=(F,)).flush()
        if dist is not None:
            dist.barrier()
        rows = np.load(rows_path, mmap_mode="r+")
        det = np.load(det_path, mmap_mode="r+") if detector is not None else None
        if me.n_frames > 0:
            done = staged(iq, me.sample0, me.sample1, rows[me.frame0:me.frame1], None if det is None else det[me.frame0:me.frame1])
            if done != me.n_frames:
                raise RuntimeError(f"rank {rank}: shard of {me.n_frames} frames produced {done} rows")
        if dist is not None:
            dist.barrier()
        if rank != 0:
            return None, None, shards
        return rows, det, shards
    if compute is None:
        compute = default_compute(bins, overlap, sample_rate, bands, detector, rank if device is None else device)
    if me.n_frames > 0:
        rows, det = compute(np.ascontiguousarray(iq[2 * me.sample0:2 * me.sample1]))
        if rows.shape[0] != me.n_frames:
            raise RuntimeError(f"rank {rank}: shard of {me.n_frames} frames produced {rows.shape[0]} rows")
    else:
        rows, det = None, None
    if world == 1:
        return rows, det, shards
    parts = [None] * world if rank == 0 else None
    dist.gather_object((rows, det), parts, dst=0)
    if rank != 0:
        return None, None, shards
    have = [p for p in parts if p[0] is not None]
    all_rows = np.concatenate([p[0] for p in have]) if have else None
    all_det = np.concatenate([p[1] for p in have]) if have and have[0][1] is not None else None
    return all_rows, all_det, shards


# ---------------------------------------------------------------------------
# 8f N4  event text formats (reference src/BolidRecorder.cpp:215-257, src/CsvLog.cpp:12-58)
# ---------------------------------------------------------------------------
def _g(x) -> str:
    """operator<< of a float / double on a default ostream: %g with 6 significant digits."""
    return "%g" % float(x)


def fft_samples_to_raw(count: int, fft_sample_rate_: np.float32, sample_rate: int) -> int:
    """Recorder::fftSamplesToRaw with getFFTSampleRate() truncated to int
    (reference src/WaterfallBackend.cpp:29-32, src/WaterfallBackend.h:84-87)."""
    rate = int(fft_sample_rate_)
    if rate == 0:
        return 0            # the reference divides by zero here
    return int((float(count) / float(rate)) * float(sample_rate))


def event_fields(ev: "BolidEvent", *, sample_rate: int, bins: int, overlap: int,
                 low_detect_freq: float, hi_detect_freq: float) -> dict:
    """What BolidRecorder reports for an event: peak frequency, the +-quarter-band frequency window
    (float arithmetic), duration in seconds and the raw-sample length of the snapshot."""
    lo, hi = sorted((f32(low_detect_freq), f32(hi_detect_freq)))
    pf = bin_to_frequency(ev.peak_bin, sample_rate, bins)
    q = f32(f32(hi - lo) / f32(4))
    fsr = fft_sample_rate(sample_rate, bins, clamp_overlap(bins, overlap))
    return dict(noise=f32(ev.noise), peak_freq=pf, magnitude=f32(ev.magnitude), min_freq=f32(pf - q), max_freq=f32(pf + q),
                duration=f32(ev.duration_s), end_sample=fft_samples_to_raw(ev.length, fsr, sample_rate))


def bolid_csv_line(file_basename: str, fields: dict) -> str:
    """A line of the hourly *_meta.csv (header "# file name; noise; peak f.; mag.; duration")."""
    return "%s;%s;%s;%s;%s" % (file_basename, _g(fields["noise"]), _g(fields["peak_freq"]), _g(fields["magnitude"]),
                               _g(fields["duration"]))


def met_line(t_sec: int, t_usec: int, fields: dict) -> str:
    """The stdout line station tooling parses: met;[<s>s, <us>us];noise;peak;mag;fmin;fmax;duration;end#"""
    return "met;[%ds, %dus];%s;%s;%s;%s;%s;%s;%d#" % (
        t_sec, t_usec, _g(fields["noise"]), _g(fields["peak_freq"]), _g(fields["magnitude"]), _g(fields["min_freq"]),
        _g(fields["max_freq"]), _g(fields["duration"]), fields["end_sample"])
