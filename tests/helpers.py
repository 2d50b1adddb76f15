"""Shared test helpers: tolerance rule, golden loading, tie classification."""
import json
import os

import numpy as np

from oracle import ref_numpy as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Parity tolerance for row amplitudes, FP32 GPU FFT against the FP64 (FFTW-double) reference.
# north_star allows "relative 1e-5 in linear power, or 1e-3 dB"; per bin we require
#     |gpu - ref| <= max(REL_TOL * |ref|, FLOOR_PEAK * max|ref row|)
# REL_TOL    = 10**(1e-3/20) - 1 = 1.1513e-4   (1e-3 dB in amplitude)
# FLOOR_PEAK = 2**-22                          (four FP32 epsilons of the row's strongest line)
# Why a floor relative to the PEAK and not to the RMS: in an FP32 transform the last butterflies
# add partial sums of magnitude ~peak, so the bins that share a butterfly with a strong carrier
# carry an absolute error of ~1 eps32 * peak whatever their own value (measured: <= 0.8 eps32 *
# peak for N = 1k..512k).  int16 input has 96 dB of dynamic range; the floor sits 132 dB below
# the strongest line.  STRICT_REL (5e-6 = 1e-5 in linear power) is additionally REPORTED as the
# fraction of bins that meet it, and is the gate for the FP64 kernel variant.
REL_TOL = 10.0 ** (1e-3 / 20.0) - 1.0
FLOOR_PEAK = 2.0 ** -22
STRICT_REL = 5e-6
STRICT_ABS_RMS = 1e-6


def row_tolerance(ref: np.ndarray) -> np.ndarray:
    """Per-bin absolute tolerance for rows `ref` (last axis = bins)."""
    ref = np.abs(np.asarray(ref, dtype=np.float64))
    peak = ref.max(axis=-1, keepdims=True) if ref.shape[-1] else ref
    return np.maximum(REL_TOL * ref, FLOOR_PEAK * peak)


def row_errors(gpu: np.ndarray, ref: np.ndarray, peak=None):
    """Return (worst ratio err/tolerance, fraction of bins within the strict 5e-6 / 1e-6*RMS rule).
    `peak`: per-row peak of the FULL row when `ref` is only a crop of it."""
    gpu = np.asarray(gpu, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    tol = row_tolerance(ref)
    if peak is not None:
        tol = np.maximum(REL_TOL * np.abs(ref), FLOOR_PEAK * np.asarray(peak, dtype=np.float64).reshape(-1, 1))
    err = np.abs(gpu - ref)
    rms = np.sqrt(np.mean(ref * ref, axis=-1, keepdims=True))
    strict = np.maximum(STRICT_REL * np.abs(ref), STRICT_ABS_RMS * rms)
    return float((err / np.maximum(tol, 1e-300)).max()), float((err <= strict).mean())


def assert_rows_close(gpu, ref, what="", peak=None):
    assert gpu.shape == ref.shape, (gpu.shape, ref.shape)
    if gpu.size == 0:
        return
    worst, frac_strict = row_errors(gpu, ref, peak)
    print(f"[parity] {what}: worst err/tol = {worst:.3f}, bins within 5e-6 rel / 1e-6 RMS = {100 * frac_strict:.2f} %")
    assert worst <= 1.0, f"{what}: worst err/tol = {worst:.3f} (strict-rule fraction {frac_strict:.4f})"


def assert_rows_strict(gpu, ref, what=""):
    """The FP64 kernel variant: 5e-6 relative (1e-5 in linear power) or 1e-6 * RMS(row)."""
    gpu = np.asarray(gpu, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert gpu.shape == ref.shape, (gpu.shape, ref.shape)
    rms = np.sqrt(np.mean(ref * ref, axis=-1, keepdims=True))
    tol = np.maximum(STRICT_REL * np.abs(ref), STRICT_ABS_RMS * rms)
    worst = float((np.abs(gpu - ref) / np.maximum(tol, 1e-300)).max()) if gpu.size else 0.0
    print(f"[parity-strict] {what}: worst err/tol = {worst:.4f}")
    assert worst <= 1.0, f"{what}: worst err/strict tol = {worst:.3f}"


def load_golden(name):
    d = os.path.join(GOLDEN, name)
    meta = json.load(open(os.path.join(d, "meta.json")))
    z = np.load(os.path.join(d, "input.npz"))
    g = dict(meta=meta, iq=z["iq"], sample_rate=int(z["sample_rate"]),
             config=json.load(open(os.path.join(d, "config.json"))),
             rows_idx=np.load(os.path.join(d, "rows_idx.npy")),
             rows_sub=np.load(os.path.join(d, "rows_sub.npy")),
             bands=np.load(os.path.join(d, "bands.npy")),
             npa=np.load(os.path.join(d, "npa.npy")),
             rowinfo=np.load(os.path.join(d, "rowinfo.npy")),
             bolid=open(os.path.join(d, "bolid.txt")).read(),
             fits=json.load(open(os.path.join(d, "fits.json"))))
    return g


def recorder_cfg(g, factory):
    for r in g["config"]["configurations"][0]["children"][0]["children"]:
        if r["factory"] == factory:
            return r
    return None


def bands_from_cfg(g):
    """DetectorBands for the golden case's bolid recorder."""
    b = recorder_cfg(g, "bolid")
    m = g["meta"]
    return R.detector_bands(
        m["sample_rate"], m["bins"], m["overlap"],
        low_detect_freq=b["low_detect_freq"], hi_detect_freq=b["hi_detect_freq"],
        low_noise_freq=b["low_noise_freq"], hi_noise_freq=b["hi_noise_freq"],
        advance_time=b["advance_time"], jitter_time=b["jitter_time"], avg_freq_range=b["avg_freq_range"])


def det_dict(d: "R.DetectorBands"):
    return dict(det_lo=d.low_detect_bin, det_w=d.detect_width, noise_lo=d.low_noise_bin,
                noise_w=d.noise_width, avg_bins=d.avg_bins)


def classify_det_mismatch(ref_rows, d, gpu_det, ref_npa):
    """Compare GPU detector records (computed on GPU rows) with the oracle's
    (computed on oracle rows).  Differences are allowed only where they are
    explained by the row tolerance: returns (n_exact, n_tie, n_bad).

      noise/avg: must agree within the row tolerance (they are order statistics
                 / means of values that each carry the row tolerance);
      peak:      may differ only if the oracle row's values at the two candidate
                 bins are within twice the row tolerance of each other.
    """
    n_exact = n_tie = n_bad = 0
    rn, rp, ra = ref_npa
    for r in range(len(ref_rows)):
        row = ref_rows[r].astype(np.float64)
        floor = FLOOR_PEAK * np.abs(row).max()
        gn, gp, ga = float(gpu_det["noise"][r]), int(gpu_det["peak"][r]), float(gpu_det["avg"][r])
        # noise = 2 x an order statistic: twice the tolerance of one bin
        tol_n = 2 * max(REL_TOL * abs(float(rn[r])) / 2, floor) * 1.01
        ok_n = abs(gn - float(rn[r])) <= tol_n
        if gp == int(rp[r]):
            tol_a = max(REL_TOL * abs(float(ra[r])), floor) * 1.01
            ok_a = abs(ga - float(ra[r])) <= tol_a
            if ok_n and ok_a:
                if gn == float(rn[r]) and ga == float(ra[r]):
                    n_exact += 1
                else:
                    n_tie += 1          # same decision inputs up to rounding
            else:
                n_bad += 1
        else:
            a = row[d.low_detect_bin + gp]
            b = row[d.low_detect_bin + int(rp[r])]
            tol = 2 * max(REL_TOL * abs(b), floor)
            if abs(a - b) <= tol and ok_n:
                n_tie += 1
            else:
                n_bad += 1
    return n_exact, n_tie, n_bad


# ----------------------------------------------------------------------------
# raw-I/Q snapshot files: the part that does not depend on thread timing
# ----------------------------------------------------------------------------
def raws_head_rows(naxis2, sr, bins, overlap):
    """Leading samples of a `raws` file that were in the raw ring when the snapshot was queued.

    SnapshotRecorder::writeRaw (reference src/WaterfallBackend.cpp:214-267) reads fftSamplesToRaw(rows)
    samples, and the recorder's getFFTSampleRate() is an `int` (src/WaterfallBackend.h:77-87): 5.859
    rows/s becomes 5, so a 65-row snapshot asks for 13 s of samples where the rows span 11.1 s.  The
    tail lies AHEAD of the producer when the snapshot is queued; the writer thread gets there later and
    finds whatever the producer has pushed by then -- the samples, or the ring's initial zeros.  Seen in
    the stock reference under load (first zero row 1146879 .. 1263886 of 1.2-1.3 M), so only the head is
    a fixed function of the input."""
    hop = bins - overlap
    rate = int(sr / hop)
    rows = int(round(naxis2 * rate / float(sr))) if rate > 0 else 0
    return max(0, min(int(naxis2), (rows - 1) * hop))


def assert_raws_equal(a, b, head, what=""):
    """Two `raws` images ([NAXIS2, 2] float32, any shape with 2 values per sample): identical over the
    head; past it a sample may differ only where one side has not been written yet (0, 0)."""
    a = np.asarray(a).reshape(-1, 2)
    b = np.asarray(b).reshape(-1, 2)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert np.array_equal(a[:head], b[:head]), "%s: raw samples differ inside the first %d" % (what, head)
    diff = np.nonzero((a[head:] != b[head:]).any(axis=1))[0]
    if len(diff):
        unwritten = (a[head:][diff] == 0).all(axis=1) | (b[head:][diff] == 0).all(axis=1)
        assert unwritten.all(), "%s: %d raw samples past the head differ and neither side is unwritten" % (what, int((~unwritten).sum()))
    return len(diff)
