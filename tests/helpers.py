"""Shared test helpers: tolerance rule, golden loading, tie classification."""
import json
import os

import numpy as np

from oracle import ref_numpy as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Parity tolerance for row amplitudes (north_star: 1e-5
# relative in linear power == 5e-6 in amplitude): per bin
#     |gpu - ref| <= max(5e-6 * |ref|, 1e-6 * RMS(ref row))
# The GPU computes the FFT in FP32; the reference in FP64 (FFTW double).
REL_TOL = 5e-6
ABS_TOL_RMS = 1e-6


def row_errors(gpu: np.ndarray, ref: np.ndarray):
    """Return (worst ratio err/tolerance, max rel err where rel applies)."""
    gpu = np.asarray(gpu, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    rms = np.sqrt(np.mean(ref * ref, axis=-1, keepdims=True))
    tol = np.maximum(REL_TOL * np.abs(ref), ABS_TOL_RMS * rms)
    err = np.abs(gpu - ref)
    return float((err / np.maximum(tol, 1e-300)).max()), float((err / np.maximum(rms, 1e-300)).max())


def assert_rows_close(gpu, ref, what=""):
    assert gpu.shape == ref.shape, (gpu.shape, ref.shape)
    if gpu.size == 0:
        return
    worst, rel_rms = row_errors(gpu, ref)
    assert worst <= 1.0, f"{what}: worst err/tol = {worst:.3f} (max err / row RMS = {rel_rms:.3e})"


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
        rms = np.sqrt(np.mean(row * row))
        gn, gp, ga = float(gpu_det["noise"][r]), int(gpu_det["peak"][r]), float(gpu_det["avg"][r])
        tol_n = 2 * max(REL_TOL * abs(float(rn[r])), 2 * ABS_TOL_RMS * rms)
        ok_n = abs(gn - float(rn[r])) <= tol_n
        if gp == int(rp[r]):
            tol_a = max(REL_TOL * abs(float(ra[r])), ABS_TOL_RMS * rms) * 1.01
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
            tol = 2 * max(REL_TOL * abs(b), ABS_TOL_RMS * rms)
            if abs(a - b) <= tol and ok_n:
                n_tie += 1
            else:
                n_bad += 1
    return n_exact, n_tie, n_bad
