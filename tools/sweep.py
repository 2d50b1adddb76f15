#!/usr/bin/env python3
"""tools/sweep.py -- BASELINE config 5: FFT size 1k..64k x overlap 0/50/75 % on synthetic int16 I/Q.

Times the device-resident hot path (fused FFT kernel; detector off) through the C ABI with the
library's own CUDA-event profiling, and prints one JSON object per shape: frames/s, complex
Msamples/s, algorithmic GB/s (4*hop + 4*N bytes per frame) and its fraction of the
measured HBM peak.  Needs a GPU; nothing here touches oracle/.
  python tools/sweep.py [--log2-samples 26] [--reps 5] [--sizes 10..16]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import radio_observer_b200 as ro                      # noqa: E402
from radio_observer_b200.capi import DeviceBuffer     # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2-samples", type=int, default=26)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--lo", type=int, default=10)
    ap.add_argument("--hi", type=int, default=16)
    ap.add_argument("--fmt", default="int16", choices=["int16", "float32"])
    ap.add_argument("--tuning", type=int, default=0, help="ro_wf_config.tuning bits (A/B switches)")
    ap.add_argument("--scratch-mb", type=int, default=0, help="ro_wf_config.scratch_mb (long-frame path), 0 = default")
    a = ap.parse_args()
    peak = 6650.0
    pp = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pp):
        peak = float(json.load(open(pp))["hbm_gbs"])
    S = 1 << a.log2_samples
    rng = np.random.default_rng(5)
    if a.fmt == "int16":
        iq = rng.integers(-3000, 3000, size=2 * S, dtype=np.int16)
        fmt, bps = ro.IN_INT16, 4
    else:
        iq = rng.standard_normal(2 * S).astype(np.float32)
        fmt, bps = ro.IN_FLOAT32, 8
    d_iq = DeviceBuffer(iq.nbytes)
    d_iq.upload(iq)
    d_rows = DeviceBuffer(4 * (4 * S + (1 << 20)))          # up to 4 rows' worth per hop (75 % overlap)
    for lg in range(a.lo, a.hi + 1):
        N = 1 << lg
        for ov in (0, N // 2, 3 * N // 4):
            wf = ro.Waterfall(N, ov, sample_rate=96000, in_format=fmt, tuning=a.tuning, scratch_mb=a.scratch_mb)
            F = wf.frame_count(S)
            wf.process_resident(d_iq.ptr, S, d_rows.ptr, None, sync=True)          # warm-up
            wf.set_profiling(True)
            wf.kernel_times()
            for _ in range(a.reps):
                wf.process_resident(d_iq.ptr, S, d_rows.ptr, None, sync=False)
            ms, n = wf.kernel_times()
            wf.set_profiling(False)
            t = ms[0] / max(1, n[0]) * 1e-3
            hop = N - ov
            alg = (bps * ((F - 1) * hop + N) + 4 * N * F)
            print(json.dumps({"bins": N, "overlap": ov, "fmt": a.fmt, "frames": F, "kernel_ms": round(t * 1e3, 4),
                              "Mframes_per_s": round(F / t / 1e6, 3), "Msamples_per_s": round(F * hop / t / 1e6, 1),
                              "alg_GBps": round(alg / t / 1e9, 1), "hbm_frac": round(alg / t / 1e9 / peak, 4),
                              "fft_TFLOPs_5NlogN": round(5.0 * N * lg * F / t / 1e12, 2)}), flush=True)
            wf.close()


if __name__ == "__main__":
    main()
