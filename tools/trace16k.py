#!/usr/bin/env python3
"""tools/trace16k.py -- per-warp timeline of the staged 16k kernel (measurement tooling, not product).

Needs the instrumented build of ro_v2.cu (-DRO_TRACE=1, see ro_kernels2.cuh) linked as
radio_observer_b200/csrc/_obj/libro_cuda_trace.so and RO_CUDA_LIB pointing at it.  Lane 0 of every
warp of CTAs 0 and 73 stamps clock64() at 17 phase boundaries of 8 consecutive frames; this script
prints, per phase, how long the warps of each chain level spend in it, and where in the frame (relative
to the CTA-wide barrier after pass 1) each warp passes each boundary.
  RO_CUDA_LIB=.../libro_cuda_trace.so python tools/trace16k.py [--overlap 12288] [--json out.json]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import radio_observer_b200 as ro                      # noqa: E402
from radio_observer_b200.capi import DeviceBuffer     # noqa: E402

NF, NW, NP = 8, 16, 20
NAMES = ["frame start", "samples landed (TMA)", "loaded + windowed", "pass 1 done", "rd3 passed", "pass-1 rows written",
         "barrier A passed", "chain 2 passed", "pass-2 loads issued", "pass 2 done", "rd2 passed", "pass-2 rows written",
         "barrier C passed", "chain 3 passed", "pass-3 loads issued", "pass 3 + magnitudes done", "rows stored"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bins", type=int, default=16384, help="16384: the staged kernel; 65536 ..: the split path's 16k items")
    ap.add_argument("--overlap", type=int, default=12288)
    ap.add_argument("--log2-samples", type=int, default=26)
    ap.add_argument("--tuning", type=int, default=0)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    lib = ro.load_library()
    lib.ro_trace_dump.restype = C.c_int
    lib.ro_trace_dump.argtypes = [C.c_void_p, C.c_size_t]
    S = 1 << a.log2_samples
    iq = np.random.default_rng(5).integers(-3000, 3000, size=2 * S, dtype=np.int16)
    d_iq = DeviceBuffer(iq.nbytes)
    d_iq.upload(iq)
    wf = ro.Waterfall(a.bins, a.overlap, sample_rate=96000, in_format=ro.IN_INT16, tuning=a.tuning)
    F = wf.frame_count(S)
    d_rows = DeviceBuffer(4 * a.bins * F)
    for _ in range(3):
        wf.process_resident(d_iq.ptr, S, d_rows.ptr, None, sync=True)
    wf.set_profiling(True)
    wf.kernel_times()
    for _ in range(5):
        wf.process_resident(d_iq.ptr, S, d_rows.ptr, None, sync=False)
    ms, n = wf.kernel_times()
    print("kernel %s: %.1f us per launch, %.2f M frames/s (instrumented build)" % (wf.kernel_name(), 1e3 * ms[0] / n[0], F / (ms[0] / n[0]) / 1e3))
    buf = np.zeros(2 * NF * NW * NP, dtype=np.uint64)
    rc = lib.ro_trace_dump(buf.ctypes.data, buf.size)
    assert rc == 0, rc
    tr = buf.reshape(2, NF, NW, NP).astype(np.int64)
    out = {}
    for cta in range(2):
        T = tr[cta]
        out["cta%d" % cta] = T.tolist()
        if T.min() == 0:
            print("cta slot %d: not every boundary is stamped in this mode (raw trace in the JSON)" % cta)
            continue
        frame_len = np.diff(T[:, :, 6], axis=0)            # barrier A to barrier A, per warp
        print("\n== CTA slot %d: frame period (barrier A to barrier A) %.0f clk mean, min %d max %d" %
              (cta, frame_len.mean(), frame_len.min(), frame_len.max()))
        # durations of each segment i -> i+1 per chain level (warp w: scheduler w % 4, level w // 4), frames 1..NF-2
        seg = np.diff(T[1:-1], axis=2)                      # [frames, warps, 16]
        print("segment (clk)                               level0  level1  level2  level3")
        for i in range(16):
            lv = [seg[:, l * 4:(l + 1) * 4, i].mean() for l in range(4)]
            print("%-42s %7.0f %7.0f %7.0f %7.0f" % ("%s -> %s" % (NAMES[i][:18], NAMES[i + 1][:18]), *lv))
        nxt = T[2:, :, 0] - T[1:-1, :, 16]                 # rows stored -> next frame start
        print("%-42s %7.0f" % ("rows stored -> next frame start", nxt.mean()))
        # when each warp passes each boundary, relative to the frame's barrier A release (its earliest warp)
        rel = T[1:-1] - T[1:-1, :, 6].min(axis=1)[:, None, None]
        print("\nboundary passed at (clk after barrier A of the frame), mean over frames; warps by level")
        for i in list(range(6, 17)):
            lv = [rel[:, l * 4:(l + 1) * 4, i].mean() for l in range(4)]
            print("%-28s %7.0f %7.0f %7.0f %7.0f" % (NAMES[i], *lv))
        rel2 = T[2:] - T[1:-1, :, 6].min(axis=1)[:, None, None]
        for i in list(range(0, 7)):
            lv = [rel2[:, l * 4:(l + 1) * 4, i].mean() for l in range(4)]
            print("%-28s %7.0f %7.0f %7.0f %7.0f   (next frame)" % (NAMES[i], *lv))
    if a.json:
        json.dump(dict(names=NAMES, trace=out), open(a.json, "w"))


if __name__ == "__main__":
    main()
