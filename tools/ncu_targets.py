#!/usr/bin/env python3
"""tools/ncu_targets.py WHAT -- small device-resident runs whose kernels `ncu` captures for profiles/:
  detect   16k / 75 %, Bolidozor detector bands: wf_fft2_kernel + wf_detect_kernel
  reduce   16k / 75 %, row_avg 4 x bin_decim 8:   wf_fft2_kernel + wf_reduce_kernel
  size N OV  any transform size (e.g. `size 65536 49152` = wf_split_kernel + wf_fft2_kernel<MODE_SPLIT>)
Needs a GPU; nothing here touches oracle/."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                          # noqa: E402
import radio_observer_b200 as ro                      # noqa: E402
from radio_observer_b200.capi import DeviceBuffer     # noqa: E402


def main():
    what = sys.argv[1]
    S = 1 << 25
    rng = np.random.default_rng(5)
    iq = rng.integers(-3000, 3000, size=2 * S, dtype=np.int16)
    d_iq = DeviceBuffer(iq.nbytes)
    d_iq.upload(iq)
    if what == "size":
        N, ov = int(sys.argv[2]), int(sys.argv[3])
        wf = ro.Waterfall(N, ov, sample_rate=96000)
    elif what == "detect":
        N, ov = 16384, 12288
        band, det = bench.bolidozor_bands()
        wf = ro.Waterfall(N, ov, sample_rate=96000, detector=det)
    elif what == "reduce":
        N, ov = 16384, 12288
        wf = ro.Waterfall(N, ov, sample_rate=96000, row_avg=4, bin_decim=8)
    else:
        raise SystemExit(__doc__)
    F = wf.row_count(S)
    d_rows = DeviceBuffer(4 * (wf.frame_count(S) * N + 1024))
    d_det = DeviceBuffer(16 * (F + 1))
    for _ in range(3):
        wf.process_resident(d_iq.ptr, S, d_rows.ptr, d_det.ptr if what == "detect" else None, sync=True)
    print(what, wf.kernel_name(), "rows", F)
    wf.close()


if __name__ == "__main__":
    main()
