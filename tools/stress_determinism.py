#!/usr/bin/env python3
"""tools/stress_determinism.py -- race hunt: the same input through the device-resident path many
times; every repetition must give bit-identical rows (prints one sha1 set per shape).  32k frames run
twice, through the fused kernel and (cfg.tuning = RO_TUNE_NO_FUSED) through the split path.  Needs a GPU."""
import hashlib
import sys
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import radio_observer_b200 as ro                      # noqa: E402
from radio_observer_b200.capi import DeviceBuffer     # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    rng = np.random.default_rng(11)
    S = 1 << 24
    iq = rng.integers(-3000, 3000, size=2 * S, dtype=np.int16)
    d_iq = DeviceBuffer(iq.nbytes)
    d_iq.upload(iq)
    ok = True
    for N, ov, tuning in [(4096, 3072, 0), (8192, 6144, 0), (16384, 12288, 0), (16384, 0, 0), (16384, 8192, 0),
                          (32768, 24576, 0), (32768, 24576, ro.TUNE_NO_FUSED), (65536, 49152, 0)]:
        wf = ro.Waterfall(N, ov, sample_rate=96000, tuning=tuning)
        F = wf.frame_count(S)
        d_rows = DeviceBuffer(4 * F * N)
        digests = set()
        for _ in range(reps):
            wf.lib.ro_dev_memset(0, d_rows.ptr, 0, d_rows.nbytes)
            wf.process_resident(d_iq.ptr, S, d_rows.ptr, None, sync=True)
            rows = d_rows.download(np.float32, F * N)
            digests.add(hashlib.sha1(rows.tobytes()).hexdigest()[:16])
        print(N, ov, F, sorted(digests), flush=True)
        ok = ok and len(digests) == 1
        wf.close()
        d_rows.free()
    print("deterministic" if ok else "NON-DETERMINISTIC")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
