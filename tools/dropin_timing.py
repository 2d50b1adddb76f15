#!/usr/bin/env python
"""Wall / user / system time of the reference's Pipeline on one WAV, three ways:
stock backend, CUDA adapter behind WAVStream, CUDA adapter behind NarrowWAVStream (-N).
Each variant is repeated; the spread tells a fluke from a bound.

usage: python tools/dropin_timing.py [SECONDS_OF_IQ] [REPEATS] [BATCH_FRAMES] [KEEP_DIR]
"""
import json
import os
import resource
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from radio_observer_b200 import synth  # noqa: E402


def main():
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 1800.0
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    keep = sys.argv[4] if len(sys.argv) > 4 else None
    n = int(secs * bench.SR)
    n -= n % 1024
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    tmp = tempfile.mkdtemp(prefix="ro_dropin_timing_", dir=base)
    out = []
    ref = None
    try:
        wav = os.path.join(tmp, "in.wav")
        synth.write_wav(wav, bench.host_station_iq(n, 0x5EED0002), bench.SR)
        variants = (("stock", bench.REF_BIN, "waterfall", None, ()),
                    ("cuda", bench.CUDA_BIN, "waterfall_cuda", {"cuda_batch_frames": batch}, ()),
                    ("cuda_narrow", bench.CUDA_BIN, "waterfall_cuda", {"cuda_batch_frames": batch}, ("-N",)))
        for rep in range(reps):
            for tag, binary, factory, extra, flags in variants:
                if tag == "stock" and rep > 0:
                    continue
                d = os.path.join(tmp, "%s_%d" % (tag, rep))
                os.makedirs(os.path.join(d, "out"))
                os.makedirs(os.path.join(d, "dump"))
                open(os.path.join(d, "cfg.json"), "w").write(bench.ref_config_json(os.path.join(d, "out"), factory, extra))
                for attempt in range(25):
                    r0 = resource.getrusage(resource.RUSAGE_CHILDREN)
                    t0 = time.perf_counter()
                    rc = subprocess.call([binary, "-c", os.path.join(d, "cfg.json"), "-d", os.path.join(d, "dump"), "-q", *flags, wav],
                                         stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=d)
                    dt = time.perf_counter() - t0
                    r1 = resource.getrusage(resource.RUSAGE_CHILDREN)
                    if rc == 0:
                        break
                timing = {}
                try:
                    timing = json.load(open(os.path.join(d, "dump", "timing.json")))
                except Exception:
                    pass
                names = sorted(os.listdir(os.path.join(d, "out")))
                try:
                    bolid = open(os.path.join(d, "dump", "bolid.txt")).read()
                    noise = open(os.path.join(d, "dump", "noise.bin"), "rb").read()
                    rowinfo = open(os.path.join(d, "dump", "rowinfo.bin"), "rb").read()
                except Exception:
                    bolid, noise, rowinfo = "", b"", b""
                if tag == "stock":
                    ref = {"names": set(names), "bolid": bolid, "noise": noise, "rowinfo": rowinfo}
                # events: peak frequency, band and raw-sample span must be identical; noise and magnitude come
                # from float32 rows here and from double-precision rows in the stock backend
                ev_same, ev_rel = True, 0.0
                ea, eb = ref["bolid"].strip().splitlines(), bolid.strip().splitlines()
                if len(ea) != len(eb):
                    ev_same = False
                for x, y in zip(ea, eb):
                    x, y = x.split(";"), y.split(";")
                    if x[1] != y[1] or x[3:] != y[3:]:
                        ev_same = False
                    for k in (0, 2):
                        ev_rel = max(ev_rel, abs(float(x[k]) - float(y[k])) / max(abs(float(x[k])), 1e-30))
                npa = np.dtype([("noise", "<f4"), ("peakfreq", "<f4"), ("mag", "<f4")])
                da, db = np.frombuffer(ref["noise"], dtype=npa), np.frombuffer(noise, dtype=npa)
                det = {}
                if len(da) == len(db) and len(da):
                    det = {"rows": int(len(da)), "same_peak_bin": float((da["peakfreq"] == db["peakfreq"]).mean()),
                           "noise_max_rel": float(np.max(np.abs(da["noise"].astype(np.float64) - db["noise"]) / np.abs(da["noise"])))}
                same = {"events": bolid.count("\n"), "event_freq_band_span_identical": ev_same,
                        "event_noise_mag_max_rel": ev_rel,
                        "rowinfo_equal_stock": rowinfo == ref["rowinfo"],
                        "detector_records": det,
                        "names_only_stock": sorted(ref["names"] - set(names))[:6],
                        "names_only_here": sorted(set(names) - ref["names"])[:6],
                        "n_only_stock": len(ref["names"] - set(names)), "n_only_here": len(set(names) - ref["names"])}
                if keep:
                    os.makedirs(keep, exist_ok=True)
                    open(os.path.join(keep, "%s_%d_names.txt" % (tag, rep)), "w").write("\n".join(names) + "\n")
                    open(os.path.join(keep, "%s_%d_bolid.txt" % (tag, rep)), "w").write(bolid)
                rec = {"variant": tag, "against_stock": same, "rep": rep, "attempts": attempt + 1, "wall_s": round(dt, 3),
                       "user_s": round(r1.ru_utime - r0.ru_utime, 3), "sys_s": round(r1.ru_stime - r0.ru_stime, 3),
                       "msamples_s": round(n / dt / 1e6, 2), "pipeline": timing,
                       "files": len(os.listdir(os.path.join(d, "out")))}
                print(json.dumps(rec), flush=True)
                out.append(rec)
                shutil.rmtree(d, ignore_errors=True)
        # The archive path on the same samples (radio_observer_b200/archive.py + recorders.py: whole-array
        # C-ABI call, trigger state machine and file planning in Python): which snapshot and bolid files it
        # would write, and its events, against what the stock Pipeline wrote.
        if ref is not None:
            from radio_observer_b200 import archive as A, recorders as P
            iq = synth.read_wav(wav)[0] if hasattr(synth, "read_wav") else bench.host_station_iq(n, 0x5EED0002)
            keys = dict(low_detect_freq=26450.0, hi_detect_freq=26550.0, low_noise_freq=26000.0, hi_noise_freq=26300.0,
                        avg_freq_range=40.0, advance_time=5.0, jitter_time=2.0)
            b = A.bolid_bands(bench.SR, bench.BINS, bench.OVERLAP, **keys)
            band = (A.frequency_to_bin(26200.0, bench.SR, bench.BINS), A.frequency_to_bin(26800.0, bench.SR, bench.BINS))
            t0 = time.perf_counter()
            rows, det, _ = A.reprocess_archive(iq, bench.BINS, bench.OVERLAP, sample_rate=bench.SR, bands=[band], detector=b)
            trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"],
                                  fft_sample_rate=A.fft_sample_rate(bench.SR, bench.BINS, bench.OVERLAP))
            trig.feed(det)
            F = len(rows)
            times = P.frame_times(int(F), bench.BINS, bench.OVERLAP, bench.SR, block=1024)
            common = dict(sample_rate=bench.SR, bins=bench.BINS, overlap=bench.OVERLAP, origin="bench")
            snaps = P.plan_snapshots(int(F), times, snapshot_length=60, **common)
            blids = P.plan_bolids(trig.events, times, snapshot_length=60, n_rows=int(F), **common)
            dt = time.perf_counter() - t0
            mine = set(p.file_name for p in snaps + blids)
            theirs = set(x for x in ref["names"] if x.endswith("_snap.fits") or x.endswith("_blid.fits"))
            ev_same = len(trig.events) == ref["bolid"].count("\n")
            for ev, line in zip(trig.events, ref["bolid"].strip().splitlines()):
                f = A.event_fields(ev, sample_rate=bench.SR, bins=bench.BINS, overlap=bench.OVERLAP,
                                   low_detect_freq=keys["low_detect_freq"], hi_detect_freq=keys["hi_detect_freq"])
                x = line.split(";")
                if ("%.9g" % f["peak_freq"], "%.9g" % f["min_freq"], "%.9g" % f["max_freq"], "%d" % f["end_sample"]) != \
                        (x[1], x[3], x[4], x[6]):
                    ev_same = False
            rec = {"variant": "archive_path", "wall_s": round(dt, 3), "msamples_s": round(n / dt / 1e6, 1), "rows": int(F),
                   "against_stock": {"events": len(trig.events), "event_freq_band_span_identical": ev_same,
                                     "snap_blid_files": len(mine), "n_only_stock": len(theirs - mine),
                                     "n_only_here": len(mine - theirs),
                                     "names_only_stock": sorted(theirs - mine)[:6], "names_only_here": sorted(mine - theirs)[:6]}}
            print(json.dumps(rec), flush=True)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
