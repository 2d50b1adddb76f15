"""radio_observer_b200.recorders: frame times, snapshot slicing and bolid files of the archive path
against the oracle's models (pinned to the reference by test_oracle_golden) and against the names,
lengths and header cards of the files the reference itself wrote for the golden runs."""
import os

import numpy as np
import pytest

from oracle import ref_numpy as R
from radio_observer_b200 import archive as A
from radio_observer_b200 import recorders as P
from radio_observer_b200 import fits as F
from radio_observer_b200.capi import DET_DTYPE
from helpers import load_golden, recorder_cfg

SKIP = ("DATE    =", "COMMENT Local time", "COMMENT Git version")


@pytest.mark.parametrize("bins,overlap,sr,block,n", [
    (16384, 12288, 96000, 1024, 768000), (2048, 1024, 48000, 1024, 200000), (32768, 24576, 48000, 1024, 400000),
    (4096, 2048, 192000, 4096, 300000), (1024, 1000, 44100, 1024, 30000), (4096, 0, 96000, 1000, 100000),
])
def test_frame_times_follow_the_reference_chain(bins, overlap, sr, block, n):
    want = R.frame_times(n, bins, overlap, sr, block=block)
    got = P.frame_times(len(want), bins, overlap, sr, block=block)
    assert [(t.sec, t.usec) for t in want] == got
    # and with a stream start time
    t0 = R.WFTime(1700000000, 999999)
    want = R.frame_times(n, bins, overlap, sr, block=block, t0=t0)
    assert [(t.sec, t.usec) for t in want] == P.frame_times(len(want), bins, overlap, sr, block=block, t0=(1700000000, 999999))


@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_planned_files_match_what_the_reference_wrote(name, tmp_path):
    g = load_golden(name)
    m = g["meta"]
    sr, bins, ov, n_rows = m["sample_rate"], m["bins"], m["overlap"], m["rows"]
    snap, bol = recorder_cfg(g, "snapshot"), recorder_cfg(g, "bolid")
    times = P.frame_times(n_rows, bins, ov, sr, block=1024)
    # --- snapshot recorder
    plan = P.plan_snapshots(n_rows, times, sample_rate=sr, bins=bins, overlap=ov,
                            snapshot_length=snap["snapshot_length"], origin="gold")
    want_names = sorted(k for k in g["fits"] if k.endswith("_snap.fits"))
    assert [p.file_name for p in plan] == want_names
    lo, hi = A.frequency_to_bin(snap["low_freq"], sr, bins), A.frequency_to_bin(snap["hi_freq"], sr, bins)
    b0, b1 = m["band"]
    band_rows = g["bands"][:, lo - b0:hi - b0]
    import hashlib
    for p in plan:
        want = g["fits"][p.file_name]
        assert want["shape"] == [hi - lo, p.length]
        path = str(tmp_path / p.file_name)
        P.write_planned(path, p, band_rows, left_bin=lo, sample_rate=sr, bins=bins, overlap=ov,
                        low_freq=snap["low_freq"], origin="gold")
        cards, hdr, data = F.read_image(path)
        assert [c for c in cards if not c.startswith(SKIP)] == [c for c in want["cards"] if not c.startswith(SKIP)]
        assert hashlib.sha256(np.ascontiguousarray(data).tobytes()).hexdigest() == want["sha256"]
    # --- bolid recorder, driven by the reference's own per-row records
    keys = {k: bol[k] for k in ("low_detect_freq", "hi_detect_freq", "low_noise_freq", "hi_noise_freq",
                                "advance_time", "jitter_time", "avg_freq_range")}
    b = A.bolid_bands(sr, bins, ov, **keys)
    npa = g["npa"]
    recs = np.zeros(n_rows, dtype=DET_DTYPE)
    recs["noise"], recs["avg"] = npa["noise"][:n_rows], npa["mag"][:n_rows]
    recs["peak"] = [R.frequency_to_bin(f, sr, bins) - b["det_lo"] for f in npa["peakfreq"][:n_rows]]
    trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"], fft_sample_rate=A.fft_sample_rate(sr, bins, ov))
    trig.feed(recs)
    bplan = P.plan_bolids(trig.events, times, sample_rate=sr, bins=bins, overlap=ov,
                          snapshot_length=bol["snapshot_length"], origin="gold")
    want_names = sorted(k for k in g["fits"] if k.endswith("_blid.fits"))
    assert sorted(p.file_name for p in bplan) == want_names
    blo, bhi = A.frequency_to_bin(bol["low_freq"], sr, bins), A.frequency_to_bin(bol["hi_freq"], sr, bins)
    bband = g["bands"][:, blo - b0:bhi - b0]
    n_px = 0
    for p in bplan:
        want = g["fits"][p.file_name]
        assert want["shape"] == [bhi - blo, p.length]
        if p.start_row < 0:
            continue                                   # rows before the archive: the reference wrote ring garbage there
        path = str(tmp_path / p.file_name)
        P.write_planned(path, p, bband, left_bin=blo, sample_rate=sr, bins=bins, overlap=ov, low_freq=bol["low_freq"], origin="gold")
        cards, hdr, data = F.read_image(path)
        assert [c for c in cards if not c.startswith(SKIP)] == [c for c in want["cards"] if not c.startswith(SKIP)]
        assert hashlib.sha256(np.ascontiguousarray(data).tobytes()).hexdigest() == want["sha256"]
        n_px += 1
    assert n_px >= 1
    # --- 8f N3: the raw-I/Q file that goes with every bolid file, cut from the archive's own samples
    rows_buf = 8 * max(P.snapshot_rows(snap["snapshot_length"], A.fft_sample_rate(sr, bins, ov)),
                       P.snapshot_rows(bol["snapshot_length"], A.fft_sample_rate(sr, bins, ov)))
    cap = P.raw_ring_capacity(rows_buf, sample_rate=sr, bins=bins, overlap=ov)
    n_raw = 0
    for p in bplan:
        name = P.raw_file_name(p, "gold")
        want = g["fits"][name]
        first, count = P.raw_extent(p, sample_rate=sr, bins=bins, overlap=ov, ring_capacity=cap)
        assert want["shape"] == [2, count], name
        path = str(tmp_path / name)
        P.write_raw(path, p, g["iq"], sample_rate=sr, bins=bins, overlap=ov, ring_capacity=cap, origin="gold")
        cards, hdr, data = F.read_image(path)
        assert [c for c in cards if not c.startswith(SKIP)] == [c for c in want["cards"] if not c.startswith(SKIP)], name
        if first + count <= len(g["iq"]) // 2:         # beyond the end of the stream the reference wrote what its ring held
            assert hashlib.sha256(np.ascontiguousarray(data).tobytes()).hexdigest() == want["sha256"], name
            n_raw += 1
    assert n_raw >= 1


def test_event_cut_short_by_the_end_of_the_stream_gets_no_file(tmp_path):
    """A bolid snapshot reaches `advance` rows past the last detection but is queued `jitter` rows after
    it; the reference's worker thread holds it back until the rows are there and drops it when the stream
    ends first (src/WaterfallBackend.cpp:60-104).  200 s of the bench's station recording end inside such
    an event: the stock Pipeline (oracle/_ref, the checker) reports 34 events and writes 33 bolid files.
    The oracle's model and the archive path's planner must both leave that file out."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    from radio_observer_b200 import synth
    if not os.path.exists(bench.REF_BIN):
        pytest.fail("oracle/_ref/ref_oracle missing: run __graft_entry__.build() in the build container")
    sr, bins, ov = bench.SR, bench.BINS, bench.OVERLAP
    n = 200 * sr
    wav = str(tmp_path / "in.wav")
    synth.write_wav(wav, bench.host_station_iq(n, 0x5EED0002), sr)
    out, dump = tmp_path / "out", tmp_path / "dump"
    cfg = tmp_path / "cfg.json"
    cfg.write_text(bench.ref_config_json(str(out), "waterfall", None))
    for attempt in range(50):              # the reference's start-up race, see test_dropin_adapter.run_pipeline
        for d in (out, dump):
            if d.exists():
                import shutil
                shutil.rmtree(d)
            d.mkdir()
        rc = subprocess.call([bench.REF_BIN, "-c", str(cfg), "-d", str(dump), "-q", wav], cwd=str(tmp_path),
                             stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        if rc == 0 and os.listdir(out):
            break
    else:
        raise AssertionError("reference run failed")
    wrote = sorted(f for f in os.listdir(out) if f.endswith("_blid.fits"))
    reported = (dump / "bolid.txt").read_text().count("\n")
    assert reported == len(wrote) + 1 >= 10
    npa = np.fromfile(str(dump / "noise.bin"), dtype=np.dtype([("noise", "<f4"), ("peakfreq", "<f4"), ("mag", "<f4")]))
    n_rows = len(npa)
    assert n_rows == A.frame_count(n, bins, ov)
    keys = dict(low_detect_freq=26450.0, hi_detect_freq=26550.0, low_noise_freq=26000.0, hi_noise_freq=26300.0,
                avg_freq_range=40.0, advance_time=5.0, jitter_time=2.0)
    b = A.bolid_bands(sr, bins, ov, **keys)
    peaks = np.array([R.frequency_to_bin(f, sr, bins) - b["det_lo"] for f in npa["peakfreq"]], dtype=np.int32)
    # the checker's model
    model = R.BolidModel(sample_rate=sr, bins=bins, overlap=ov, origin="bench", output_type="blid", snapshot_length=60,
                         low_freq=26200.0, hi_freq=26800.0, **keys)
    R.run_recorders(None, n_rows, R.frame_times(n, bins, ov, sr, block=1024), [model], bins,
                    npa=(npa["noise"], peaks, npa["mag"]))
    assert len(model.events) == reported
    assert sorted(d.file_name.split("/")[-1] for d in model.out if d.written) == wrote
    assert [d.written for d in model.out].count(False) == 1 and not model.out[-1].written
    # the archive path's planner
    recs = np.zeros(n_rows, dtype=DET_DTYPE)
    recs["noise"], recs["peak"], recs["avg"] = npa["noise"], peaks, npa["mag"]
    trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"], fft_sample_rate=A.fft_sample_rate(sr, bins, ov))
    trig.feed(recs)
    assert len(trig.events) == reported
    times = P.frame_times(n_rows, bins, ov, sr, block=1024)
    common = dict(sample_rate=sr, bins=bins, overlap=ov, snapshot_length=60, origin="bench")
    assert sorted(p.file_name for p in P.plan_bolids(trig.events, times, n_rows=n_rows, **common)) == wrote
    assert len(P.plan_bolids(trig.events, times, **common)) == reported
    # the snapshot recorder of the same run
    snaps = P.plan_snapshots(n_rows, times, sample_rate=sr, bins=bins, overlap=ov, snapshot_length=60, origin="bench")
    assert sorted(p.file_name for p in snaps) == sorted(f for f in os.listdir(out) if f.endswith("_snap.fits"))
