"""radio_observer_b200.fits against the FITS files the reference's own
SnapshotRecorder / FITSWriter wrote for the golden cases (over the oracle's cfitsio shim):
every header card and every pixel byte."""
import hashlib
import os

import numpy as np
import pytest

from oracle import ref_numpy as R
from radio_observer_b200 import archive as A
from radio_observer_b200 import fits as F
from helpers import load_golden, recorder_cfg

SKIP = ("DATE    =", "COMMENT Local time", "COMMENT Git version")


@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_snapshot_files_byte_for_byte(name, tmp_path):
    g = load_golden(name)
    m = g["meta"]
    sr, bins, ov = m["sample_rate"], m["bins"], m["overlap"]
    snap = recorder_cfg(g, "snapshot")
    # which rows go into which file, and the file's time, is the recorder's sequential logic: from the checker
    model = R.SnapshotModel(sample_rate=sr, bins=bins, overlap=ov, origin="gold", output_type="snap",
                            snapshot_length=snap["snapshot_length"], low_freq=snap["low_freq"], hi_freq=snap["hi_freq"])
    times = R.frame_times(m["n_samples"], bins, ov, sr, block=1024)
    R.run_recorders(None, m["rows"], times, [model], bins)
    lo, hi = model.left_bin, model.right_bin
    assert lo == A.frequency_to_bin(snap["low_freq"], sr, bins) and hi == A.frequency_to_bin(snap["hi_freq"], sr, bins)
    b0, b1 = m["band"]                      # bands.npy holds the reference's rows over [b0, b1) for every row
    assert b0 <= lo and hi <= b1
    n_files = 0
    for d in model.out:
        fn = os.path.basename(d.file_name)
        want = g["fits"][fn]
        assert fn == F.file_basename(d.time.sec, d.time.usec, "gold", "snap")
        if d.global_start < 0:
            continue                        # the first snapshot's rows predate row 0 in the ring: nothing to rebuild from
        pixels = g["bands"][d.global_start:d.global_start + d.length, lo - b0:hi - b0]
        cards = F.snapshot_cards(hi - lo, d.length, origin="gold", time_sec=d.time.sec, time_usec=d.time.usec,
                                 fft_sample_rate=A.fft_sample_rate(sr, bins, ov),
                                 left_frequency=np.float32(snap["low_freq"]),
                                 bin_width=A.bin_to_frequency(1, sr, bins) - A.bin_to_frequency(0, sr, bins))
        path = str(tmp_path / fn)
        F.write_image(path, cards, pixels)
        assert os.path.getsize(path) % 2880 == 0
        got_cards, hdr, data = F.read_image(path)
        assert [c for c in got_cards if not c.startswith(SKIP)] == [c for c in want["cards"] if not c.startswith(SKIP)]
        assert list(data.shape[::-1]) == want["shape"]
        assert hashlib.sha256(np.ascontiguousarray(data).tobytes()).hexdigest() == want["sha256"]
        assert np.array_equal(data.astype(np.float32), pixels)
        n_files += 1
    assert n_files >= 1


def test_value_formatting():
    assert F._fmt_real(26200.0, 7) == "26200." and F._fmt_real(5.859375, 7) == "5.859375"
    assert F._fmt_real(1.0, 7) == "1." and F._fmt_real(42.666666666666664, 15) == "42.6666666666667"
    assert F._fmt_real(1e20, 7) == "1.0E+20" and F._fmt_real(-46875.0, 7) == "-46875."
    assert F._card_string("ORIGIN", "gold").rstrip() == "ORIGIN  = 'gold    '"
    assert F._card_string("X", "it's").startswith("X       = 'it''s   '")
    assert len(F._card_value("CRVAL2", "0", "x" * 100)) == 80
