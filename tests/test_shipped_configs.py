"""The three station configurations the reference ships, end to end.

tests/golden/shipped/<name>.json are the shipped files themselves (comments, trailing commas, unread
keys and all) with only the output directories replaced by a placeholder and snapshot_length cut to
20 s (tests/golden/make_shipped.py, which also ran them through the UNMODIFIED reference and committed
what it wrote).  Here:

  CPU   the stock reference (oracle/_ref/ref_oracle, prebuilt) on the regenerated input reproduces the
        committed fixtures: the oracle and the input recipe are pinned;
  GPU   the same text with "factory": "waterfall" -> "waterfall_cuda" as the ONLY edit runs through
        radio-observer-cuda: rows, handles, detector records, events, FITS cards (CRVAL1 / CDELT1 /
        CRVAL2 / CDELT2 ...) and pixels, CSV lines against the stock run -- with the WAVStream frontend
        and with the narrow-ingest one.
"""
import hashlib
import json
import os
import shutil
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import make_shipped as M                                    # noqa: E402
import test_dropin_adapter as D                             # noqa: E402
from helpers import FLOOR_PEAK, REL_TOL, assert_raws_equal                     # noqa: E402

SHIPPED = os.path.join(ROOT, "tests", "golden", "shipped")
NAMES = ["radio-observer", "Bolidozor", "Ionozor"]


def load(name):
    d = os.path.join(SHIPPED, name)
    meta = json.load(open(os.path.join(d, "meta.json")))
    c = M.CASES[name]
    iq = M.station_iq(c)
    assert hashlib.sha256(iq.tobytes()).hexdigest() == meta["input_sha256"], "the input recipe no longer gives the committed input"
    return dict(text=open(os.path.join(SHIPPED, name + ".json")).read(), meta=meta, case=c, iq=iq,
                rowinfo=np.load(os.path.join(d, "rowinfo.npy")), npa=np.load(os.path.join(d, "npa.npy")),
                bolid=open(os.path.join(d, "bolid.txt")).read(), fits=json.load(open(os.path.join(d, "fits.json"))),
                band_rows=np.load(os.path.join(d, "band_rows.npy")))


def tmpdir():
    return tempfile.mkdtemp(prefix="ro_shipped_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)


def check_stock_against_fixtures(g, ref):
    meta = g["meta"]
    assert len(ref["info"]) == meta["rows"]
    assert np.array_equal(ref["info"], g["rowinfo"])
    assert np.array_equal(ref["npa"], g["npa"])
    assert ref["bolid"] == g["bolid"]
    lo, hi = meta["band"]
    assert np.array_equal(ref["rows"][::8, lo:hi], g["band_rows"])
    assert sorted(ref["fits"]) == sorted(g["fits"])
    for fn, f in g["fits"].items():
        assert ref["fits"][fn]["cards"] == f["cards"], fn
        assert ref["fits"][fn]["shape"] == f["shape"], fn
        px = ref["fits"][fn]["data"]
        if "sha256_head" not in f:
            assert hashlib.sha256(px.tobytes()).hexdigest() == f["sha256"], fn
            continue
        # raw I/Q: the head is a fixed function of the input; the tail is read ahead of the producer
        # (helpers.raws_head_rows) and holds the input's samples or, where the writer thread got there
        # first, the ring's zeros
        px = px.reshape(-1, 2)
        head = f["head_rows"]
        assert hashlib.sha256(px[:head].tobytes()).hexdigest() == f["sha256_head"], fn
        if hashlib.sha256(px.tobytes()).hexdigest() != f["sha256"]:
            src = g["iq"].reshape(-1, 2)[f["first_sample"]:f["first_sample"] + len(px)].astype(np.float32)
            late = assert_raws_equal(px, src, head, fn)
            print("[shipped] %s: %d of %d samples past the head not yet pushed when the stock writer read them" % (fn, late, len(px) - head))


@pytest.mark.parametrize("name", NAMES)
def test_stock_reference_reproduces_the_committed_fixtures(name):
    if not os.path.exists(M.REF):
        pytest.fail("oracle/_ref/ref_oracle missing: run __graft_entry__.build() in the build container")
    g = load(name)
    # the shipped syntax really is the extended one: comments and trailing commas survive in the fixture
    assert "//" in g["text"] and ",\n\t\t}" in g["text"].replace(" ", "") or ",\n" in g["text"]
    assert '"snapshot_length": 20' in g["text"].replace("  ", " ") or "20" in g["text"]
    tmp = tmpdir()
    try:
        def once():
            shutil.rmtree(os.path.join(tmp, "OUT"), ignore_errors=True)
            ref = M.run_reference(M.REF, g["text"], g["iq"], g["case"]["sr"], tmp)
            check_stock_against_fixtures(g, ref)
            n_snap = len([f for f in ref["fits"] if f.endswith("_snap.fits")])
            assert n_snap >= 3
        D._retry_reference_races(once)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name,narrow", [(n, w) for n in NAMES for w in (False, True)])
def test_shipped_config_through_the_cuda_backend(name, narrow):
    if not (os.path.exists(M.REF) and os.path.exists(D.CUDA_BIN)):
        pytest.fail("prebuilt binaries missing: run __graft_entry__.build() in the build container")
    g = load(name)
    cuda_text = g["text"].replace('"factory": "waterfall"', '"factory": "waterfall_cuda"')
    assert cuda_text != g["text"] and cuda_text.count("waterfall_cuda") == g["text"].count('"factory": "waterfall"')
    gg = dict(meta=dict(name="%s%s" % (name, " narrow" if narrow else ""), sample_rate=g["case"]["sr"], bins=g["case"]["bins"],
                        overlap=g["case"]["overlap"]), config=None)
    # the bolid recorder's keys, for the tie classification of compare_runs (radio-observer, Bolidozor)
    rec = {"radio-observer": dict(low_detect_freq=10300, hi_detect_freq=10900, low_noise_freq=9000, hi_noise_freq=9600,
                                   advance_time=2, jitter_time=5, avg_freq_range=40),
           "Bolidozor": dict(low_detect_freq=26450, hi_detect_freq=26550, low_noise_freq=26000, hi_noise_freq=26300,
                             advance_time=5, jitter_time=2, avg_freq_range=40)}.get(name)
    gg["config"] = {"configurations": [{"children": [{"children": ([dict(factory="bolid", **rec)] if rec else [])}]}]}
    for attempt in range(2):
        ta, tb = tmpdir(), tmpdir()
        try:
            ref = M.run_reference(M.REF, g["text"], g["iq"], g["case"]["sr"], ta)
            check_stock_against_fixtures(g, ref)
            got = M.run_reference(D.CUDA_BIN, cuda_text, g["iq"], g["case"]["sr"], tb, flags=("-N",) if narrow else (),
                                  dump_rows=False)
            assert got["timing"]["backend"] == "waterfall_cuda" and got["timing"]["full_rows"] is False
            racy, n_files, n_events = D.compare_runs(ref, got, gg, rawmark_exact=True, rows_available=False, gpu_detector=bool(rec))
            if not racy:
                break
            D.RETRIES["name_race"] += 1
            print("[shipped] gmtime race in names / DATE-OBS, repeating once: %s" % (racy[:3],))
        finally:
            shutil.rmtree(ta, ignore_errors=True)
            shutil.rmtree(tb, ignore_errors=True)
    else:
        raise AssertionError("names / DATE-OBS differ in two consecutive runs: %r" % (racy[:6],))
    assert n_files == g["meta"]["fits"] and n_events == g["meta"]["events"]
    print("[shipped] %s: %d rows, %d files, %d events identical to the stock backend's (band %s of %d bins to the host)"
          % (gg["meta"]["name"], len(ref["info"]), n_files, n_events, got["timing"]["band"], g["case"]["bins"]))
