"""Drop-in test: the reference's own Pipeline / WAVStream / recorders / FITS writer driven
by CudaWaterfallBackend (factory "waterfall_cuda") against the same binary's stock
"waterfall" backend, on the committed golden inputs.

Both programs are built from the reference's sources in place (oracle/Makefile,
radio_observer_b200/host/Makefile) and travel to the GPU box prebuilt.  The stock one is
the checker; the adapter one is the product path (it fails if libro_cuda.so is missing).

Compared: number of rows, per-row (ring mark, raw mark, stream time) handles -- exactly;
full rows -- tests/helpers.py tolerance; per-row (noise, peak frequency, magnitude) -- exact
or a documented tie; bolid events; names, header cards and shapes of every FITS file written
-- exactly (CRVAL1/CDELT1/CRVAL2/CDELT2 included); FITS pixels -- row tolerance.
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

from radio_observer_b200 import synth
from helpers import FLOOR_PEAK, REL_TOL, assert_rows_close, load_golden
from make_golden import read_fits

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")
CUDA_BIN = os.path.join(ROOT, "radio_observer_b200", "host", "_build", "radio-observer-cuda")

ROWINFO = np.dtype([("mark", "<i4"), ("rawmark", "<i4"), ("sec", "<i8"), ("usec", "<i8")])
NPA = np.dtype([("noise", "<f4"), ("peakfreq", "<f4"), ("mag", "<f4")])


def run_pipeline(binary, g, factory, tmp, extra=None, flags=(), tag=None, iq=None, dump_rows=True):
    d = os.path.join(tmp, tag or factory)
    os.makedirs(os.path.join(d, "OUT"))
    os.makedirs(os.path.join(d, "dump"))
    cfg = json.loads(json.dumps(g["config"]))
    be = cfg["configurations"][0]["children"][0]
    be["factory"] = factory
    be.update(extra or {})
    txt = json.dumps(cfg, indent=1).replace('"OUT"', '"%s"' % os.path.join(d, "OUT"))
    open(os.path.join(d, "cfg.json"), "w").write(txt)
    wav = os.path.join(tmp, "in.wav")
    if not os.path.exists(wav):
        synth.write_wav(wav, g["iq"] if iq is None else iq, g["sample_rate"])
    # the reference starts recorder threads from a base-class constructor and now and then
    # aborts with "pure virtual method called" before reading any sample: retry
    for attempt in range(25):
        for sub in ("OUT", "dump"):
            shutil.rmtree(os.path.join(d, sub))
            os.makedirs(os.path.join(d, sub))
        r = subprocess.run([binary, "-c", os.path.join(d, "cfg.json"), "-d", os.path.join(d, "dump"),
                            *(("-r",) if dump_rows else ()), "-q", *flags, wav],
                           cwd=d, capture_output=True, text=True)
        if r.returncode == 0:
            break
    else:
        raise AssertionError("%s failed rc=%d: %s" % (binary, r.returncode, r.stderr[-2000:]))
    bins = g["meta"]["bins"]
    out = dict(
        rows=(np.fromfile(os.path.join(d, "dump", "rows.f32"), dtype=np.float32).reshape(-1, bins)
              if dump_rows else None),
        info=np.fromfile(os.path.join(d, "dump", "rowinfo.bin"), dtype=ROWINFO),
        npa=np.fromfile(os.path.join(d, "dump", "noise.bin"), dtype=NPA),
        bolid=open(os.path.join(d, "dump", "bolid.txt")).read(),
        fits={}, csv={}, stderr=r.stderr)
    for fn in sorted(os.listdir(os.path.join(d, "OUT"))):
        p = os.path.join(d, "OUT", fn)
        if fn.endswith(".fits"):
            cards, hdr, shape, data = read_fits(p)
            # wall-clock stamps and the build label are not outputs of the path
            skip = ("DATE    =", "COMMENT Local time", "COMMENT Git version")
            out["fits"][fn] = dict(cards=[c for c in cards if not c.startswith(skip)], shape=shape,
                                   data=np.array(data, dtype=np.float32))
        elif fn.endswith(".csv"):
            out["csv"][fn] = open(p).read()
    return out


def _retry_reference_races(fn, *args, attempts=4):
    """Both binaries are the REFERENCE's Pipeline, and the reference formats file names and DATE-OBS
    with the non-reentrant gmtime() from several threads at once (src/WFTime.cpp:25-30): now and then a
    name or a date card of either run carries another thread's time.  That race is not on the path
    under test; a comparison that trips over it is repeated from scratch."""
    for i in range(attempts):
        try:
            return fn(*args)
        except AssertionError:
            if i == attempts - 1:
                raise


@pytest.mark.gpu
@pytest.mark.parametrize("name,batch", [("n2048", 1), ("c2b_16k", 1), ("c2b_16k", 7)])
def test_cuda_backend_is_a_drop_in(name, batch):
    _retry_reference_races(_dropin_once, name, batch)


@pytest.mark.gpu
@pytest.mark.parametrize("name,batch", [("n2048", 1), ("c2b_16k", 1), ("c2b_16k", 64)])
def test_narrow_ingest_is_a_drop_in(name, batch):
    """NarrowWAVStream -> CudaWaterfallBackend::processRaw (int16 all the way to the GPU)
    against the stock WAVStream -> WaterfallBackend, same checks as the Complex path."""
    _retry_reference_races(_dropin_once, name, batch, True)


def _dropin_once(name, batch, narrow=False):
    if not (os.path.exists(REF_BIN) and os.path.exists(CUDA_BIN)):
        pytest.fail("prebuilt binaries missing: run __graft_entry__.build() in the build container")
    g = load_golden(name)
    tmp = tempfile.mkdtemp(prefix="ro_dropin_")
    try:
        ref = run_pipeline(REF_BIN, g, "waterfall", tmp)
        # File names are formatted with the non-reentrant gmtime() from the producer thread and
        # from every recorder's writer thread at once (reference src/WFTime.cpp:25-30,
        # src/WFTime.h:131,150).  With the DSP on the GPU the producer no longer spends its time
        # in FFTs, the threads collide more often, and now and then a name carries another
        # thread's date (seen: 19700101000000 + 394 ms instead of 19700101000004 + 394 ms).
        # That race is the reference's; rerun when only the names differ.
        for attempt in range(6):
            got = run_pipeline(CUDA_BIN, g, "waterfall_cuda", tmp, {"cuda_batch_frames": batch},
                               flags=("-N",) if narrow else ())
            if sorted(got["fits"]) == sorted(ref["fits"]):
                break
            shutil.rmtree(os.path.join(tmp, "waterfall_cuda"))
        # the stock run must reproduce the committed golden (same reference code)
        assert len(ref["rows"]) == g["meta"]["rows"]
        assert np.array_equal(ref["info"], g["rowinfo"])
        # 1. framing, row numbering, handles: exact
        assert len(got["rows"]) == len(ref["rows"])
        assert np.array_equal(got["info"]["mark"], ref["info"]["mark"])
        assert np.array_equal(got["info"]["sec"], ref["info"]["sec"])
        assert np.array_equal(got["info"]["usec"], ref["info"]["usec"])
        if batch == 1:      # raw-ring marks are taken at frame time; identical when rows are not delayed
            assert np.array_equal(got["info"]["rawmark"], ref["info"]["rawmark"])
        # 2. rows
        assert_rows_close(got["rows"], ref["rows"], "%s adapter rows" % name)
        # 3. detector statistics computed by the UNMODIFIED BolidRecorder on the GPU rows
        assert len(got["npa"]) == len(ref["npa"])
        peak = np.abs(ref["rows"]).max(axis=1)
        same_peak = got["npa"]["peakfreq"] == ref["npa"]["peakfreq"]
        tol = np.maximum(REL_TOL * np.abs(ref["npa"]["mag"]), 2 * FLOOR_PEAK * peak)
        assert (np.abs(got["npa"]["mag"].astype(np.float64) - ref["npa"]["mag"])[same_peak] <= tol[same_peak]).all()
        tol_n = np.maximum(REL_TOL * np.abs(ref["npa"]["noise"]), 4 * FLOOR_PEAK * peak)
        assert (np.abs(got["npa"]["noise"].astype(np.float64) - ref["npa"]["noise"]) <= tol_n).all()
        print("[dropin] %s: rows %d, identical peak bin in %d rows" % (name, len(ref["rows"]), int(same_peak.sum())))
        assert same_peak.mean() > 0.97          # argmax ties between near-equal noise bins only
        # 4. bolid events: same count; start/end samples and frequencies identical
        eb = [l.split(";") for l in ref["bolid"].strip().splitlines()]
        gb = [l.split(";") for l in got["bolid"].strip().splitlines()]
        assert len(eb) == len(gb) and len(eb) >= 1
        for e, q in zip(eb, gb):
            assert e[3:5] == q[3:5]             # minFreq, maxFreq
            if batch == 1:
                assert e[5:] == q[5:]           # startSample, endSample (raw-ring positions)
            assert abs(float(e[1]) - float(q[1])) <= 2 * float(ref_df(g))     # peak frequency within one bin
        # 5. FITS files: same names, same header cards (CRVAL/CDELT/CRPIX/CTYPE/NAXIS...), pixels in tolerance
        assert sorted(got["fits"]) == sorted(ref["fits"]), (sorted(got["fits"]), sorted(ref["fits"]), got["bolid"], ref["bolid"])
        n_checked = 0
        for fn in ref["fits"]:
            a, b = ref["fits"][fn], got["fits"][fn]
            assert a["shape"] == b["shape"], fn
            assert a["cards"] == b["cards"], (fn, [x for x in zip(a["cards"], b["cards"]) if x[0] != x[1]])
            if "raws" in fn:
                assert np.array_equal(a["data"], b["data"]), fn       # raw I/Q: untouched by the GPU
            else:
                w = a["shape"][0]
                ra, rb = a["data"].reshape(-1, w), b["data"].reshape(-1, w)
                # floor from the full row's peak is not available here; the band's own peak is smaller -> stricter
                err = np.abs(ra.astype(np.float64) - rb)
                fl = FLOOR_PEAK * np.abs(ref["rows"]).max()
                assert (err <= np.maximum(REL_TOL * np.abs(ra), fl)).all(), fn
            n_checked += 1
        assert n_checked >= 2
        assert sorted(got["csv"]) == sorted(ref["csv"]) or len(got["csv"]) == len(ref["csv"])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.gpu
@pytest.mark.parametrize("narrow", [False, True])
def test_long_run_that_laps_the_row_ring(narrow):
    _retry_reference_races(_lapping_once, narrow)


def _lapping_once(narrow):
    """The recorders' worker threads read the row ring, the handle table and the raw ring after the
    fact (reference src/WaterfallBackend.cpp:60-108); nothing in the reference stops the producer from
    overwriting what a queued snapshot still needs.  300 s of station I/Q lap the golden configuration's
    ring (752 rows = 32 s) nine times, and behind the CUDA backend the producer outruns the FITS
    writers: without the adapter's back-pressure (CudaWaterfallBackend::waitForWriters) files come out
    with rows, times and names of one ring capacity later.  Same files as the stock backend expected."""
    if not (os.path.exists(REF_BIN) and os.path.exists(CUDA_BIN)):
        pytest.fail("prebuilt binaries missing: run __graft_entry__.build() in the build container")
    g = load_golden("c2b_16k")
    sr = g["sample_rate"]
    spec = synth.SynthSpec(sample_rate=sr, n_samples=300 * sr, seed=0x1A9, carrier_freq=26700.3,
                           meteor_band=(26450.0, 26550.0))
    iq = synth.generate_int16(spec)
    tmp = tempfile.mkdtemp(prefix="ro_dropin_lap_")
    try:
        # no -r row dump for 7000 rows of 64 KB: rows are compared through the FITS files
        ref = run_pipeline(REF_BIN, g, "waterfall", tmp, iq=iq, dump_rows=False)
        for attempt in range(6):            # the reference's gmtime race, see _dropin_once
            got = run_pipeline(CUDA_BIN, g, "waterfall_cuda", tmp, {"cuda_batch_frames": 64},
                               flags=("-N",) if narrow else (), iq=iq, dump_rows=False)
            if sorted(got["fits"]) == sorted(ref["fits"]):
                break
            shutil.rmtree(os.path.join(tmp, "waterfall_cuda"))
        assert len(ref["info"]) > 8 * 752
        for k in ("mark", "sec", "usec"):
            assert np.array_equal(got["info"][k], ref["info"][k])
        eb = [l.split(";") for l in ref["bolid"].strip().splitlines()]
        gb = [l.split(";") for l in got["bolid"].strip().splitlines()]
        assert len(eb) == len(gb) and len(eb) >= 10
        for e, q in zip(eb, gb):
            assert e[3:5] == q[3:5]
        only_ref = sorted(set(ref["fits"]) - set(got["fits"]))
        only_got = sorted(set(got["fits"]) - set(ref["fits"]))
        assert not only_ref and not only_got, (len(ref["fits"]), only_ref[:5], only_got[:5])
        peak = max(float(np.abs(v["data"]).max()) for fn, v in ref["fits"].items() if "raws" not in fn)
        for fn in ref["fits"]:
            a, b = ref["fits"][fn], got["fits"][fn]
            assert a["shape"] == b["shape"], fn
            assert a["cards"] == b["cards"], (fn, [x for x in zip(a["cards"], b["cards"]) if x[0] != x[1]])
            if "raws" in fn:
                assert np.array_equal(a["data"], b["data"]), fn
            else:
                # the full row's peak is not in the file: the largest pixel of any file of this run (the carrier
                # is inside the recorded band) stands in for it
                err = np.abs(a["data"].astype(np.float64) - b["data"])
                assert (err <= np.maximum(REL_TOL * np.abs(a["data"]), FLOOR_PEAK * peak)).all(), fn
        print("[dropin lap] %d rows, %d events, %d FITS files identical in name, header and raw I/Q"
              % (len(ref["info"]), len(eb), len(ref["fits"])))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_cuda_backend_fp64_reproduces_the_stock_backend(name):
    _retry_reference_races(_dropin_fp64_once, name)


def _dropin_fp64_once(name):
    """cuda_precision = 1: the reference's Pipeline with the CUDA backend in double precision gives the
    stock backend's rows to the last float ulp, hence the same detector statistics, events and files."""
    if not (os.path.exists(REF_BIN) and os.path.exists(CUDA_BIN)):
        pytest.fail("prebuilt binaries missing: run __graft_entry__.build() in the build container")
    g = load_golden(name)
    tmp = tempfile.mkdtemp(prefix="ro_dropin64_")
    try:
        ref = run_pipeline(REF_BIN, g, "waterfall", tmp)
        for attempt in range(6):            # the reference's gmtime race, see above
            got = run_pipeline(CUDA_BIN, g, "waterfall_cuda", tmp, {"cuda_batch_frames": 1, "cuda_precision": 1})
            if sorted(got["fits"]) == sorted(ref["fits"]):
                break
            shutil.rmtree(os.path.join(tmp, "waterfall_cuda"))
        assert got["rows"].shape == ref["rows"].shape
        ulp = np.spacing(np.abs(ref["rows"]))
        assert (np.abs(got["rows"] - ref["rows"]) <= ulp).all()
        same = float((got["rows"] == ref["rows"]).mean())
        print("[dropin fp64] %s: %.4f %% of pixels bit-identical, the rest one ulp apart" % (name, 100 * same))
        assert same > 0.99
        for k in ("mark", "sec", "usec", "rawmark"):
            assert np.array_equal(got["info"][k], ref["info"][k])
        # detector statistics of the unmodified BolidRecorder: identical unless a one-ulp pixel is involved
        assert np.array_equal(got["npa"]["peakfreq"], ref["npa"]["peakfreq"])
        assert np.allclose(got["npa"]["noise"], ref["npa"]["noise"], rtol=2e-7, atol=0)
        assert np.allclose(got["npa"]["mag"], ref["npa"]["mag"], rtol=2e-7, atol=0)
        assert got["bolid"].strip().splitlines().__len__() == ref["bolid"].strip().splitlines().__len__() >= 1
        for e, q in zip(ref["bolid"].strip().splitlines(), got["bolid"].strip().splitlines()):
            e, q = e.split(";"), q.split(";")
            assert e[1] == q[1] and e[3:] == q[3:]          # peak frequency, band, start/end samples
        assert sorted(got["fits"]) == sorted(ref["fits"])
        for fn in ref["fits"]:
            a, b = ref["fits"][fn], got["fits"][fn]
            assert a["shape"] == b["shape"] and a["cards"] == b["cards"], fn
            assert (np.abs(a["data"].astype(np.float64) - b["data"]) <= np.spacing(np.abs(a["data"]))).all(), fn
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.parametrize("tail", [0, 1, 300, 512])
def test_narrow_wav_frontend_reads_what_wavstream_reads(tail):
    """The frontend alone, on the CPU: NarrowWAVStream in front of the reference's own stock backend
    (which gets the widened block) against WAVStream in front of the same backend.  Everything is
    bit-identical, including the tail block of a data chunk that is not a multiple of 4096 bytes,
    for which WAVStream passes on twice the samples it read (src/WAVStream.cpp:126-139)."""
    if not (os.path.exists(REF_BIN) and os.path.exists(CUDA_BIN)):
        pytest.fail("prebuilt binaries missing: run __graft_entry__.build() in the build container")
    g = load_golden("n2048")
    iq = g["iq"]
    if tail:
        iq = iq[: 2 * ((len(iq) // 2 // 1024 - 3) * 1024 + tail)]
    tmp = tempfile.mkdtemp(prefix="ro_narrow_")
    try:
        def once():
            for sub in ("a", "b"):
                shutil.rmtree(os.path.join(tmp, sub), ignore_errors=True)
            a = run_pipeline(REF_BIN, g, "waterfall", tmp, tag="a", iq=iq)
            b = run_pipeline(CUDA_BIN, g, "waterfall", tmp, flags=("-N",), tag="b", iq=iq)
            assert len(a["rows"]) > 0 and np.array_equal(a["rows"], b["rows"])
            assert np.array_equal(a["info"], b["info"])
            assert np.array_equal(a["npa"], b["npa"])
            assert a["bolid"] == b["bolid"]
            assert sorted(a["fits"]) == sorted(b["fits"])
            for fn in a["fits"]:
                assert a["fits"][fn]["cards"] == b["fits"][fn]["cards"], fn
                assert np.array_equal(a["fits"][fn]["data"], b["fits"][fn]["data"]), fn
            if tail:
                n_in = len(iq) // 2 + tail        # the doubled tail block
                hop = g["meta"]["bins"] - g["meta"]["overlap"]
                assert len(a["rows"]) == (n_in - g["meta"]["bins"]) // hop + 1
        _retry_reference_races(once)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def ref_df(g):
    return g["meta"]["sample_rate"] / g["meta"]["bins"]
