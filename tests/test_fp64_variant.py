"""cfg.precision = 1: the FP64 transform.  north_star's first tolerance -- relative 1e-5 in linear
power = 5e-6 in the stored amplitude -- must hold on EVERY bin (helpers.assert_rows_strict), against
the numpy oracle and against rows produced by the reference's own classes (golden fixtures)."""
import numpy as np
import pytest

from oracle import ref_numpy as R
from helpers import assert_rows_strict, load_golden

pytestmark = pytest.mark.gpu


def _iq(n, seed=5):
    from radio_observer_b200 import synth
    return synth.generate_int16(synth.SynthSpec(sample_rate=96000, n_samples=n, seed=seed))


@pytest.mark.parametrize("bins,overlap", [(1024, 0), (2048, 1536), (4096, 2048), (16384, 12288), (32768, 24576),
                                          (65536, 49152), (524288, 262144)])
def test_fp64_rows_meet_the_strict_tolerance_int16(bins, overlap):
    import radio_observer_b200 as ro
    iq = _iq(bins * 4 + 99 if bins <= 65536 else bins * 2)
    rows, _, n = ro.Waterfall(bins, overlap, precision=ro.PREC_FP64).process_host(iq)
    ref = R.waterfall_rows(iq, bins, overlap)
    assert n == ref.shape[0] >= 1
    assert_rows_strict(rows, ref.astype(np.float32), f"fp64 {bins}/{overlap}")
    # far inside the rule in fact: at most an ulp of float32 apart, and identical almost everywhere
    ulp = np.spacing(np.abs(ref.astype(np.float32)))
    assert (np.abs(rows - ref.astype(np.float32)) <= ulp).all()
    assert (rows == ref.astype(np.float32)).mean() > 0.99


def test_fp64_float32_input_and_gain():
    import radio_observer_b200 as ro
    from radio_observer_b200 import synth
    iq = synth.generate_float32(synth.SynthSpec(sample_rate=96000, n_samples=16384 * 3 + 5, seed=8))
    rows, _, n = ro.Waterfall(16384, 8192, in_format=ro.IN_FLOAT32, iq_gain=0.125, precision=1).process_host(iq)
    ref = R.waterfall_rows(iq, 16384, 8192, iq_gain=0.125).astype(np.float32)
    assert_rows_strict(rows, ref, "fp64 float32 + gain")


@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_fp64_against_rows_from_the_reference(name):
    import radio_observer_b200 as ro
    g = load_golden(name)
    m = g["meta"]
    rows, _, n = ro.Waterfall(m["bins"], m["overlap"], sample_rate=m["sample_rate"], precision=1).process_host(g["iq"])
    assert n == m["rows"]
    lo, hi = m["band"]
    assert_rows_strict(rows[g["rows_idx"]], g["rows_sub"], f"fp64 golden {name} full rows")
    assert_rows_strict(rows[:, lo:hi], g["bands"], f"fp64 golden {name} band")


def test_fp64_streaming_and_detector_exact():
    """Detector records on FP64 rows equal the checker's on the same rows, and the streaming path agrees."""
    import radio_observer_b200 as ro
    bins, overlap = 4096, 3072
    iq = _iq(bins * 20)
    det = dict(det_lo=2100, det_w=60, noise_lo=1800, noise_w=200, avg_bins=9)
    wf = ro.Waterfall(bins, overlap, precision=1, detector=det, max_batch_frames=5)
    rows, recs, n = wf.process_host(iq)
    d = R.DetectorBands(low_detect_bin=2100, detect_width=60, low_noise_bin=1800, noise_width=200, avg_bins=9, advance=1, jitter=1)
    nn, pp, aa = R.detector_rows(rows, d)
    assert np.array_equal(nn, recs["noise"]) and np.array_equal(pp, recs["peak"]) and np.array_equal(aa, recs["avg"])
    wf.push(iq[:2 * 30001])
    wf.push(iq[2 * 30001:])
    wf.flush()
    rows_s, recs_s, first = wf.poll_all()
    assert first == 0 and np.array_equal(rows_s, rows) and np.array_equal(recs_s["peak"], recs["peak"])
