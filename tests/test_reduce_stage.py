"""The row_avg / bin_decim extension stage (wf_reduce_kernel).  The reference has no such stage
(SURVEY.md D2): the checker is the definition restated in oracle/ref_numpy.reduce_rows, and the
default 1/1 must leave the reference-matching rows untouched, bit for bit."""
import numpy as np
import pytest

from oracle import ref_numpy as R
from helpers import assert_rows_close

pytestmark = pytest.mark.gpu


def _iq(n_samples, seed=3):
    from radio_observer_b200 import synth
    return synth.generate_int16(synth.SynthSpec(sample_rate=96000, n_samples=n_samples, seed=seed))


def test_default_is_identity_bit_for_bit():
    import radio_observer_b200 as ro
    iq = _iq(4096 * 9 + 100)
    a, _, na = ro.Waterfall(4096, 2048).process_host(iq)
    b, _, nb = ro.Waterfall(4096, 2048, row_avg=1, bin_decim=1).process_host(iq)
    c, _, nc = ro.Waterfall(4096, 2048, row_avg=0, bin_decim=0).process_host(iq)
    assert na == nb == nc and np.array_equal(a, b) and np.array_equal(a, c)


@pytest.mark.parametrize("bins,overlap,ravg,decim", [
    (4096, 2048, 4, 1), (4096, 2048, 1, 8), (4096, 0, 3, 2), (16384, 12288, 2, 128), (16384, 12288, 5, 16),
    (1024, 512, 7, 4), (32768, 24576, 2, 2),
])
def test_reduced_rows_match_the_definition(bins, overlap, ravg, decim):
    import radio_observer_b200 as ro
    iq = _iq(bins * 8 + 321)
    full, _, nf = ro.Waterfall(bins, overlap).process_host(iq)
    wf = ro.Waterfall(bins, overlap, row_avg=ravg, bin_decim=decim)
    assert wf.out_bins == bins // decim == wf.row_width
    red, _, n = wf.process_host(iq)
    assert n == nf // ravg == wf.row_count(iq.size // 2) and red.shape == (n, bins // decim)
    # the stage itself, against the definition applied to the GPU's own full rows: FP32 sums of R*D squares
    want = R.reduce_rows(full, ravg, decim)
    assert np.allclose(red, want, rtol=2e-6, atol=0.0), float(np.abs(red / want - 1).max())
    # end to end against the oracle's rows: the row tolerance carries over (an RMS of values within
    # tolerance is within tolerance), with the floor taken from the strongest line of the FULL rows
    ref = R.reduce_rows(R.waterfall_rows(iq, bins, overlap).astype(np.float32), ravg, decim)
    peak = full[:n * ravg].reshape(n, ravg * bins).max(axis=1)
    assert_rows_close(red, ref, f"reduce {bins}/{overlap} R={ravg} D={decim}", peak=peak)


def test_resident_streaming_and_host_paths_agree():
    import radio_observer_b200 as ro
    from radio_observer_b200.capi import DeviceBuffer
    bins, overlap, ravg, decim = 4096, 3072, 3, 4
    iq = _iq(bins * 30 + 17)
    S = iq.size // 2
    bands = [(100, 300), (700, 1024)]
    det = dict(det_lo=400, det_w=40, noise_lo=200, noise_w=100, avg_bins=5)
    host = ro.Waterfall(bins, overlap, row_avg=ravg, bin_decim=decim, bands=bands, detector=det)
    rows_h, det_h, n = host.process_host(iq)
    assert n == host.frame_count(S) // ravg and rows_h.shape == (n, 200 + 324)
    # resident: full reduced rows on the device
    wf = ro.Waterfall(bins, overlap, row_avg=ravg, bin_decim=decim, detector=det)
    d_iq, d_rows, d_det = DeviceBuffer(iq.nbytes), DeviceBuffer(4 * n * wf.out_bins), DeviceBuffer(16 * n)
    d_iq.upload(iq)
    assert wf.process_resident(d_iq.ptr, S, d_rows.ptr, d_det.ptr) == n
    rows_r = d_rows.download(np.float32, n * wf.out_bins).reshape(n, wf.out_bins)
    assert np.array_equal(np.concatenate([rows_r[:, 100:300], rows_r[:, 700:1024]], axis=1), rows_h)
    det_r = d_det.download(np.dtype(det_h.dtype), n)
    for k in ("noise", "peak", "avg"):
        assert np.array_equal(det_r[k], det_h[k])
    # the detector ran on the REDUCED rows: exact against the checker on those rows
    d = R.DetectorBands(low_detect_bin=400, detect_width=40, low_noise_bin=200, noise_width=100, avg_bins=5, advance=1, jitter=1)
    nn, pp, aa = R.detector_rows(rows_r, d)
    assert np.array_equal(nn, det_r["noise"]) and np.array_equal(pp, det_r["peak"]) and np.array_equal(aa, det_r["avg"])
    # streaming: ragged pushes, batches of 7 -> 9 frames (whole groups of 3); the incomplete last group is dropped
    st = ro.Waterfall(bins, overlap, row_avg=ravg, bin_decim=decim, bands=bands, detector=det, max_batch_frames=7)
    rng = np.random.default_rng(1)
    pos = 0
    while pos < S:
        k = int(rng.integers(1, 9000))
        st.push(iq[2 * pos:2 * min(S, pos + k)])
        pos += k
    st.flush()
    rows_s, det_s, first = st.poll_all()
    assert first == 0 and rows_s.shape == rows_h.shape
    assert np.array_equal(rows_s, rows_h)
    for k in ("noise", "peak", "avg"):
        assert np.array_equal(det_s[k], det_h[k])


def test_reduce_stage_alone_and_bad_configs():
    import radio_observer_b200 as ro
    from radio_observer_b200.capi import DeviceBuffer
    rng = np.random.default_rng(9)
    rows = np.abs(rng.standard_normal((10, 2048))).astype(np.float32) * 1000
    wf = ro.Waterfall(2048, 0, row_avg=3, bin_decim=32)
    d_in, d_out = DeviceBuffer(rows.nbytes), DeviceBuffer(4 * 3 * 64)
    d_in.upload(rows)
    assert wf.reduce_resident(d_in.ptr, 10, d_out.ptr) == 3
    got = d_out.download(np.float32, 3 * 64).reshape(3, 64)
    assert np.allclose(got, R.reduce_rows(rows, 3, 32), rtol=2e-6, atol=0)
    for kw in (dict(bin_decim=3), dict(bin_decim=256), dict(row_avg=-1), dict(bin_decim=8, bands=[(0, 2048)]),
               dict(bin_decim=2, detector=dict(det_lo=1000, det_w=100, noise_lo=0, noise_w=10, avg_bins=3))):
        with pytest.raises(ro.RoError):
            ro.Waterfall(2048, 0, **kw)
