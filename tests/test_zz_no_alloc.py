"""ro_wf_long_rows_static -- the allocation-free entry point (include/ro_cuda.h, csrc/ro_static.cu).

Collected last on purpose: everything before it has run by the time
this file is reached, and it passes alone in a fresh process as well.

The new entry point is called through the C ABI on inputs and outputs the test owns and compared with
the existing entry point, ro_wf_process_resident of a handle with the same bins / overlap / encoding /
gain, on two calls in a row that share the library's static scratch: the second call has other input,
fewer frames, and (in one case) another frame size, so anything left behind by the first would show.
Tolerance: the rule every row comparison of this suite uses (helpers.assert_rows_close), applied with
the existing entry point's rows as the reference; where both run the same two kernels (every case
but 32k, which the handle gives to the fused kernel) the rows are also required to be bit-identical.
"""
import numpy as np
import pytest

import radio_observer_b200 as ro
from radio_observer_b200 import capi, synth
from radio_observer_b200.capi import DeviceBuffer

from helpers import assert_rows_close

pytestmark = pytest.mark.gpu


def station(n, seed, fmt):
    spec = synth.SynthSpec(sample_rate=96000, n_samples=n, seed=seed, carrier_freq=0.228 * 96000,
                           meteors=[synth.Meteor(start=n // 3, amp=4000.0, tau=0.2, freq=0.2146 * 96000, rate=100.0)])
    iq = synth.generate_int16(spec)
    if fmt == ro.IN_FLOAT32:
        return (iq.astype(np.float32) / np.float32(32768.0)).astype(np.float32)
    return iq


def existing_entry_point(iq, n, bins, overlap, fmt, gain):
    """Rows of ro_wf_process_resident (handle-owned scratch and tables)."""
    wf = ro.Waterfall(bins, overlap, sample_rate=96000, in_format=fmt, iq_gain=gain)
    try:
        F = wf.frame_count(n)
        d_in = DeviceBuffer(iq.nbytes)
        d_rows = DeviceBuffer(F * bins * 4)
        d_in.upload(iq)
        assert wf.process_resident(d_in.ptr, n, d_rows.ptr, None) == F
        rows = d_rows.download(np.float32, F * bins).reshape(F, bins)
        name = wf.kernel_name()
        d_in.free()
        d_rows.free()
        return rows, name
    finally:
        wf.close()


def static_entry_point(iq, n, bins, overlap, fmt, gain, d_in, d_rows, F):
    """Rows of ro_wf_long_rows_static into buffers the caller brought along."""
    d_in.upload(iq)
    capi.load_library().ro_dev_memset(0, d_rows.ptr, 0xFF, d_rows.nbytes)      # NaN everywhere: a row not written shows
    got = capi.long_rows_static(0, bins, overlap, fmt, gain, d_in.ptr, n, d_rows.ptr, F)
    assert got == F
    return d_rows.download(np.float32, F * bins).reshape(F, bins)


@pytest.mark.parametrize("first,second", [
    # (bins, overlap, encoding, iq_gain, frames)
    ((65536, 49152, ro.IN_INT16, 0.0, 9), (65536, 49152, ro.IN_INT16, 0.0, 4)),
    ((131072, 65536, ro.IN_INT16, 9.0, 5), (65536, 49153, ro.IN_FLOAT32, -0.25, 7)),
    ((32768, 24576, ro.IN_INT16, 0.0, 12), (32768, 24576, ro.IN_INT16, 0.0, 3)),
])
def test_static_entry_point_matches_the_existing_one_twice_in_a_row(first, second):
    # the caller's buffers, sized for the larger of the two calls, allocated before either call
    need_in = need_rows = 0
    calls = []
    for k, (bins, overlap, fmt, gain, frames) in enumerate((first, second)):
        hop = bins - overlap
        n = bins + (frames - 1) * hop + min(7, hop - 1)
        iq = station(n, seed=1000 * k + bins + overlap + frames, fmt=fmt)
        calls.append((iq, n, bins, overlap, fmt, gain, frames))
        need_in = max(need_in, iq.nbytes)
        need_rows = max(need_rows, frames * bins * 4)
    d_in, d_rows = DeviceBuffer(need_in), DeviceBuffer(need_rows)
    try:
        got = [static_entry_point(iq, n, bins, overlap, fmt, gain, d_in, d_rows, frames).copy()
               for (iq, n, bins, overlap, fmt, gain, frames) in calls]           # two calls in a row ...
        for k, (iq, n, bins, overlap, fmt, gain, frames) in enumerate(calls):    # ... then the existing entry point
            ref, kernel = existing_entry_point(iq, n, bins, overlap, fmt, gain)
            assert ref.shape == got[k].shape == (frames, bins)
            assert np.isfinite(got[k]).all()
            assert_rows_close(got[k], ref.astype(np.float64), "static entry point, call %d, N=%d" % (k + 1, bins))
            if "MODE_SPLIT" in kernel:
                assert np.array_equal(got[k], ref), "same kernels, other scratch: rows must be bit-identical"
    finally:
        d_in.free()
        d_rows.free()


def test_static_entry_point_rejects_what_does_not_fit():
    bins, overlap = 65536, 49152
    cap = capi.STATIC_SCRATCH_BYTES // (8 * bins)
    frames = cap + 1
    n = bins + (frames - 1) * (bins - overlap)
    d_in, d_rows = DeviceBuffer(4 * n), DeviceBuffer(16)
    try:
        with pytest.raises(ro.RoError, match="static array"):
            capi.long_rows_static(0, bins, overlap, ro.IN_INT16, 0.0, d_in.ptr, n, d_rows.ptr, frames)
        # frame sizes outside 32768 .. RO_STATIC_MAX_BINS belong to the handle's entry points
        with pytest.raises(ro.RoError, match="long frames only"):
            capi.long_rows_static(0, 16384, 12288, ro.IN_INT16, 0.0, d_in.ptr, n, d_rows.ptr, 1)
        with pytest.raises(ro.RoError, match="long frames only"):
            capi.long_rows_static(0, 2 * capi.STATIC_MAX_BINS, 0, ro.IN_INT16, 0.0, d_in.ptr, n, d_rows.ptr, 1)
    finally:
        d_in.free()
        d_rows.free()
