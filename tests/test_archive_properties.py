"""Randomised cross-checks of the archive path's host logic against the oracle's models
(hypothesis): shard plans, the frame-time chain, the trigger state machine and the file planning."""
import numpy as np
from hypothesis import given, settings, strategies as st

from oracle import ref_numpy as R
from radio_observer_b200 import archive as A
from radio_observer_b200 import recorders as P
from radio_observer_b200.capi import DET_DTYPE


@settings(max_examples=150, deadline=None)
@given(log2n=st.integers(10, 16), ov_frac=st.floats(0, 0.999), n_mult=st.floats(0.5, 40), world=st.integers(1, 9))
def test_shard_plan_properties(log2n, ov_frac, n_mult, world):
    bins = 1 << log2n
    overlap = int(bins * ov_frac)
    n = int(bins * n_mult)
    shards = A.plan_shards(n, bins, overlap, world)
    F = R.frame_count(n, bins, overlap)
    hop = bins - R.clamp_overlap(bins, overlap)
    assert sum(s.n_frames for s in shards) == F
    pos = 0
    for s in shards:
        assert s.frame0 == pos
        pos = s.frame1
        if s.n_frames:
            assert s.sample0 == s.frame0 * hop and s.sample1 == (s.frame1 - 1) * hop + bins <= n


@settings(max_examples=60, deadline=None)
@given(log2n=st.integers(10, 15), ov_frac=st.floats(0, 0.99), sr=st.sampled_from([44100, 48000, 96000, 192000]),
       block=st.sampled_from([1000, 1024, 4096, 333]), n_frames=st.integers(1, 60),
       t0=st.tuples(st.integers(0, 2 * 10**9), st.integers(0, 999999)))
def test_frame_times_equal_the_oracle_chain(log2n, ov_frac, sr, block, n_frames, t0):
    bins = 1 << log2n
    overlap = int(bins * ov_frac)
    hop = bins - overlap
    n = (n_frames - 1) * hop + bins + 17
    want = R.frame_times(n, bins, overlap, sr, block=block, t0=R.WFTime(*t0))
    got = P.frame_times(len(want), bins, overlap, sr, block=block, t0=t0)
    assert [(t.sec, t.usec) for t in want] == got


@settings(max_examples=80, deadline=None)
@given(seed=st.integers(0, 10**6), n_rows=st.integers(1, 400), advance=st.floats(0.05, 3), jitter=st.floats(0.05, 3),
       p_hit=st.floats(0.02, 0.6))
def test_trigger_and_bolid_planning_equal_the_oracle_model(seed, n_rows, advance, jitter, p_hit):
    sr, bins, ov = 96000, 4096, 2048
    rng = np.random.default_rng(seed)
    keys = dict(low_detect_freq=26450.0, hi_detect_freq=26950.0, low_noise_freq=25000.0, hi_noise_freq=26300.0,
                advance_time=advance, jitter_time=jitter, avg_freq_range=200.0)
    b = A.bolid_bands(sr, bins, ov, **keys)
    noise = rng.uniform(50, 100, n_rows).astype(np.float32)
    # runs of detections and gaps
    hit = np.zeros(n_rows, bool)
    i = 0
    while i < n_rows:
        run = int(rng.geometric(0.3))
        if rng.random() < p_hit:
            hit[i:i + run] = True
        i += run
    avg = np.where(hit, noise * 2 * rng.uniform(1.01, 4, n_rows), noise * 2 * rng.uniform(0.1, 1.0, n_rows)).astype(np.float32)
    peak = rng.integers(0, b["det_w"], n_rows).astype(np.int32)
    model = R.BolidModel(sample_rate=sr, bins=bins, overlap=ov, origin="x", output_type="blid", snapshot_length=3,
                         low_freq=26000.0, hi_freq=27000.0, **keys)
    times = R.frame_times((n_rows - 1) * (bins - ov) + bins, bins, ov, sr, block=1024)
    R.run_recorders(None, n_rows, times, [model], bins, npa=(noise, peak, avg))
    recs = np.zeros(n_rows, dtype=DET_DTYPE)
    recs["noise"], recs["peak"], recs["avg"] = noise, peak, avg
    trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"], fft_sample_rate=A.fft_sample_rate(sr, bins, ov))
    cut = int(rng.integers(0, n_rows + 1))
    trig.feed(recs[:cut])
    trig.feed(recs[cut:])
    assert len(trig.events) == len(model.events)
    ptimes = P.frame_times(n_rows, bins, ov, sr, block=1024)
    plan = P.plan_bolids(trig.events, ptimes, sample_rate=sr, bins=bins, overlap=ov, snapshot_length=3, origin="x")
    for e, o, d, pl in zip(trig.events, model.events, model.out, plan):
        assert (e.emit_row, e.length, e.noise, e.magnitude, e.duration_s) == (o.row, o.length, o.noise, o.magnitude, o.duration)
        assert A.bin_to_frequency(e.peak_bin, sr, bins) == o.peak_freq
        assert pl.length == d.length
        # the model's ring has not wrapped here, so its ring index is the global row
        if model.ring.pushed <= model.ring.capacity:
            assert pl.file_name == d.file_name.split("/")[-1] and pl.time == (d.time.sec, d.time.usec)
    # files the reference's worker thread gets to write before the stream ends (see run_recorders)
    if model.ring.pushed <= model.ring.capacity:
        kept = P.plan_bolids(trig.events, ptimes, sample_rate=sr, bins=bins, overlap=ov, snapshot_length=3, origin="x",
                             n_rows=n_rows)
        assert [(pl.file_name, pl.length) for pl in kept] == \
            [(d.file_name.split("/")[-1], d.length) for d in model.out if d.written]


@settings(max_examples=80, deadline=None)
@given(n_rows=st.integers(1, 500), snapshot_length=st.integers(1, 6), log2n=st.sampled_from([11, 12, 13]),
       ov_frac=st.sampled_from([0.0, 0.5, 0.75]))
def test_snapshot_planning_equals_the_oracle_model(n_rows, snapshot_length, log2n, ov_frac):
    sr = 96000
    bins = 1 << log2n
    ov = int(bins * ov_frac)
    model = R.SnapshotModel(sample_rate=sr, bins=bins, overlap=ov, origin="x", output_type="snap",
                            snapshot_length=snapshot_length, low_freq=26000.0, hi_freq=27000.0)
    times = R.frame_times((n_rows - 1) * (bins - ov) + bins, bins, ov, sr, block=1024)
    ring = R.run_recorders(None, n_rows, times, [model], bins)
    if ring.pushed > ring.capacity - 2:
        return                        # the model's ring wrapped: outside the archive path's stated assumption
    ptimes = P.frame_times(n_rows, bins, ov, sr, block=1024)
    plan = P.plan_snapshots(n_rows, ptimes, sample_rate=sr, bins=bins, overlap=ov, snapshot_length=snapshot_length, origin="x")
    assert [(p.start_row, p.length, p.file_name, p.time) for p in plan] == \
        [(d.global_start if d.start else 0, d.length, d.file_name.split("/")[-1], (d.time.sec, d.time.usec)) for d in model.out]
