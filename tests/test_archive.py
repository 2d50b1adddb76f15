"""Time-sharded archive reprocessing: shard plan, host-side trigger state machine,
two-rank merge over gloo (CPU), and shard-vs-whole bit-exactness on the GPU."""
import os
import sys

import numpy as np
import pytest

from oracle import ref_numpy as R
from radio_observer_b200 import archive as A
from helpers import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_samples,bins,overlap,world", [
    (768000, 16384, 12288, 1), (768000, 16384, 12288, 2), (768000, 16384, 12288, 8),
    (100000, 4096, 2048, 3), (65536, 65536, 49152, 4), (5000, 4096, 0, 2), (1 << 20, 1024, 1023, 8),
])
def test_plan_shards_partitions_the_frames(n_samples, bins, overlap, world):
    shards = A.plan_shards(n_samples, bins, overlap, world)
    F = R.frame_count(n_samples, bins, overlap)
    hop = bins - R.clamp_overlap(bins, overlap)
    assert len(shards) == world
    assert shards[0].frame0 == 0 and shards[-1].frame1 == F
    for a, b in zip(shards, shards[1:]):
        assert a.frame1 == b.frame0                      # contiguous, no frame twice, none lost
    for s in shards:
        if s.n_frames:
            assert s.sample0 == s.frame0 * hop and s.sample1 == (s.frame1 - 1) * hop + bins
            assert s.sample1 <= n_samples
            assert R.frame_count(s.n_samples, bins, overlap) == s.n_frames
        else:
            assert s.n_samples == 0
    sizes = [s.n_frames for s in shards]
    assert max(sizes) - min(sizes) <= 1                  # balanced
    # halo: the next non-empty shard re-reads exactly `overlap` samples of this one
    ne = [s for s in shards if s.n_frames]
    for a, b in zip(ne, ne[1:]):
        assert a.sample1 - b.sample0 == R.clamp_overlap(bins, overlap)


@pytest.mark.parametrize("sr,bins,overlap,keys", [
    (48000, 32768, 24576, dict(low_detect_freq=10300, hi_detect_freq=10900, low_noise_freq=9000, hi_noise_freq=9600,
                                 avg_freq_range=40, advance_time=2, jitter_time=5)),          # radio-observer.json
    (96000, 65536, 49152, dict(low_detect_freq=26450, hi_detect_freq=26550, low_noise_freq=26000, hi_noise_freq=26300,
                                 avg_freq_range=40, advance_time=5, jitter_time=2)),          # Bolidozor.json
    (96000, 16384, 12288, dict(low_detect_freq=26450, hi_detect_freq=26550, low_noise_freq=26000, hi_noise_freq=26300,
                                 avg_freq_range=40, advance_time=0.5, jitter_time=0.5)),
    (192000, 4096, 2048, dict(low_detect_freq=26550, hi_detect_freq=26450, low_noise_freq=26300, hi_noise_freq=26000,
                                avg_freq_range=100, advance_time=1, jitter_time=1)),          # swapped edges
])
def test_bands_match_the_oracle(sr, bins, overlap, keys):
    d = R.detector_bands(sr, bins, overlap, **keys)
    b = A.bolid_bands(sr, bins, overlap, **keys)
    assert (b["det_lo"], b["det_w"], b["noise_lo"], b["noise_w"], b["avg_bins"], b["advance"], b["jitter"]) == \
        (d.low_detect_bin, d.detect_width, d.low_noise_bin, d.noise_width, d.avg_bins, d.advance, d.jitter)
    for f in (0.0, 26500.0, -47999.0, 1e9, -1e9, 10100.5):
        assert A.frequency_to_bin(f, sr, bins) == R.frequency_to_bin(f, sr, bins)
    for k in (0, 1, bins // 2, bins - 1):
        assert A.bin_to_frequency(k, sr, bins) == R.bin_to_frequency(k, sr, bins)
    assert A.fft_sample_rate(sr, bins, overlap) == R.fft_sample_rate(sr, bins, overlap)


def _oracle_bolid(gold):
    cfg = gold["config"]["configurations"][0]["children"][0]
    rec = [c for c in cfg["children"] if c["factory"] == "bolid"][0]
    m = gold["meta"]
    keys = {k: rec[k] for k in ("low_detect_freq", "hi_detect_freq", "low_noise_freq", "hi_noise_freq",
                                "advance_time", "jitter_time", "avg_freq_range")}
    return rec, keys, m


@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_trigger_state_machine_matches_the_oracle(name):
    """BolidTrigger on the reference's own per-row (noise, peak, avg) records reproduces the events of
    the oracle's BolidModel (itself pinned to the reference's bolid.txt / CSV by test_oracle_golden)."""
    gold = load_golden(name)
    rec, keys, m = _oracle_bolid(gold)
    sr, bins, ov = m["sample_rate"], m["bins"], m["overlap"]
    npa = gold["npa"]
    model = R.BolidModel(sample_rate=sr, bins=bins, overlap=ov, origin="gold", output_type="blid",
                         snapshot_length=rec["snapshot_length"], low_freq=rec["low_freq"], hi_freq=rec["hi_freq"], **keys)
    times = R.frame_times(m["n_samples"], bins, ov, sr, block=1024)
    n_rows = m["rows"]
    b = A.bolid_bands(sr, bins, ov, **keys)
    # the reference logs the peak as a frequency: back to the offset inside the detect band
    peaks = np.array([R.frequency_to_bin(f, sr, bins) - b["det_lo"] for f in npa["peakfreq"]], dtype=np.int32)
    R.run_recorders(None, n_rows, times, [model], bins, npa=(npa["noise"], peaks, npa["mag"]))
    assert len(model.events) >= 1
    from radio_observer_b200.capi import DET_DTYPE
    recs = np.zeros(n_rows, dtype=DET_DTYPE)
    recs["noise"], recs["peak"], recs["avg"] = npa["noise"][:n_rows], peaks[:n_rows], npa["mag"][:n_rows]
    npa = recs
    trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"],
                          fft_sample_rate=A.fft_sample_rate(sr, bins, ov))
    # feed in two pieces: state carries across calls
    trig.feed(npa[:n_rows // 3])
    trig.feed(npa[n_rows // 3:n_rows])
    assert len(trig.events) == len(model.events)
    for e, o, desc in zip(trig.events, model.events, model.out):
        assert e.emit_row == o.row and e.length == o.length
        assert e.noise == o.noise and e.magnitude == o.magnitude and e.duration_s == o.duration
        assert A.bin_to_frequency(e.peak_bin, sr, bins) == o.peak_freq
        if e.start_row >= 0:
            assert e.start_row == desc.global_start


def _rank_main(rank, world, port, path, bins, overlap, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    iq = np.load(path, mmap_mode="r")

    def compute(x):                       # the checker stands in for the CUDA callable on a GPU-less host
        return R.waterfall_rows(np.asarray(x), bins, overlap).astype(np.float32), None

    rows, det, shards = A.reprocess_archive(iq, bins, overlap, rank=rank, world=world, dist=dist, compute=compute)
    if rank == 0:
        q.put((rows, [(s.frame0, s.frame1) for s in shards]))
    dist.barrier()
    dist.destroy_process_group()


def _rank_main_shared(rank, world, port, path, bins, overlap, shared, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    iq = np.load(path, mmap_mode="r")

    def staged(iq_map, s0, s1, rows_out, det_out):      # the checker in place of staged_compute() on a GPU-less host
        rows = R.waterfall_rows(np.asarray(iq_map[2 * s0:2 * s1]), bins, overlap).astype(np.float32)
        rows_out[:] = rows[:, 100:300]
        return len(rows)
    staged.row_width = 200

    rows, det, shards = A.reprocess_archive(iq, bins, overlap, rank=rank, world=world, dist=dist, shared_dir=shared, staged=staged)
    if rank == 0:
        q.put((np.array(rows), [(s.frame0, s.frame1) for s in shards]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_write_their_shards_in_place_into_shared_arrays(tmp_path, world):
    """The archive driver's merge: rank 0 creates the merged arrays as files every rank maps, each rank
    writes its own frame range, two barriers, nothing gathered."""
    import torch.multiprocessing as mp
    bins, overlap = 1024, 512
    rng = np.random.default_rng(12)
    iq = rng.integers(-3000, 3000, size=2 * 30000, dtype=np.int16)
    path = str(tmp_path / "iq.npy")
    np.save(path, iq)
    whole = R.waterfall_rows(iq, bins, overlap).astype(np.float32)[:, 100:300]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_rank_main_shared, args=(r, world, port, path, bins, overlap, str(tmp_path), q)) for r in range(world)]
    for p in procs:
        p.start()
    rows, frames = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert frames[0][0] == 0 and frames[-1][1] == whole.shape[0]
    assert np.array_equal(rows, whole)


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_merge_to_the_unsharded_result_over_gloo(tmp_path, world):
    import torch.multiprocessing as mp
    bins, overlap = 1024, 768
    rng = np.random.default_rng(11)
    iq = rng.integers(-3000, 3000, size=2 * 20000, dtype=np.int16)
    path = str(tmp_path / "iq.npy")
    np.save(path, iq)
    whole = R.waterfall_rows(iq, bins, overlap).astype(np.float32)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, path, bins, overlap, q)) for r in range(world)]
    for p in procs:
        p.start()
    rows, frames = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert frames[0][0] == 0 and frames[-1][1] == whole.shape[0]
    assert rows.shape == whole.shape
    assert np.array_equal(rows, whole)    # frames are independent: sharding changes nothing, bit for bit


@pytest.mark.gpu
@pytest.mark.parametrize("bins,overlap,world", [(16384, 12288, 2), (4096, 2048, 3), (32768, 24576, 2)])
def test_shards_equal_the_whole_on_the_gpu(bins, overlap, world):
    """Each shard through the CUDA library (one after the other on this GPU) == one unsharded run, bit for bit,
    rows and detector records."""
    import radio_observer_b200 as ro
    from radio_observer_b200 import synth
    spec = synth.SynthSpec(sample_rate=96000, n_samples=bins * 12 + 777, seed=21)
    iq = synth.generate_int16(spec)
    b = A.bolid_bands(96000, bins, overlap, low_detect_freq=26450, hi_detect_freq=26550, low_noise_freq=26000,
                      hi_noise_freq=26300, avg_freq_range=40, advance_time=0.5, jitter_time=0.5)
    band = (A.frequency_to_bin(26200, 96000, bins), A.frequency_to_bin(26800, 96000, bins))
    whole_rows, whole_det, shards1 = A.reprocess_archive(iq, bins, overlap, bands=[band], detector=b, device=0)
    assert whole_rows.shape[0] == A.frame_count(spec.n_samples, bins, overlap) == shards1[0].n_frames
    compute = A.default_compute(bins, overlap, 96000, [band], b, 0)
    parts = []
    for s in A.plan_shards(spec.n_samples, bins, overlap, world):
        r, d = compute(np.ascontiguousarray(iq[2 * s.sample0:2 * s.sample1]))
        assert r.shape[0] == s.n_frames
        parts.append((r, d))
    rows = np.concatenate([p[0] for p in parts])
    det = np.concatenate([p[1] for p in parts])
    assert np.array_equal(rows, whole_rows)
    for k in ("noise", "peak", "avg"):
        assert np.array_equal(det[k], whole_det[k])


@pytest.mark.parametrize("name", ["n2048", "c2b_16k"])
def test_event_text_formats_match_the_reference_output(name):
    """8f N4: the CSV metadata lines and the stdout `met;` lines the reference printed for the golden
    run, rebuilt from the trigger's events (file names and the wall-clock stamp taken from the golden)."""
    gold = load_golden(name)
    rec, keys, m = _oracle_bolid(gold)
    sr, bins, ov = m["sample_rate"], m["bins"], m["overlap"]
    b = A.bolid_bands(sr, bins, ov, **keys)
    npa = gold["npa"]
    from radio_observer_b200.capi import DET_DTYPE
    n_rows = m["rows"]
    recs = np.zeros(n_rows, dtype=DET_DTYPE)
    recs["noise"], recs["avg"] = npa["noise"][:n_rows], npa["mag"][:n_rows]
    recs["peak"] = [R.frequency_to_bin(f, sr, bins) - b["det_lo"] for f in npa["peakfreq"][:n_rows]]
    trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"], fft_sample_rate=A.fft_sample_rate(sr, bins, ov))
    trig.feed(recs)
    csv = [v["text"] for k, v in gold["fits"].items() if k.endswith(".csv") and "_blid.fits" in v["text"]][0]
    csv_lines = [l for l in csv.strip().split("\n") if "_blid.fits" in l]
    met_lines = [l for l in m["stdout"].strip().split("\n") if l.startswith("met;")]
    dump = [l.split(";") for l in gold["bolid"].strip().split("\n") if l]
    assert len(trig.events) == len(csv_lines) == len(met_lines) == len(dump) >= 1
    for ev, cl, ml, d in zip(trig.events, csv_lines, met_lines, dump):
        f = A.event_fields(ev, sample_rate=sr, bins=bins, overlap=ov, low_detect_freq=keys["low_detect_freq"],
                           hi_detect_freq=keys["hi_detect_freq"])
        assert A.bolid_csv_line(cl.split(";")[0], f) == cl
        stamp = ml.split(";")[1]                                   # "[<s>s, <us>us]": WFTime::now(), not an output of the path
        ts, tu = int(stamp[1:].split("s,")[0]), int(stamp.split(",")[1].strip()[:-3])
        assert A.met_line(ts, tu, f) == ml
        # the BolidMessage fields at full precision
        assert [np.float32(x) for x in d[:5]] == [f["noise"], f["peak_freq"], f["magnitude"], f["min_freq"], f["max_freq"]]
        assert int(d[6]) == f["end_sample"]


@pytest.mark.gpu
@pytest.mark.parametrize("chunk_frames,threads", [(64, 1), (100, 3)])
def test_staged_reader_equals_the_one_shot_call(tmp_path, chunk_frames, threads):
    """staged_compute (chunks through two pinned staging buffers, copy threads, in-place output) against
    one ro_wf_process_host call on the whole shard: identical rows and records, chunk seams included."""
    from radio_observer_b200 import synth
    from radio_observer_b200.capi import DET_DTYPE
    bins, overlap, sr = 16384, 12288, 96000
    hop = bins - overlap
    n = bins + 700 * hop + 11
    iq = synth.generate_int16(synth.SynthSpec(sample_rate=sr, n_samples=n, seed=77, carrier_freq=26700.3,
                                              meteor_band=(26450.0, 26550.0)))
    path = str(tmp_path / "iq.npy")
    np.save(path, iq)
    iq_map = np.load(path, mmap_mode="r")
    keys = dict(low_detect_freq=26450.0, hi_detect_freq=26550.0, low_noise_freq=26000.0, hi_noise_freq=26300.0,
                avg_freq_range=40.0, advance_time=5.0, jitter_time=2.0)
    b = A.bolid_bands(sr, bins, overlap, **keys)
    band = (A.frequency_to_bin(26200.0, sr, bins), A.frequency_to_bin(26800.0, sr, bins))
    one = A.default_compute(bins, overlap, sr, [band], b, 0)
    s0, s1 = 5 * hop, n - 3
    want_rows, want_det = one(np.ascontiguousarray(iq_map[2 * s0:2 * s1]))
    staged = A.staged_compute(bins, overlap, sr, [band], b, 0, chunk_frames=chunk_frames, copy_threads=threads)
    F = A.frame_count(s1 - s0, bins, overlap)
    rows = np.zeros((F, band[1] - band[0]), dtype=np.float32)
    det = np.zeros(F, dtype=DET_DTYPE)
    assert staged(iq_map, s0, s1, rows, det) == F == len(want_rows)
    assert np.array_equal(rows, want_rows)
    for k in ("noise", "peak", "avg"):
        assert np.array_equal(det[k], want_det[k])
    staged.close()
