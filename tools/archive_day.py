#!/usr/bin/env python3
"""tools/archive_day.py -- BASELINE config 4: reprocess H hours of synthetic 96 kHz Bolidozor-station
I/Q, time-sharded over the ranks of one box with overlap halos, no data-path collective.

  python tools/archive_day.py --hours 2                                   # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
         --master-port 29533 tools/archive_day.py --hours 24              # eight GPUs

Rank 0 builds the archive in shared memory (a 10 min synthetic station recording -- carrier, chirped
meteor echoes, noise -- tiled to the requested length), every rank memory-maps it, processes its own
shard through ro_wf_process_host (host int16 in, recorder-band rows + detector records out) and hands
the results to rank 0 over a gloo group; rank 0 concatenates in frame order, runs the trigger state
machine over the merged records and re-computes the frames around every shard boundary on its own
GPU to check that the seams are invisible (bit-identical rows).  One JSON line on stdout.
Needs GPUs; nothing here touches oracle/.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from radio_observer_b200 import archive as A, synth      # noqa: E402
from radio_observer_b200.capi import DET_DTYPE as A_DET  # noqa: E402

SR, BINS, OVERLAP = 96000, 16384, 12288
KEYS = dict(low_detect_freq=26450.0, hi_detect_freq=26550.0, low_noise_freq=26000.0, hi_noise_freq=26300.0,
            avg_freq_range=40.0, advance_time=5.0, jitter_time=2.0)        # Bolidozor.json:74-93


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hours", type=float, default=1.0)
    ap.add_argument("--dir", default="/dev/shm" if os.path.isdir("/dev/shm") else "/tmp")
    ap.add_argument("--out", default=None, help="write the snapshot / bolid FITS files and the metadata CSV here")
    ap.add_argument("--max-files", type=int, default=200, help="bound on FITS files written per recorder")
    ap.add_argument("--keep-archive", action="store_true", help="reuse an archive file of the same length and leave it behind")
    ap.add_argument("--chunk-frames", type=int, default=16384, help="frames per pinned staging buffer")
    ap.add_argument("--copy-threads", type=int, default=0,
                    help="threads reading the archive into the staging buffer (0 = min(8, cores / ranks))")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if a.copy_threads <= 0:
        a.copy_threads = max(2, min(8, (os.cpu_count() or 8) // world))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    out_fd = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")           # host objects only: the data path has no collective
    path = os.path.join(a.dir, "ro_archive_%dh.npy" % int(a.hours * 100))
    n_total = int(a.hours * 3600 * SR)
    t_gen = 0.0
    if rank == 0 and not (a.keep_archive and os.path.exists(path)):
        t0 = time.time()
        base = synth.generate_int16(synth.SynthSpec(sample_rate=SR, n_samples=600 * SR, seed=4, carrier_freq=26700.3,
                                                    meteor_band=(26450.0, 26550.0)))
        mm = np.lib.format.open_memmap(path, mode="w+", dtype=np.int16, shape=(2 * n_total,))
        for s in range(0, 2 * n_total, base.size):
            e = min(2 * n_total, s + base.size)
            mm[s:e] = base[:e - s]
        mm.flush()
        del mm
        t_gen = time.time() - t0
    if dist is not None:
        dist.barrier()
    iq = np.load(path, mmap_mode="r")
    b = A.bolid_bands(SR, BINS, OVERLAP, **KEYS)
    band = (A.frequency_to_bin(26200.0, SR, BINS), A.frequency_to_bin(26800.0, SR, BINS))
    compute = A.default_compute(BINS, OVERLAP, SR, [band], b, local)      # seams, below
    staged = A.staged_compute(BINS, OVERLAP, SR, [band], b, local, chunk_frames=a.chunk_frames, copy_threads=a.copy_threads)
    warm_r = np.empty((8, band[1] - band[0]), dtype=np.float32)
    warm_d = np.zeros(8, dtype=A_DET)
    staged(iq, 0, BINS + 7 * (BINS - OVERLAP), warm_r, warm_d)      # warm-up: context, buffers
    if dist is not None:
        dist.barrier()
    t0 = time.time()
    rows, det, shards = A.reprocess_archive(iq, BINS, OVERLAP, sample_rate=SR, bands=[band], detector=b,
                                            rank=rank, world=world, dist=dist, shared_dir=a.dir, staged=staged)
    t_all = time.time() - t0
    if rank == 0:
        F = A.frame_count(n_total, BINS, OVERLAP)
        assert rows.shape == (F, band[1] - band[0]) and det.shape == (F,)
        trig = A.BolidTrigger(det_lo=b["det_lo"], advance=b["advance"], jitter=b["jitter"],
                              fft_sample_rate=A.fft_sample_rate(SR, BINS, OVERLAP))
        t1 = time.time()
        trig.feed(det)
        t_trig = time.time() - t1
        # seams: 16 frames around every shard boundary, recomputed here from the archive
        hop = BINS - OVERLAP
        seams_ok = True
        for s in shards[1:]:
            f0 = max(0, s.frame0 - 8)
            f1 = min(F, s.frame0 + 8)
            r2, d2 = compute(np.ascontiguousarray(iq[2 * f0 * hop:2 * ((f1 - 1) * hop + BINS)]))
            seams_ok = seams_ok and np.array_equal(r2, rows[f0:f1]) and np.array_equal(d2["peak"], det["peak"][f0:f1]) \
                and np.array_equal(d2["noise"], det["noise"][f0:f1])
        written = None
        if a.out:
            from radio_observer_b200 import recorders as P
            os.makedirs(a.out, exist_ok=True)
            t2 = time.time()
            times = P.frame_times(int(F), BINS, OVERLAP, SR, block=1024)
            common = dict(sample_rate=SR, bins=BINS, overlap=OVERLAP, origin="archive")
            snaps = P.plan_snapshots(int(F), times, snapshot_length=60, **common)
            blids = P.plan_bolids(trig.events, times, snapshot_length=60, n_rows=int(F), **common)
            with open(os.path.join(a.out, "archive_meta.csv"), "w") as csv:
                csv.write("# file name; noise; peak f.; mag.; duration\n")
                for p in blids:         # every reported event has its line, also one that starts before row 0
                    f = A.event_fields(p.event, sample_rate=SR, bins=BINS, overlap=OVERLAP,
                                       low_detect_freq=KEYS["low_detect_freq"], hi_detect_freq=KEYS["hi_detect_freq"])
                    csv.write(A.bolid_csv_line(p.file_name, f) + "\n")
            cap = P.raw_ring_capacity(8 * P.snapshot_rows(60, A.fft_sample_rate(SR, BINS, OVERLAP)), sample_rate=SR, bins=BINS, overlap=OVERLAP)
            n_raw = 0
            for p in snaps[:a.max_files]:
                P.write_planned(os.path.join(a.out, p.file_name), p, rows, left_bin=band[0], low_freq=26200.0, **common)
            for p in blids[:a.max_files]:
                if p.start_row >= 0:    # rows before the archive do not exist (the reference writes ring garbage there)
                    P.write_planned(os.path.join(a.out, p.file_name), p, rows, left_bin=band[0], low_freq=26200.0, **common)
                P.write_raw(os.path.join(a.out, P.raw_file_name(p, "archive")), p, iq, ring_capacity=cap, **common)
                n_raw += 1
            written = {"snapshots_planned": len(snaps), "bolids_planned": len(blids),
                       "fits_written": min(len(snaps), a.max_files) + min(len(blids), a.max_files), "raw_iq_files": n_raw,
                       "seconds": round(time.time() - t2, 2)}
        line = {"tool": "archive_day", "hours": a.hours, "n_gpus": world, "samples": n_total, "frames": int(F),
                "bins": BINS, "overlap": OVERLAP, "seconds_compute_and_merge": round(t_all, 3),
                "Msamples_per_s_end_to_end": round(n_total / t_all / 1e6, 1),
                "realtime_factor": round(a.hours * 3600 / t_all, 1),
                "events": len(trig.events), "seconds_trigger_fsm": round(t_trig, 3),
                "seams_bit_identical": bool(seams_ok), "shards": [[s.frame0, s.frame1] for s in shards],
                "seconds_generate_archive": round(t_gen, 1), "outputs": written,
                "chunk_frames": a.chunk_frames, "copy_threads": a.copy_threads,
                "note": "host int16 archive in shared memory -> per rank: copy threads -> two pinned staging buffers -> "
                        "ro_wf_process_host -> rows of the recorder band + detector records written in place into arrays "
                        "all ranks map (no gather); gloo for two barriers; no data-path collective"}
        os.write(out_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        for f in [path] if not a.keep_archive else []:
            try:
                os.remove(f)
            except OSError:
                pass
        for f in os.listdir(a.dir):            # the merged arrays of reprocess_archive
            if f.startswith("ro_rows_") or f.startswith("ro_det_"):
                try:
                    os.remove(os.path.join(a.dir, f))
                except OSError:
                    pass
    return 0


if __name__ == "__main__":
    sys.exit(main())
