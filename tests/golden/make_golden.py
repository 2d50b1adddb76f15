#!/usr/bin/env python3
"""tests/golden/make_golden.py -- TEST INFRASTRUCTURE.

Generates the committed golden fixtures by running the UNMODIFIED reference
classes (oracle/_ref/ref_oracle, built in place from /root/reference by
oracle/Makefile) on seeded synthetic WAV input.  Run in the build container:

    make -C oracle ref && python tests/golden/make_golden.py

Each case directory holds:
  input.npz      the int16 I/Q fed to the reference (and its sample rate)
  config.json    the radio-observer config used (reference JSON dialect)
  rows_sub.npy   full rows for a subset of row indices (rows_idx.npy)
  bands.npy      band-cropped rows [lo, hi) for ALL rows (bands_range in meta.json)
  npa.npy        per-row (noise, peakFreq, magnitude) from NoiseMessage
  rowinfo.npy    per-row (mark, rawMark, sec, usec) handles
  bolid.txt      BolidMessage dump
  fits.json      name -> {header cards, data sha256, shape} of every FITS written
  meta.json      row count etc.
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from radio_observer_b200 import synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")


def config_json(bins, overlap, snap, bolid, origin="gold"):
    rec = []
    if snap:
        rec.append(dict(key="recorder", factory="snapshot", output_dir="OUT", output_type="snap",
                        compress_output=False, **snap))
    if bolid:
        rec.append(dict(key="recorder", factory="bolid", output_dir="OUT", output_type="blid",
                        compress_output=False, **bolid))
    return {
        "configuration": "default",
        "configurations": [{
            "key": "default", "factory": "pipeline",
            "children": [{
                "key": "backend", "factory": "waterfall", "bins": bins, "overlap": overlap,
                "origin": origin, "iq_gain": 0, "iq_phase_shift": 0, "metadata_path": "OUT",
                "children": rec}]}]}


def read_fits(path):
    b = open(path, "rb").read()
    cards, pos, end = [], 0, False
    while not end:
        block = b[pos:pos + 2880]
        pos += 2880
        for i in range(0, 2880, 80):
            c = block[i:i + 80].decode("ascii")
            if c.startswith("END"):
                end = True
                break
            cards.append(c.rstrip())
    hdr = {}
    for c in cards:
        if c[8:10] == "= ":
            hdr[c[:8].strip()] = c[10:].split(" /")[0].strip()
    naxis = int(hdr.get("NAXIS", "0"))
    shape = [int(hdr["NAXIS%d" % (i + 1)]) for i in range(naxis)]
    n = int(np.prod(shape)) if shape else 0
    data = np.frombuffer(b, dtype=">f4", count=n, offset=pos)
    return cards, hdr, shape, data


CASES = {
    # small shape: bit-level pinning of the whole chain incl. recorders
    "n2048": dict(
        sr=48000, bins=2048, overlap=1536, seconds=6.0, seed=0x5EED0001,
        carrier=11400.3,
        meteors=[(0.9, 5000.0, 0.25, 10400.0, 150.0), (2.6, 2500.0, 0.10, 10700.0, -100.0),
                 (4.2, 6000.0, 0.40, 10550.0, 60.0)],
        snap=dict(snapshot_length=1, low_freq=10100, hi_freq=11000),
        bolid=dict(snapshot_length=2, low_freq=9000, hi_freq=12000, low_detect_freq=10300,
                   hi_detect_freq=10900, low_noise_freq=9000, hi_noise_freq=10000,
                   advance_time=0.2, jitter_time=0.3, avg_freq_range=300),
        band=(1400, 1600), sub_step=7),
    # the benchmark shape (BASELINE config 2: Bolidozor bands at 16k/75 %)
    "c2b_16k": dict(
        sr=96000, bins=16384, overlap=12288, seconds=8.0, seed=0x5EED0003,
        carrier=26700.3,
        meteors=[(1.5, 5000.0, 0.3, 26500.0, 120.0), (5.0, 3000.0, 0.2, 26480.0, -80.0)],
        snap=dict(snapshot_length=2, low_freq=26200, hi_freq=26800),
        bolid=dict(snapshot_length=4, low_freq=26200, hi_freq=26800, low_detect_freq=26450,
                   hi_detect_freq=26550, low_noise_freq=26000, hi_noise_freq=26300,
                   advance_time=0.5, jitter_time=0.5, avg_freq_range=40),
        band=(12600, 12800), sub_step=23),
}


def make_case(name, c):
    out = os.path.join(HERE, name)
    shutil.rmtree(out, ignore_errors=True)
    os.makedirs(out)
    n = int(c["seconds"] * c["sr"])
    n -= n % 1024                      # whole 1024-frame WAV blocks (reference tail quirk Q1)
    spec = synth.SynthSpec(
        sample_rate=c["sr"], n_samples=n, seed=c["seed"], carrier_freq=c["carrier"],
        meteors=[synth.Meteor(start=int(t * c["sr"]), amp=a, tau=tau, freq=f, rate=r)
                 for (t, a, tau, f, r) in c["meteors"]])
    iq = synth.generate_int16(spec)
    tmp = tempfile.mkdtemp(prefix="ro_gold_")
    try:
        os.makedirs(os.path.join(tmp, "OUT"))
        os.makedirs(os.path.join(tmp, "dump"))
        cfg = config_json(c["bins"], c["overlap"], c["snap"], c["bolid"])
        txt = json.dumps(cfg, indent=1).replace('"OUT"', '"%s"' % os.path.join(tmp, "OUT"))
        open(os.path.join(tmp, "cfg.json"), "w").write(txt)
        synth.write_wav(os.path.join(tmp, "in.wav"), iq, c["sr"])
        # The reference starts its recorder threads from a base-class constructor
        # (cppapp/cppapp/Thread.cpp:47-56), so a run occasionally dies with "pure
        # virtual method called" before processing anything: retry.
        for attempt in range(20):
            r = subprocess.run([REF, "-c", os.path.join(tmp, "cfg.json"), "-d", os.path.join(tmp, "dump"),
                                "-r", "-q", os.path.join(tmp, "in.wav")], cwd=tmp,
                               capture_output=True, text=True)
            if r.returncode == 0:
                break
        else:
            raise SystemExit("ref_oracle failed rc=%d: %s %s" % (r.returncode, r.stdout[-500:], r.stderr[-2000:]))
        bins = c["bins"]
        rows = np.fromfile(os.path.join(tmp, "dump", "rows.f32"), dtype=np.float32).reshape(-1, bins)
        info = np.fromfile(os.path.join(tmp, "dump", "rowinfo.bin"),
                           dtype=np.dtype([("mark", "<i4"), ("rawmark", "<i4"), ("sec", "<i8"), ("usec", "<i8")]))
        npa = np.fromfile(os.path.join(tmp, "dump", "noise.bin"),
                          dtype=np.dtype([("noise", "<f4"), ("peakfreq", "<f4"), ("mag", "<f4")]))
        idx = np.arange(0, len(rows), c["sub_step"])
        np.savez_compressed(os.path.join(out, "input.npz"), iq=iq, sample_rate=c["sr"])
        json.dump(cfg, open(os.path.join(out, "config.json"), "w"), indent=1)
        np.save(os.path.join(out, "rows_idx.npy"), idx)
        np.save(os.path.join(out, "rows_sub.npy"), rows[idx])
        lo, hi = c["band"]
        np.save(os.path.join(out, "bands.npy"), rows[:, lo:hi])
        np.save(os.path.join(out, "npa.npy"), npa)
        np.save(os.path.join(out, "rowinfo.npy"), info)
        shutil.copy(os.path.join(tmp, "dump", "bolid.txt"), os.path.join(out, "bolid.txt"))
        fits = {}
        for fn in sorted(os.listdir(os.path.join(tmp, "OUT"))):
            p = os.path.join(tmp, "OUT", fn)
            if fn.endswith(".fits"):
                cards, hdr, shape, data = read_fits(p)
                fits[fn] = dict(cards=[x for x in cards if not x.startswith("DATE    =")],
                                shape=shape, sha256=hashlib.sha256(data.tobytes()).hexdigest())
            elif fn.endswith(".csv"):
                fits[fn] = dict(text=open(p).read())
        json.dump(fits, open(os.path.join(out, "fits.json"), "w"), indent=1)
        json.dump(dict(rows=int(len(rows)), bins=bins, overlap=c["overlap"], sample_rate=c["sr"],
                       band=list(c["band"]), n_samples=n, stdout=r.stdout[-4000:],
                       input_sha256=hashlib.sha256(iq.tobytes()).hexdigest()),
                  open(os.path.join(out, "meta.json"), "w"), indent=1)
        print(name, "rows", len(rows), "fits", len([k for k in fits if k.endswith(".fits")]),
              "bolids", open(os.path.join(out, "bolid.txt")).read().count("\n"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    for name, c in CASES.items():
        if len(sys.argv) > 1 and name not in sys.argv[1:]:
            continue
        make_case(name, c)
