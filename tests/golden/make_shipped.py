#!/usr/bin/env python3
"""tests/golden/make_shipped.py -- TEST INFRASTRUCTURE.

The three station configurations the reference ships, run end to end through the UNMODIFIED reference
(oracle/_ref/ref_oracle) on seeded synthetic station I/Q, and what it produced committed as fixtures:

  shipped/<name>.json     the shipped file (radio-observer.json, Bolidozor.json, Ionozor.json of
                          /root/reference) byte for byte -- comments, trailing commas, unread keys and
                          all -- except: every "output_dir" / "metadata_path" value -> the placeholder
                          directory OUT, "snapshot_length" -> 20 (so that two minutes of input give five
                          snapshots; the FFT, band and detector keys are untouched)
  shipped/<name>/         meta.json (input recipe + sha256, row count), rowinfo.npy, npa.npy, bolid.txt,
                          fits.json (every file: cards, shape, sha256 of the pixels; raw-I/Q files also
                          first_sample (its place in the input),
                          head_rows and the sha256 of that many leading samples),
                          band_rows.npy (the snapshot band of every 8th row)

Run in the build container:  make -C oracle ref && python tests/golden/make_shipped.py
The tests (tests/test_shipped_configs.py) regenerate the input from the recipe, check its sha256, and
run the same text through the stock backend and -- with "factory": "waterfall" -> "waterfall_cuda" as
the only edit -- through the CUDA backend.
"""
import hashlib
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from radio_observer_b200 import synth  # noqa: E402
from make_golden import read_fits      # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import raws_head_rows     # noqa: E402

REFERENCE = "/root/reference"
REF = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")
OUT = os.path.join(HERE, "shipped")

ROWINFO = np.dtype([("mark", "<i4"), ("rawmark", "<i4"), ("sec", "<i8"), ("usec", "<i8")])
NPA = np.dtype([("noise", "<f4"), ("peakfreq", "<f4"), ("mag", "<f4")])

# (t s, amplitude, tau s, start Hz, chirp Hz/s)
CASES = {
    "radio-observer": dict(   # C1: 48 kHz WAV, 32768 / 24576, snapshot 10100-11000 Hz + bolid recorder
        sr=48000, seconds=120.0, seed=0x5EED00C1, carrier=10950.3, bins=32768, overlap=24576,
        snap_band_hz=(10100.0, 11000.0),
        meteors=[(14.0, 5000.0, 0.6, 10420.0, 90.0), (47.5, 2600.0, 0.3, 10700.0, -120.0), (83.0, 6000.0, 1.1, 10555.0, 40.0)]),
    "Bolidozor": dict(        # C2p: 96 kHz, 65536 / 49152, Bolidozor bands
        sr=96000, seconds=120.0, seed=0x5EED00C2, carrier=26700.3, bins=65536, overlap=49152,
        snap_band_hz=(26200.0, 26800.0),
        meteors=[(11.0, 5200.0, 0.5, 26500.0, 100.0), (52.0, 3000.0, 0.25, 26480.0, -80.0), (90.0, 5600.0, 0.9, 26520.0, 30.0)]),
    "Ionozor": dict(          # C3p: the "VLF" configuration, 96 kHz, 32768 / 16384, snapshot recorder only
        sr=96000, seconds=120.0, seed=0x5EED00C3, carrier=22000.3, bins=32768, overlap=16384,
        snap_band_hz=(15000.0, 30000.0),
        meteors=[(20.0, 4000.0, 0.8, 18000.0, 150.0), (70.0, 4500.0, 0.4, 27000.0, -200.0)]),
}


def shipped_text(name):
    txt = open(os.path.join(REFERENCE, name + ".json")).read()
    txt = re.sub(r'("(?:output_dir|metadata_path)"\s*:\s*)"[^"]*"', r'\1"OUT"', txt)
    txt = re.sub(r'("snapshot_length"\s*:\s*)\d+', r'\g<1>20', txt)
    return txt


def station_iq(c):
    n = int(c["seconds"] * c["sr"])
    n -= n % 1024
    spec = synth.SynthSpec(sample_rate=c["sr"], n_samples=n, seed=c["seed"], carrier_freq=c["carrier"],
                           meteors=[synth.Meteor(start=int(t * c["sr"]), amp=a, tau=tau, freq=f, rate=r)
                                    for (t, a, tau, f, r) in c["meteors"]])
    return synth.generate_int16(spec)


def run_reference(binary, cfg_text, iq, sr, tmp, flags=(), dump_rows=True):
    """One Pipeline run; returns rows (or None), rowinfo, npa, bolid text, {file: (cards, shape, data)}, csv texts."""
    out, dump = os.path.join(tmp, "OUT"), os.path.join(tmp, "dump")
    wav = os.path.join(tmp, "in.wav")
    if not os.path.exists(wav):
        synth.write_wav(wav, iq, sr)
    open(os.path.join(tmp, "cfg.json"), "w").write(cfg_text.replace('"OUT"', '"%s"' % out))
    for attempt in range(60):     # the reference's start-up race (cppapp/cppapp/Thread.cpp:47-56), seen in bursts
        for d in (out, dump):
            shutil.rmtree(d, ignore_errors=True)
            os.makedirs(d)
        r = subprocess.run([binary, "-c", os.path.join(tmp, "cfg.json"), "-d", dump, *(("-r",) if dump_rows else ()), "-q",
                            *flags, wav], cwd=tmp, capture_output=True, text=True)
        if r.returncode == 0:
            break
    else:
        raise RuntimeError("%s failed rc=%d: %s" % (binary, r.returncode, r.stderr[-2000:]))
    timing = json.load(open(os.path.join(dump, "timing.json")))
    rows = None
    if dump_rows:
        rows = np.fromfile(os.path.join(dump, "rows.f32"), dtype=np.float32).reshape(-1, timing["bins"])
    info = np.fromfile(os.path.join(dump, "rowinfo.bin"), dtype=ROWINFO)
    npa = np.fromfile(os.path.join(dump, "noise.bin"), dtype=NPA)
    bolid = open(os.path.join(dump, "bolid.txt")).read()
    fits, csv = {}, {}
    skip = ("DATE    =", "COMMENT Local time", "COMMENT Git version")
    for fn in sorted(os.listdir(out)):
        p = os.path.join(out, fn)
        if fn.endswith(".fits"):
            cards, hdr, shape, data = read_fits(p)
            fits[fn] = dict(cards=[c for c in cards if not c.startswith(skip)], shape=shape, data=np.array(data, dtype=np.float32))
        elif fn.endswith(".csv"):
            csv[fn] = open(p).read()
    return dict(rows=rows, info=info, npa=npa, bolid=bolid, fits=fits, csv=csv, timing=timing, stderr=r.stderr)


def _card_value(cards, key):
    for card in cards:
        if card.startswith(key.ljust(8) + "="):
            return card.split("=", 1)[1].split("/")[0].strip()
    raise KeyError(key)


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, c in CASES.items():
        if len(sys.argv) > 1 and name not in sys.argv[1:]:
            continue
        txt = shipped_text(name)
        open(os.path.join(OUT, name + ".json"), "w").write(txt)
        d = os.path.join(OUT, name)
        shutil.rmtree(d, ignore_errors=True)
        os.makedirs(d)
        iq = station_iq(c)
        tmp = tempfile.mkdtemp(prefix="ro_shipped_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        try:
            r = run_reference(REF, txt, iq, c["sr"], tmp)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
        np.save(os.path.join(d, "rowinfo.npy"), r["info"])
        np.save(os.path.join(d, "npa.npy"), r["npa"])
        open(os.path.join(d, "bolid.txt"), "w").write(r["bolid"])
        fits = {fn: dict(cards=f["cards"], shape=f["shape"], sha256=hashlib.sha256(f["data"].tobytes()).hexdigest())
                for fn, f in r["fits"].items()}
        # raw-I/Q files: the tail is read ahead of the producer (helpers.raws_head_rows), so the part that is
        # a fixed function of the input gets its own hash; the full hash above is of a run whose tail was
        # complete (checked: the file is the input, sample for sample)
        for fn, f in r["fits"].items():
            if fn.endswith("_raws.fits"):
                head = raws_head_rows(f["shape"][1], c["sr"], c["bins"], c["overlap"])
                px = f["data"].reshape(-1, 2)
                src = iq.reshape(-1, 2).astype(np.float32)
                t = int(_card_value(f["cards"], "CRVAL2")) * c["sr"] // 1000
                at = [k for k in range(max(0, t - 2 * c["bins"]), t + 2 * c["bins"]) if np.array_equal(src[k:k + 64], px[:64])]
                assert len(at) == 1 and np.array_equal(src[at[0]:at[0] + len(px)], px), "%s: tail not complete, run again" % fn
                fits[fn].update(first_sample=at[0], head_rows=head, sha256_head=hashlib.sha256(px[:head].tobytes()).hexdigest())
        json.dump(fits, open(os.path.join(d, "fits.json"), "w"), indent=1)
        # the snapshot band of every 8th row (bins from the first snap file's NAXIS1 and the config's low_freq)
        snap = [f for fn, f in r["fits"].items() if fn.endswith("_snap.fits") and f["shape"][0] < c["bins"]]
        width = snap[0]["shape"][0]
        f32 = np.float32
        q = f32(f32(c["snap_band_hz"][0]) / f32(c["sr"]))
        lo = min(max(int(float(f32(c["bins"])) * (float(q) + 0.5)), 0), c["bins"] - 1)
        np.save(os.path.join(d, "band_rows.npy"), r["rows"][::8, lo:lo + width])
        json.dump(dict(rows=int(len(r["rows"])), input_sha256=hashlib.sha256(iq.tobytes()).hexdigest(), band=[lo, lo + width],
                       fits=len(fits), events=r["bolid"].count("\n"), timing=r["timing"],
                       recipe={k: v for k, v in c.items()}),
                  open(os.path.join(d, "meta.json"), "w"), indent=1)
        print(name, "rows", len(r["rows"]), "fits", len(fits), "events", r["bolid"].count("\n"), "band", lo, lo + width)


if __name__ == "__main__":
    main()
