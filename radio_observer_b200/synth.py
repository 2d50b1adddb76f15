"""Synthetic I/Q station data: carrier tone + Doppler-chirped,
exponentially decaying meteor echoes + Gaussian noise, as interleaved int16
[I0, Q0, I1, Q1, ...] (the WAV frontend's encoding, reference
src/WAVStream.cpp:115-121) or float32 pairs (raw frontend,
src/RawStream.cpp:59-63).

The noise comes from a counter-based generator (numpy Philox, key = seed) so a
run is reproducible from (seed, n_samples).  Parity runs generate the bytes once
and feed the same buffer to both sides.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

SEED_PARITY = 0x5EED0001
SEED_BENCH = 0x5EED0002


@dataclass
class Meteor:
    start: int          # sample index of the echo head
    amp: float          # peak amplitude, LSB
    tau: float          # decay time constant, s
    freq: float         # start frequency, Hz
    rate: float         # chirp rate, Hz/s


@dataclass
class SynthSpec:
    sample_rate: int = 48000
    n_samples: int = 1 << 20
    seed: int = SEED_PARITY
    sigma: float = 300.0                 # noise LSB per component
    carrier_amp: float = 3000.0
    carrier_freq: float = 10950.3
    meteor_band: Tuple[float, float] = (10300.0, 10900.0)
    meteor_mean_interval: float = 20.0   # s, Poisson arrivals
    meteor_amp: Tuple[float, float] = (600.0, 6000.0)
    meteor_tau: Tuple[float, float] = (0.05, 2.0)
    meteor_rate: float = 200.0           # |chirp| upper bound, Hz/s
    meteors: Optional[List[Meteor]] = None   # explicit list overrides the Poisson draw


def draw_meteors(spec: SynthSpec) -> List[Meteor]:
    if spec.meteors is not None:
        return list(spec.meteors)
    rng = np.random.Generator(np.random.Philox(key=spec.seed ^ 0xA5A5A5A5))
    out: List[Meteor] = []
    t = 0.0
    dur = spec.n_samples / spec.sample_rate
    while True:
        t += rng.exponential(spec.meteor_mean_interval)
        if t >= dur:
            break
        lo, hi = spec.meteor_amp
        amp = float(np.exp(rng.uniform(np.log(lo), np.log(hi))))
        out.append(Meteor(
            start=int(t * spec.sample_rate),
            amp=amp,
            tau=float(rng.uniform(*spec.meteor_tau)),
            freq=float(rng.uniform(*spec.meteor_band)),
            rate=float(rng.uniform(-spec.meteor_rate, spec.meteor_rate)),
        ))
    return out


def generate_complex(spec: SynthSpec) -> np.ndarray:
    """The station signal as complex128, before quantisation."""
    n = spec.n_samples
    sr = float(spec.sample_rate)
    rng = np.random.Generator(np.random.Philox(key=spec.seed))
    x = np.empty(n, dtype=np.complex128)
    block = 1 << 20
    for s in range(0, n, block):
        e = min(n, s + block)
        g = rng.standard_normal(2 * (e - s))
        x[s:e] = spec.sigma * (g[0::2] + 1j * g[1::2])
        idx = np.arange(s, e, dtype=np.float64)
        x[s:e] += spec.carrier_amp * np.exp(2j * np.pi * ((spec.carrier_freq / sr) * idx % 1.0))
    for m in draw_meteors(spec):
        # echo is negligible after 12 time constants
        stop = min(n, m.start + int(12 * m.tau * sr) + 1)
        if stop <= m.start:
            continue
        k = np.arange(0, stop - m.start, dtype=np.float64)
        phase = (m.freq * k + 0.5 * m.rate * k * k / sr) / sr
        x[m.start:stop] += m.amp * np.exp(-k / (m.tau * sr)) * np.exp(2j * np.pi * (phase % 1.0))
    return x


def generate_int16(spec: SynthSpec) -> np.ndarray:
    """Interleaved int16 [I, Q] of shape (2 * n_samples,), rounded to nearest, saturated."""
    x = generate_complex(spec)
    out = np.empty(2 * spec.n_samples, dtype=np.int16)
    out[0::2] = np.clip(np.rint(x.real), -32767, 32767).astype(np.int16)
    out[1::2] = np.clip(np.rint(x.imag), -32767, 32767).astype(np.int16)
    return out


def generate_float32(spec: SynthSpec, scale: float = 1.0 / 32768.0) -> np.ndarray:
    """Interleaved float32 [I, Q]; scale mimics a normalised SDR stream."""
    x = generate_complex(spec) * scale
    out = np.empty(2 * spec.n_samples, dtype=np.float32)
    out[0::2] = x.real.astype(np.float32)
    out[1::2] = x.imag.astype(np.float32)
    return out


def write_wav(path: str, iq: np.ndarray, sample_rate: int, reference_safe: bool = True) -> None:
    """16-bit stereo PCM RIFF/WAVE, the only layout the reference's WAV frontend
    accepts (src/WAVStream.cpp:62-83,101-106).

    reference_safe: the reference's chunk walker subtracts size+4 per subchunk
    although it consumed size+8 (src/WAVStream.cpp:178, 234-238), so with the
    standard RIFF size (36 + data) it runs past EOF and re-processes the last
    block from stale buffers.  Writing the RIFF size as 28 + data makes its
    accounting end exactly at EOF; other readers ignore the field.
    """
    iq = np.ascontiguousarray(iq, dtype="<i2")
    data = iq.tobytes()
    with open(path, "wb") as f:
        f.write(b"RIFF")
        f.write(struct.pack("<I", (28 if reference_safe else 36) + len(data)))
        f.write(b"WAVE")
        f.write(b"fmt ")
        f.write(struct.pack("<IHHIIHH", 16, 1, 2, sample_rate, sample_rate * 4, 4, 16))
        f.write(b"data")
        f.write(struct.pack("<I", len(data)))
        f.write(data)


def read_wav(path: str) -> Tuple[np.ndarray, int]:
    """Minimal reader for files produced by write_wav (and plain PCM WAVs)."""
    with open(path, "rb") as f:
        b = f.read()
    assert b[0:4] == b"RIFF" and b[8:12] == b"WAVE"
    pos = 12
    sr = None
    while pos + 8 <= len(b):
        cid = b[pos:pos + 4]
        size = struct.unpack("<I", b[pos + 4:pos + 8])[0]
        if cid == b"fmt ":
            _, ch, sr, _, _, bits = struct.unpack("<HHIIHH", b[pos + 8:pos + 24])
            assert ch == 2 and bits == 16
        elif cid == b"data":
            return np.frombuffer(b, dtype="<i2", count=size // 2, offset=pos + 8).copy(), int(sr)
        pos += 8 + size
    raise ValueError("no data chunk")
