#!/usr/bin/env python3
"""bench.py -- throughput of the waterfall hot path on B200 ("Measurement").

Workload (BASELINE.json config 2): 96 kHz int16 I/Q, 16384-point
FFT, overlap 12288 (75 %), Bolidozor recorder bands, bolid detector statistics per row.
One step = one pass of the hot path (fused unpack+window+FFT+magnitude kernel, then the
detector kernel) over one batch of synthetic station data per GPU: --passes shards of
2^--log2-samples samples each, taken in turn from --shards distinct shards resident in HBM,
rows of every shard written to the same device buffer (as the library's own host path does).
The default batch makes a step about 100 ms, so that the --steps 20 the driver asks for are
2 s of steady state under the power cap, not a 50 ms burst.

  value     complex input Msamples/s, inputs resident in HBM (whole job, all ranks)
  e2e       the same through ro_wf_process_host: pinned host I/Q in, recorder-band
            rows + detector records out, H2D/D2H inside the timed region
  roofline  algorithmic HBM bytes of the fused kernel / its CUDA-event duration (a second,
            separately profiled pass: the timed loop carries no per-launch events)
  parity    the timed input's first frames against the oracle, the tests' rule and the strict one
  fp64      the same shape through the FP64 variant (cfg.precision = 1)
  cpu_baseline  the reference's own C++ (oracle/_ref/ref_oracle) on this box's host cores;
  cpu_baseline_alt  the same path restated on scipy.fft (pocketfft), the -O0 build, an FFTW probe
  dropin    the reference's Pipeline binary with "factory": "waterfall_cuda"

`--impl reference` times only the reference CPU implementation (rank 0).
`--sweep` prints BASELINE config 5 (FFT size x overlap, device-resident) with a CPU figure per point.
"""
import argparse
import ctypes
import ctypes.util
import json
import math
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 96000
BINS = 16384
OVERLAP = 12288
HOP = BINS - OVERLAP
# Bolidozor.json recorder keys (reference Bolidozor.json:74-93)
SNAP_BAND_HZ = (26200.0, 26800.0)
DETECT_HZ = (26450.0, 26550.0)
NOISE_HZ = (26000.0, 26300.0)
AVG_HZ = 40.0
WORKLOAD = "C2b: Bolidozor bands, 96 kHz int16 I/Q, 16384-pt FFT, overlap 12288 (75%)"
METRIC = "complex Msamples/s through waterfall FFT+magnitude+bolid detect"


def bench_config(log2_samples, passes, shards):
    """The `config` object of both arms (the driver compares them)."""
    S = 1 << log2_samples
    F = (S - BINS) // HOP + 1
    return {"workload": WORKLOAD, "bins": BINS, "overlap": OVERLAP, "sample_rate": SR,
            "samples_per_gpu_per_step": S * passes, "frames_per_gpu_per_step": F * passes,
            "shard_samples": S, "shards_per_step": passes, "distinct_shards": shards,
            "sharding": "time shards, one per GPU, overlap halo re-read from input; no collective",
            "l2": "every pass reads another %d MiB shard and writes %d MiB of rows: larger than L2" % (4 * S >> 20, 4 * BINS * F >> 20)}


# ----------------------------------------------------------------------------
# host-side Hz -> bin (reference src/FFTBackend.h:159-178), restated for the bench
# ----------------------------------------------------------------------------
def freq_to_bin(freq, sr, bins):
    f32 = np.float32
    q = f32(f32(freq) / f32(sr))
    v = float(f32(bins)) * (float(q) + 0.5)
    return min(max(int(v), 0), bins - 1)


def bolidozor_bands(sr=SR, bins=BINS):
    lo, hi = (freq_to_bin(f, sr, bins) for f in SNAP_BAND_HZ)
    dlo, dhi = (freq_to_bin(f, sr, bins) for f in DETECT_HZ)
    nlo, nhi = (freq_to_bin(f, sr, bins) for f in NOISE_HZ)
    avg = max(1, freq_to_bin(AVG_HZ, sr, bins) - freq_to_bin(0.0, sr, bins))
    return (lo, hi), dict(det_lo=dlo, det_w=dhi - dlo, noise_lo=nlo, noise_w=nhi - nlo, avg_bins=avg)


# ----------------------------------------------------------------------------
# synthetic station signal generated on the device (tones + chirped echoes + noise)
# ----------------------------------------------------------------------------
def gen_iq_device(torch, n_samples, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(2 * n_samples, dtype=torch.int16, device=device)
    rng = np.random.default_rng(seed)
    dur = n_samples / SR
    n_met = max(1, int(dur / 20.0))
    met = [(int(rng.uniform(0, dur) * SR), float(np.exp(rng.uniform(np.log(600), np.log(6000)))),
            float(rng.uniform(0.05, 2.0)), float(rng.uniform(26450, 26550)), float(rng.uniform(-200, 200)))
           for _ in range(n_met)]
    chunk = 1 << 22
    for s in range(0, n_samples, chunk):
        e = min(n_samples, s + chunk)
        idx = torch.arange(s, e, dtype=torch.float64, device=device)
        ph = torch.remainder(idx * (26700.3 / SR), 1.0) * (2 * math.pi)
        re = 3000.0 * torch.cos(ph)
        im = 3000.0 * torch.sin(ph)
        for (st, amp, tau, f0, rate) in met:
            stop = st + int(12 * tau * SR)
            a, b = max(s, st), min(e, stop)
            if a >= b:
                continue
            k = torch.arange(a - st, b - st, dtype=torch.float64, device=device)
            p = torch.remainder((f0 * k + 0.5 * rate * k * k / SR) / SR, 1.0) * (2 * math.pi)
            env = amp * torch.exp(-k / (tau * SR))
            re[a - s:b - s] += env * torch.cos(p)
            im[a - s:b - s] += env * torch.sin(p)
        nz = torch.randn(2, e - s, generator=g, device=device, dtype=torch.float32) * 300.0
        re = torch.clamp(torch.round(re.float() + nz[0]), -32767, 32767).to(torch.int16)
        im = torch.clamp(torch.round(im.float() + nz[1]), -32767, 32767).to(torch.int16)
        out[2 * s:2 * e:2] = re
        out[2 * s + 1:2 * e:2] = im
    return out


# ----------------------------------------------------------------------------
# clocks sampler
# ----------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, pw, smax, reasons = [], [], None, set()
        for (t, l) in self.lines:
            if t < t0 or t > t1:
                continue
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
                pw.append(float(f[3]))
            except Exception:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": float(min(sm)) if sm else None,
                "sm_max_mhz": smax, "power_w_median": float(np.median(pw)) if pw else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------
# reference CPU arm
# ----------------------------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")
REF_BIN_O0 = os.path.join(ROOT, "oracle", "_ref", "ref_oracle_O0")
CUDA_BIN = os.path.join(ROOT, "radio_observer_b200", "host", "_build", "radio-observer-cuda")
FFT_NOTE = ("FFTW and CFITSIO are absent from this image: the reference's classes are linked against a scalar "
            "radix-4 Stockham FFT (oracle/shims/fftw3_shim.cpp, about pocketfft's speed, not SIMD FFTW) and an "
            "uncompressed FITS writer; cpu_baseline_alt gives the scipy.fft figure of the same path beside it")


def ref_config_json(out_dir, factory="waterfall", extra=None):
    cfg = _ref_config(out_dir)
    be = cfg["configurations"][0]["children"][0]
    be["factory"] = factory
    be.update(extra or {})
    return json.dumps(cfg, indent=1)


def _ref_config(out_dir):
    return ({
        "configuration": "default",
        "configurations": [{
            "key": "default", "factory": "pipeline",
            "children": [{
                "key": "backend", "factory": "waterfall", "bins": BINS, "overlap": OVERLAP,
                "origin": "bench", "iq_gain": 0, "iq_phase_shift": 0, "metadata_path": out_dir,
                "children": [
                    {"key": "recorder", "factory": "snapshot", "output_dir": out_dir, "output_type": "snap",
                     "compress_output": False, "snapshot_length": 60,
                     "low_freq": SNAP_BAND_HZ[0], "hi_freq": SNAP_BAND_HZ[1]},
                    {"key": "recorder", "factory": "bolid", "output_dir": out_dir, "output_type": "blid",
                     "compress_output": False, "snapshot_length": 60, "advance_time": 5, "jitter_time": 2,
                     "low_freq": SNAP_BAND_HZ[0], "hi_freq": SNAP_BAND_HZ[1],
                     "low_detect_freq": DETECT_HZ[0], "hi_detect_freq": DETECT_HZ[1],
                     "low_noise_freq": NOISE_HZ[0], "hi_noise_freq": NOISE_HZ[1], "avg_freq_range": AVG_HZ},
                ]}]}]})


def host_station_iq(n_samples, seed):
    from radio_observer_b200 import synth
    spec = synth.SynthSpec(sample_rate=SR, n_samples=n_samples, seed=seed, carrier_freq=26700.3,
                           meteor_band=(26450.0, 26550.0))
    return synth.generate_int16(spec)


class RefRunner:
    """The reference's own Pipeline binary on `procs` concurrent copies of one WAV in tmpfs."""

    def __init__(self, sample_seconds, procs, binary=None, factory="waterfall", extra=None, wav_path=None, flags=()):
        from radio_observer_b200 import synth
        self.binary = binary or REF_BIN
        self.n = int(sample_seconds * SR)
        self.n -= self.n % 1024
        self.procs = procs
        self.flags = flags
        base = "/dev/shm" if os.path.isdir("/dev/shm") else None
        self.tmp = tempfile.mkdtemp(prefix="ro_bench_", dir=base)
        self.wav = wav_path or os.path.join(self.tmp, "in.wav")
        if not wav_path:
            synth.write_wav(self.wav, host_station_iq(self.n, 0x5EED0002), SR)
        for p in range(procs):
            d = os.path.join(self.tmp, "p%d" % p)
            os.makedirs(os.path.join(d, "out"))
            open(os.path.join(d, "cfg.json"), "w").write(ref_config_json(os.path.join(d, "out"), factory, extra))
        self.crashes = 0

    def run_once(self):
        """Returns the wall time of the slowest process, or None if one of them kept failing."""
        times = [None] * self.procs

        def worker(p):
            d = os.path.join(self.tmp, "p%d" % p)
            # the reference has a start-up race (threads started from a base-class constructor, see
            # DESIGN.md) and sometimes dies before processing: run again, counted
            for attempt in range(60):
                shutil.rmtree(os.path.join(d, "out"), ignore_errors=True)
                os.makedirs(os.path.join(d, "out"))
                t0 = time.perf_counter()
                rc = subprocess.call([self.binary, "-c", os.path.join(d, "cfg.json"), "-q", *self.flags, self.wav],
                                     stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=d)
                if rc == 0:
                    times[p] = time.perf_counter() - t0
                    return
                self.crashes += 1

        ths = [threading.Thread(target=worker, args=(p,)) for p in range(self.procs)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        wall = time.perf_counter() - t0
        return wall if all(t is not None for t in times) else None

    def close(self):
        shutil.rmtree(self.tmp, ignore_errors=True)


def run_reference_cpu(sample_seconds, procs, runs=1, **kw):
    """(Msamples/s of the best run, list of all runs) for the reference binary, or None without it."""
    binary = kw.get("binary") or REF_BIN
    if not os.path.exists(binary):
        return None
    r = RefRunner(sample_seconds, procs, **kw)
    try:
        vals = []
        for _ in range(runs):
            w = r.run_once()
            if w is not None:
                vals.append(procs * r.n / w / 1e6)
        return (max(vals), [round(v, 2) for v in vals], r.n, r.crashes) if vals else None
    finally:
        r.close()


def numpy_port_baseline(seconds):
    """oracle/ref_numpy.py, single thread (the fallback when the reference binary is absent)."""
    from oracle import ref_numpy as R
    n = int(seconds * SR)
    iq = host_station_iq(n, 0x5EED0002)
    _, det = bolidozor_bands()
    d = R.DetectorBands(det["det_lo"], det["det_w"], det["noise_lo"], det["noise_w"], det["avg_bins"], 0, 0)
    t0 = time.perf_counter()
    rows = R.waterfall_rows(iq, BINS, OVERLAP)
    R.detector_rows(rows, d)
    return n / (time.perf_counter() - t0) / 1e6


def scipy_path(iq, bins, overlap, det, workers, chunk_frames=256):
    """The hot path restated on scipy.fft (pocketfft, SIMD, double): framing, window in double, FFT,
    magnitude, fftshift, float rows, then noise / peak / mean per row.  A timing vehicle for
    cpu_baseline_alt (the parity checker is oracle/ref_numpy.py).  Returns seconds."""
    import scipy.fft
    from oracle import ref_numpy as R
    hop = bins - overlap
    x = iq.astype(np.float64).view(np.complex128) if iq.dtype == np.float64 else (
        iq.reshape(-1, 2).astype(np.float64) @ np.array([1.0, 1j]))
    w = R.nuttall_window(bins).astype(np.float64)
    F = (len(x) - bins) // hop + 1
    frames = np.lib.stride_tricks.sliding_window_view(x, bins)[::hop][:F]
    t0 = time.perf_counter()
    acc = 0.0
    for f0 in range(0, F, chunk_frames):
        blk = frames[f0:f0 + chunk_frames] * w
        spec = scipy.fft.fft(blk, axis=1, workers=workers, overwrite_x=True)
        rows = np.fft.fftshift(np.abs(spec), axes=1).astype(np.float32)
        nb = rows[:, det["noise_lo"]:det["noise_lo"] + det["noise_w"]]
        noise = 2.0 * np.partition(nb, det["noise_w"] // 4, axis=1)[:, det["noise_w"] // 4]
        db = rows[:, det["det_lo"]:det["det_lo"] + det["det_w"]]
        peak = det["det_w"] - 1 - np.argmax(db[:, ::-1], axis=1)
        acc += float(noise.sum()) + float(peak.sum())
    return time.perf_counter() - t0, F


def fft_probe():
    """Per-transform time of a 16384-point complex128 FFT on one core: scipy.fft, and FFTW3 through
    ctypes when libfftw3 is on the box (the reference links -lfftw3, src/FFTBackend.cpp:117-120)."""
    import scipy.fft
    x = (np.random.default_rng(1).standard_normal(2 * BINS)).view(np.complex128).copy()
    scipy.fft.fft(x)
    t0 = time.perf_counter()
    for _ in range(200):
        scipy.fft.fft(x)
    out = {"scipy_fft_16k_us": round((time.perf_counter() - t0) / 200 * 1e6, 1), "libfftw3": None, "fftw_16k_us": None}
    path = ctypes.util.find_library("fftw3")
    out["libfftw3"] = path
    if path:
        try:
            lib = ctypes.CDLL(path)
            lib.fftw_plan_dft_1d.restype = ctypes.c_void_p
            lib.fftw_plan_dft_1d.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_uint]
            lib.fftw_execute.argtypes = [ctypes.c_void_p]
            y = np.empty_like(x)
            plan = lib.fftw_plan_dft_1d(BINS, x.ctypes.data, y.ctypes.data, -1, 1 << 6)      # FFTW_ESTIMATE
            lib.fftw_execute(plan)
            t0 = time.perf_counter()
            for _ in range(200):
                lib.fftw_execute(plan)
            out["fftw_16k_us"] = round((time.perf_counter() - t0) / 200 * 1e6, 1)
        except Exception as e:                      # noqa: BLE001
            out["fftw_16k_us"] = "probe failed: %s" % e
    return out


def reference_arm(args, emit):
    cores = os.cpu_count() or 1
    ref_seconds = float(os.environ.get("RO_BENCH_REF_SECONDS", "60"))     # tests shorten it
    cfg = bench_config(args.log2_samples, args.passes, args.shards)
    if os.path.exists(REF_BIN):
        r = RefRunner(ref_seconds, cores)
        try:
            for _ in range(args.warmup):
                r.run_once()
            walls = []
            for _ in range(args.steps):
                w = r.run_once()
                if w is not None:
                    walls.append(w)
            n, crashes = r.n, r.crashes
        finally:
            r.close()
        if not walls:
            emit({"impl": "reference", "unavailable": "oracle/_ref/ref_oracle kept failing on this box"})
            return 0
        ms = 1e3 * float(np.mean(walls))
        v = cores * n / (ms * 1e-3) / 1e6
        kind = "reference"
        desc = ("each step: %d concurrent processes, each %.0f s of 96 kHz station I/Q (WAV in tmpfs) through the "
                "reference's own Pipeline/WAVStream/WaterfallBackend/BolidRecorder (oracle/_ref, -O2), FITS to tmpfs, "
                "process start-up included; %d start-up crashes re-run. %s" % (cores, n / SR, crashes, FFT_NOTE))
    else:
        secs = min(ref_seconds, 20.0)
        vals = [numpy_port_baseline(secs) for _ in range(min(args.steps, 3))]
        v, cores, kind = float(np.median(vals)), 1, "port"
        ms = 1e3 * secs * SR / (v * 1e6)
        desc = "%.0f s of station I/Q through oracle/ref_numpy.py (numpy pocketfft), one thread" % secs
    emit({"impl": "reference", "metric": METRIC, "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
          "steps": len(walls) if kind == "reference" else min(args.steps, 3), "warmup": args.warmup, "ms_per_step": ms,
          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
          "config": cfg,
          "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": desc},
          "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
          "gpu_launches": 0})
    return 0


# ----------------------------------------------------------------------------
# rank placement: CPU cores next to the rank's GPU, before any pinned allocation
# ----------------------------------------------------------------------------
def bind_to_gpu_cpus(local_rank, world):
    info = {"numa_node": None, "cpus": None, "how": None}
    try:
        bus = subprocess.check_output(["nvidia-smi", "-i", str(local_rank), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                                      text=True).strip().lower()
        if bus.startswith("0000"):
            bus = bus[4:]                       # nvidia-smi prints an 8-digit domain, sysfs a 4-digit one
        base = "/sys/bus/pci/devices/" + bus
        node = int(open(base + "/numa_node").read())
        cpus = open(base + "/local_cpulist").read().strip()
        ids = []
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            ids += list(range(int(a), int(b or a) + 1))
        every = sorted(os.sched_getaffinity(0))
        ids = [c for c in ids if c in every]
        info["numa_node"] = node
        if ids and len(ids) < len(every):
            os.sched_setaffinity(0, ids)
            info.update(cpus=cpus, how="the GPU's local_cpulist")
        elif world > 1 and len(every) >= world:
            # one NUMA node reported for every GPU: give each rank its own slice of the cores, so that the
            # first touch of its pinned buffers and its copy thread stay put
            per = len(every) // world
            mine = every[local_rank * per:(local_rank + 1) * per]
            os.sched_setaffinity(0, mine)
            info.update(cpus="%d-%d" % (mine[0], mine[-1]), how="even slice of the cores (no per-GPU locality reported)")
    except Exception as e:                          # noqa: BLE001
        info["how"] = "unavailable: %s" % e
    return info


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2-samples", type=int, default=28, help="complex samples per shard (2^k)")
    ap.add_argument("--passes", type=int, default=40, help="shards processed per step")
    ap.add_argument("--shards", type=int, default=4, help="distinct shards resident in HBM, taken in turn")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="BASELINE config 5 instead of the headline line")
    args = ap.parse_args()
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version
    # banner from C on the first collective): keep the real stdout aside for the result line and point
    # fd 1 at stderr for everything else.
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(result_fd, (json.dumps(obj) + "\n").encode())
    if args.warmup < 3:
        args.warmup = 3

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        if rank != 0:
            return 0
        return reference_arm(args, emit)

    t_start = time.time()

    def note(msg):
        if rank == 0:
            print("[bench %6.1f s] %s" % (time.time() - t_start, msg), file=sys.stderr, flush=True)

    placement = bind_to_gpu_cpus(local_rank, world)
    note("placement %r" % (placement,))

    import torch
    import radio_observer_b200 as ro
    from radio_observer_b200.capi import DET_DTYPE, PinnedArray

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if args.sweep:
        return sweep(args, emit, torch, ro, dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- the shards of the archive this rank owns (time-sharded, overlap halo) ----
    S = 1 << args.log2_samples                       # samples per shard
    P, G = args.passes, max(1, min(args.shards, args.passes))
    band, det = bolidozor_bands()
    wf = ro.Waterfall(BINS, OVERLAP, sample_rate=SR, in_format=ro.IN_INT16, bands=[band], detector=det,
                      device=local_rank)
    F = wf.frame_count(S)
    d_iq = [gen_iq_device(torch, S, 0x5EED0002 + 1000 * rank + g, dev) for g in range(G)]
    torch.cuda.synchronize()
    note("%d shard(s) of 2^%d samples generated on the device" % (G, args.log2_samples))
    d_rows = torch.empty(F * BINS, dtype=torch.float32, device=dev)
    d_det = torch.empty(F * 4, dtype=torch.int32, device=dev)
    # a dedicated non-default stream: the default stream's handle is 0, which the C ABI
    # reads as "use the handle's own stream" and torch events would then time nothing
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    wf.set_stream(stream.cuda_stream)

    def step_resident():
        for p in range(P):
            wf.process_resident(d_iq[p % G].data_ptr(), S, d_rows.data_ptr(), d_det.data_ptr(), sync=False)

    for _ in range(args.warmup):
        step_resident()
    torch.cuda.synchronize()
    l0 = wf.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        step_resident()
    e1.record(stream)
    torch.cuda.synchronize()
    t_wall1 = time.time()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = wf.launch_count() - l0
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    ms_per_step = ms_total / args.steps
    value = world * S * P / (ms_per_step * 1e-3) / 1e6
    note("timed region done: %.3f ms per step" % ms_per_step)

    # ---- kernel times: a separate pass with the library's per-launch CUDA events on the same stream ----
    wf.set_profiling(True)
    wf.kernel_times()                                # clear
    for _ in range(2):
        step_resident()
    kms, kn = wf.kernel_times()
    wf.set_profiling(False)

    # ---- roofline of the dominant (fused FFT) kernel ----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    used_samples = (F - 1) * HOP + BINS
    alg_bytes = 4 * used_samples + 4 * BINS * F      # int16 I/Q read once + full float32 row written once
    fft_ms = kms[0] / max(1, kn[0])
    achieved = alg_bytes / (fft_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get("dram_bytes_per_frame", 0) * F or None
        except Exception:
            traffic = None
    flops = 5.0 * BINS * math.log2(BINS) * F
    sm_clock = (clocks or {}).get("sm_mhz") or 1965.0
    fp32_peak = 148 * 128 * 2 * sm_clock * 1e6 / 1e12      # TFLOP/s at the clock seen during the timed region
    tflops = flops / (fft_ms * 1e-3) / 1e12
    roofline = {"bound": "fp32-fma (the HBM fraction is the judged figure; at 75% overlap the FMA pipe binds first)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": wf.kernel_name(), "kernel_ms": fft_ms,
                "kernel_launches_profiled": int(kn[0]),
                "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                "fft_tflops_5NlogN": tflops,
                "fp32_frac": tflops / fp32_peak,
                "fp32_peak_tflops": fp32_peak,
                "fp32_peak_source": "148 SMs x 128 lanes x 2 x the median SM clock of the timed region (pipe rate measured: profiles/r01_microbench_pipe_rates.txt)",
                "detector_kernel_ms": kms[1] / max(1, kn[1]),
                "kernel_share_of_step": (kms[0] + kms[1]) / 2.0 / ms_per_step if ms_per_step else None,
                "fp32_roof": {"lane_ops_per_frame": 1698 * 512, "fma_pipe_floor_cycles_per_frame": 6792,
                              "note": "FMA-pipe floor of the kernel's own instruction mix (812 packed + 74 scalar FP instructions per thread and frame, 128 lanes/clk/SM) = 42.8 M frames/s at 1965 MHz = 0.537 of HBM peak at 20 B/sample"},
                "note": "75% overlap: 280 flop/sample puts this shape past the FP32 ridge; see DESIGN.md section 4"}

    # ---- end to end through the host API: pinned host in, bands + detector records out ----
    h_iq = PinnedArray((2 * S,), np.int16)
    h_iq.array[:] = d_iq[0].cpu().numpy()
    h_rows = PinnedArray((F, wf.row_width), np.float32)
    h_det = PinnedArray((F,), DET_DTYPE)
    wf.set_stream(None)
    # the ceiling of this leg: every rank copying its pinned shard to its GPU at the same time
    h2d_gbs = None
    try:
        d_tmp = torch.empty(2 * S, dtype=torch.int16, device=dev)
        h_t = torch.from_numpy(h_iq.array)
        d_tmp.copy_(h_t, non_blocking=True)
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(4):
            d_tmp.copy_(h_t, non_blocking=True)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        h2d_gbs = 4 * 4 * S / dt / 1e9
        del d_tmp
    except Exception:                               # noqa: BLE001
        h2d_gbs = None
    note("concurrent H2D: %s GB/s per GPU" % h2d_gbs)
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        wf.process_host(h_iq.array, h_rows.array, h_det.array)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        wf.process_host(h_iq.array, h_rows.array, h_det.array)
    torch.cuda.synchronize()
    dt = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * S * e2e_steps / dt / 1e6
    checksum = float(h_rows.array[:, 0].sum()) + float(h_det.array["avg"].sum())
    e2e = {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 4 * used_samples,
           "d2h_bytes_per_step": F * (4 * wf.row_width + 16), "steps": e2e_steps, "samples_per_step": S,
           "output": "recorder band rows (%d bins) + detector records" % wf.row_width, "checksum": checksum,
           "h2d_gbs_per_gpu_concurrent": h2d_gbs,
           "h2d_ceiling_msamples_per_s": (world * h2d_gbs * 1e9 / 4 / 1e6) if h2d_gbs else None,
           "placement": placement}

    note("e2e done: %.1f Msamples/s" % e2e_value)
    cpu_baseline = cpu_alt = dropin = parity = fp64 = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # ---- parity of the timed input: its first frames against the oracle (the checker), the tests' rule
        # and the strict one of BASELINE.md section 5 ----
        try:
            from oracle import ref_numpy as R
            nfr = 48
            ns = BINS + (nfr - 1) * HOP
            iq_head = h_iq.array[:2 * ns].copy()
            wfp = ro.Waterfall(BINS, OVERLAP, sample_rate=SR, in_format=ro.IN_INT16, device=local_rank)
            rows_gpu, _, _ = wfp.process_host(iq_head)
            wfp.close()
            ref = R.waterfall_rows(iq_head, BINS, OVERLAP).astype(np.float64)
            g64 = rows_gpu.astype(np.float64)
            err = np.abs(g64 - ref)
            tol = np.maximum((10.0 ** (1e-3 / 20.0) - 1.0) * np.abs(ref), 2.0 ** -22 * np.abs(ref).max(axis=1, keepdims=True))
            rms = np.sqrt(np.mean(ref * ref, axis=1, keepdims=True))
            strict = np.maximum(5e-6 * np.abs(ref), 1e-6 * rms)
            parity = {"rule": "per bin |gpu - ref| <= max(1e-3 dB x |ref|, 2^-22 x max|row|) (tests/helpers.py); "
                              "strict = 5e-6 relative or 1e-6 x RMS(row) (BASELINE.md section 5)",
                      "frames": nfr, "worst_err_over_tol": float((err / tol).max()),
                      "frac_bins_strict_5e-6": float((err <= strict).mean()),
                      "worst_err_over_strict": float((err / strict).max()), "precision": "fp32"}
            wf64 = ro.Waterfall(BINS, OVERLAP, sample_rate=SR, in_format=ro.IN_INT16, device=local_rank, precision=ro.PREC_FP64)
            rows64, _, _ = wf64.process_host(iq_head)
            e64 = np.abs(rows64.astype(np.float64) - ref)
            parity["fp64"] = {"worst_err_over_strict": float((e64 / strict).max()),
                              "frac_bins_strict_5e-6": float((e64 <= strict).mean()),
                              "frac_bins_bit_identical": float((rows64 == ref.astype(np.float32)).mean())}
            # ---- throughput of the FP64 variant on a smaller shard (multi-kernel, a global scratch) ----
            S64 = 1 << 24
            F64 = wf64.frame_count(S64)
            wf64.set_stream(stream.cuda_stream)
            for _ in range(2):
                wf64.process_resident(d_iq[0].data_ptr(), S64, d_rows.data_ptr(), None, sync=False)
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            for _ in range(5):
                wf64.process_resident(d_iq[0].data_ptr(), S64, d_rows.data_ptr(), None, sync=False)
            a1.record(stream)
            torch.cuda.synchronize()
            ms64 = a0.elapsed_time(a1) / 5
            b64 = 4 * ((F64 - 1) * HOP + BINS) + 4 * BINS * F64
            fp64 = {"value": S64 / (ms64 * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_shard": ms64, "samples": S64,
                    "roofline_frac": b64 / (ms64 * 1e-3) / 1e9 / peak, "dtype": "f64",
                    "what": "cfg.precision = 1 (wf64_* kernels): the variant that meets the strict tolerance on every bin"}
            wf64.close()
        except Exception as e:                      # noqa: BLE001
            parity = {"error": "%s: %s" % (type(e).__name__, e)}

        note("parity / fp64 done: %r" % (parity,))
        # ---- CPU baselines: 30 min of station data, about 12 s of work for the reference on one core ----
        secs = 1800.0
        from radio_observer_b200 import synth
        base = "/dev/shm" if os.path.isdir("/dev/shm") else None
        wtmp = tempfile.mkdtemp(prefix="ro_bench_wav_", dir=base)
        try:
            wav = None
            if os.path.exists(REF_BIN):
                n = int(secs * SR)
                n -= n % 1024
                wav = os.path.join(wtmp, "in.wav")
                synth.write_wav(wav, host_station_iq(n, 0x5EED0002), SR)
                r = run_reference_cpu(secs, 1, wav_path=wav)
                cpu_baseline = {"value": r[0], "unit": "Msamples/s", "cores": 1, "kind": "reference",
                                "sample": "one process, %.0f s of 96 kHz station I/Q (WAV in tmpfs) through the reference's own "
                                          "Pipeline/WAVStream/WaterfallBackend/BolidRecorder (oracle/_ref, -O2), FITS to tmpfs. %s"
                                          % (n / SR, FFT_NOTE)}
            else:
                v = numpy_port_baseline(20.0)
                cpu_baseline = {"value": v, "unit": "Msamples/s", "cores": 1, "kind": "port",
                                "sample": "20 s of station I/Q through oracle/ref_numpy.py (numpy pocketfft)"}
            # the same path on scipy.fft, the as-shipped -O0 build, and what FFT library the box has
            note("cpu_baseline done: %r" % (cpu_baseline["value"],))
            cpu_alt = {"fft_probe": fft_probe()}
            try:
                iq120 = host_station_iq(120 * SR, 0x5EED0002)
                nproc = os.cpu_count() or 1
                t1, fr = scipy_path(iq120, BINS, OVERLAP, det, 1)
                tn, _ = scipy_path(iq120, BINS, OVERLAP, det, nproc)
                cpu_alt["scipy_fft"] = {"workers_1": 120 * SR / t1 / 1e6, "workers_all": 120 * SR / tn / 1e6, "cores": nproc,
                                        "unit": "Msamples/s", "sample": "120 s of station I/Q, framing + window + scipy.fft (pocketfft, "
                                        "complex128) + magnitude + shift + detector statistics in numpy, %d frames" % fr}
            except Exception as e:                  # noqa: BLE001
                cpu_alt["scipy_fft"] = {"error": str(e)}
            if os.path.exists(REF_BIN_O0):
                r0 = run_reference_cpu(300.0, 1, binary=REF_BIN_O0)
                if r0:
                    cpu_alt["reference_O0"] = {"value": r0[0], "unit": "Msamples/s", "cores": 1,
                                               "sample": "the same binary built with the reference's shipped flags (-O0 -ggdb, Makefile:26-30), 300 s of station I/Q"}
            # the same WAV through the reference's own Pipeline with one string changed in the config
            # ("factory": "waterfall_cuda"): what a station operator gets.  Process start-up and CUDA
            # initialisation (about 1.3 s) are inside every figure; each variant runs twice.
            note("cpu_baseline_alt done: %r" % (cpu_alt,))
            wf.close()
            if wav and os.path.exists(CUDA_BIN):
                def variant(**kw):
                    r = run_reference_cpu(secs, 1, runs=2, binary=CUDA_BIN, factory="waterfall_cuda", wav_path=wav, **kw)
                    return None if r is None else {"value": r[0], "unit": "Msamples/s", "cores": 1, "runs": r[1], "startup_crashes": r[3]}
                dropin = variant(flags=("-N",))
                if dropin:
                    dropin["what"] = ("reference Pipeline / recorders / FITSWriter / CsvLog, backend factory waterfall_cuda (bands + "
                                      "detector records from the GPU), frontend NarrowWAVStream (-N), same %.0f s WAV and outputs as cpu_baseline" % secs)
                    dropin["vs_cpu_baseline"] = dropin["value"] / cpu_baseline["value"]
                    w = variant()
                    if w:
                        w["what"] = "as above behind the reference's own WAVStream (vector<Complex> blocks, 16 B per sample)"
                        dropin["wavstream_frontend"] = w
                    f = variant(flags=("-N",), extra={"cuda_rows": "full"})
                    if f:
                        f["what"] = "cuda_rows = full: whole rows back to the host, statistics recomputed by the reference's BolidRecorder (round 1's adapter)"
                        dropin["full_rows_mode"] = f
        finally:
            shutil.rmtree(wtmp, ignore_errors=True)

    if rank == 0:
        line = {"metric": METRIC,
                "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": bench_config(args.log2_samples, P, G),
                "roofline": roofline, "parity": parity, "fp64": fp64, "cpu_baseline": cpu_baseline,
                "cpu_baseline_alt": cpu_alt, "dropin": dropin, "e2e": e2e, "clocks": clocks,
                "timed_region_s": t_wall1 - t_wall0,
                "gpu_launches": int(launches)}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def sweep(args, emit, torch, ro, dev):
    """BASELINE config 5: FFT size 1k..64k x overlap 0/50/75 % on synthetic int16 I/Q, device-resident,
    with the scipy.fft one-core figure of the same point beside it."""
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(peaks_path))["hbm_gbs"]) if os.path.exists(peaks_path) else 6650.0
    S = 1 << 26
    rng = np.random.default_rng(5)
    iq = rng.integers(-3000, 3000, size=2 * S, dtype=np.int16)
    d_iq = torch.from_numpy(iq).to(dev)
    d_rows = torch.empty(4 * S + (1 << 20), dtype=torch.float32, device=dev)
    points = []
    for lg in range(10, 17):
        N = 1 << lg
        _, det = bolidozor_bands(SR, N)
        for ov in (0, N // 2, 3 * N // 4):
            wf = ro.Waterfall(N, ov, sample_rate=SR, in_format=ro.IN_INT16)
            F = wf.frame_count(S)
            wf.process_resident(d_iq.data_ptr(), S, d_rows.data_ptr(), None, sync=True)
            wf.set_profiling(True)
            wf.kernel_times()
            for _ in range(5):
                wf.process_resident(d_iq.data_ptr(), S, d_rows.data_ptr(), None, sync=False)
            ms, n = wf.kernel_times()
            kname = wf.kernel_name()
            wf.close()
            t = ms[0] / max(1, n[0]) * 1e-3
            hop = N - ov
            alg = 4 * ((F - 1) * hop + N) + 4 * N * F
            cpu = None
            if not args.no_cpu_baseline:
                nf = max(8, min(64, (1 << 20) // N))
                tc, fr = scipy_path(iq[:2 * (N + (nf - 1) * hop)], N, ov, det, 1)
                cpu = fr * hop / tc / 1e6
            points.append({"bins": N, "overlap": ov, "frames": F, "kernel": kname, "kernel_ms": round(t * 1e3, 4),
                           "Mframes_per_s": round(F / t / 1e6, 3), "Msamples_per_s": round(F * hop / t / 1e6, 1),
                           "hbm_frac": round(alg / t / 1e9 / peak, 4), "fft_TFLOPs_5NlogN": round(5.0 * N * lg * F / t / 1e12, 2),
                           "cpu_scipy_fft_1core_Msamples_per_s": None if cpu is None else round(cpu, 3)})
            print(json.dumps(points[-1]), file=sys.stderr, flush=True)
    emit({"metric": "config 5: FFT size x overlap sweep, device-resident, int16 I/Q", "unit": "Msamples/s", "n_gpus": 1,
          "hbm_peak_gbs": peak, "points": points})
    return 0


if __name__ == "__main__":
    sys.exit(main())
