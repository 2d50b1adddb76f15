set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dropin_adapter.py tests/test_shipped_configs.py -x -q -m gpu > gpurun_out/r2_dropin_tests3.log 2>&1; echo rc=$? >> gpurun_out/r2_dropin_tests3.log
tail -3 gpurun_out/r2_dropin_tests3.log
python - <<'P' 2>&1 | tee gpurun_out/r2_dropin_timing.txt
import sys, os, time, tempfile, shutil
sys.path.insert(0, '.')
import bench
from radio_observer_b200 import synth
secs = 1800.0
n = int(secs * bench.SR); n -= n % 1024
tmp = tempfile.mkdtemp(dir='/dev/shm')
wav = os.path.join(tmp, 'in.wav')
synth.write_wav(wav, bench.host_station_iq(n, 0x5EED0002), bench.SR)
for name, kw in [("stock", dict()),
                 ("cuda narrow bands", dict(binary=bench.CUDA_BIN, factory="waterfall_cuda", flags=("-N",))),
                 ("cuda wavstream bands", dict(binary=bench.CUDA_BIN, factory="waterfall_cuda")),
                 ("cuda narrow full", dict(binary=bench.CUDA_BIN, factory="waterfall_cuda", flags=("-N",), extra={"cuda_rows": "full"})),
                 ("cuda narrow bands batch 512", dict(binary=bench.CUDA_BIN, factory="waterfall_cuda", flags=("-N",), extra={"cuda_batch_frames": 512}))]:
    r = bench.run_reference_cpu(secs, 1, runs=2, wav_path=wav, **kw)
    print(name, r[0], r[1], "crashes", r[3], "seconds", n / (r[0] * 1e6), flush=True)
shutil.rmtree(tmp)
P
