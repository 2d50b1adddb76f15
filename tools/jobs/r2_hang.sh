set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out /dev/shm/hang/out
python - <<'P'
import sys, json
sys.path.insert(0,'.')
import bench
from radio_observer_b200 import synth
n=900*bench.SR
synth.write_wav('/dev/shm/hang/in.wav', bench.host_station_iq(n, 0x5EED0002), bench.SR)
open('/dev/shm/hang/cfg.json','w').write(bench.ref_config_json('/dev/shm/hang/out','waterfall_cuda',None))
P
cd /dev/shm/hang
for flags in "-N" ""; do
  rm -rf out; mkdir out
  ( time timeout 90 $GRAFT_REPO_ROOT/radio_observer_b200/host/_build/radio-observer-cuda -c cfg.json $flags in.wav > run.out 2> run.err ); echo "flags=[$flags] rc=$?"
  grep -c "Writing snapshot" run.err; grep -i "waiting\|warn\|error\|overwrit" run.err | grep -v METEOR | head -10; tail -4 run.err | cut -c1-200; ls out | wc -l
done
cp run.err $GRAFT_REPO_ROOT/gpurun_out/r2_hang.err
