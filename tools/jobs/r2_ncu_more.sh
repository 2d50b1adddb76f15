set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() {  # name, kernel regex, skip, command...
  name=$1; shift; k=$1; shift; skip=$1; shift
  "$@" > gpurun_out/r2_ncu_plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o gpurun_out/r2_$name "$@" > gpurun_out/r2_ncu_$name.log 2>&1
  tail -2 gpurun_out/r2_ncu_$name.log
}
run detect wf_detect 2 python tools/ncu_targets.py detect
run reduce wf_reduce 2 python tools/ncu_targets.py reduce
run split64k_split wf_split 2 python tools/ncu_targets.py size 65536 49152
run split64k_fft wf_fft2 2 python tools/ncu_targets.py size 65536 49152
run fft4k wf_fft2 2 python tools/ncu_targets.py size 4096 3072
run fft8k wf_fft2 2 python tools/ncu_targets.py size 8192 6144
run fft16k wf_fft2 2 python tools/ncu_targets.py size 16384 12288
run fft16k_ov0 wf_fft2 2 python tools/ncu_targets.py size 16384 0
ls -la gpurun_out/*.ncu-rep
