#!/bin/bash
mkdir -p gpurun_out
run() {  # name, kernel regex, skip, command...
  name=$1; shift; k=$1; shift; skip=$1; shift
  "$@" > gpurun_out/r2_ncu_plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o gpurun_out/r2c_$name "$@" > gpurun_out/r2_ncu_$name.log 2>&1
  tail -1 gpurun_out/r2_ncu_$name.log
}
run fused64k_i16 wf_fft2 2 python tools/ncu_targets.py size 65536 49152
run split128k_fft wf_fft2 2 python tools/ncu_targets.py size 131072 98304
