#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_targets.py size 65536 49152 > gpurun_out/r2_ncu_plain_f2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wf_fft2 -s 2 -c 1 -o gpurun_out/r2b_fused64k python tools/ncu_targets.py size 65536 49152 > gpurun_out/r2_ncu_f2.log 2>&1
tail -2 gpurun_out/r2_ncu_f2.log
