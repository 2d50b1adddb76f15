#!/bin/bash
mkdir -p gpurun_out
export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_trace.so
for ov in 12288 0; do
echo "##### overlap $ov"
timeout 200 python tools/trace16k.py --overlap $ov --json gpurun_out/trace16k_ov$ov.json
done
echo "##### overlap 12288, no chain"
timeout 200 python tools/trace16k.py --overlap 12288 --tuning 4
