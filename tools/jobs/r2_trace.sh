#!/bin/bash
mkdir -p gpurun_out
export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_trace.so
echo "##### 64k, overlap 75 %: MODE_SPLIT items"
timeout 200 python tools/trace16k.py --bins 65536 --overlap 49152 --json gpurun_out/trace_split64k.json
echo "##### 16k, overlap 75 %"
timeout 200 python tools/trace16k.py --overlap 12288 --json gpurun_out/trace16k_ov12288.json
