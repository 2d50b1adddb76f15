#!/bin/bash
mkdir -p gpurun_out
export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_trace.so
echo "##### 64k, overlap 75 %: MODE_FUSED2 items (item 4j is class 0 with the load phase in front)"
timeout 200 python tools/trace16k.py --bins 65536 --overlap 49152 --json gpurun_out/trace_fused64k.json
