#!/bin/bash
mkdir -p gpurun_out
RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_b24.so timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "65536 or fused" 2>&1 | tail -2
for rep in 1 2; do
for v in prod b24; do
  if [ $v = prod ]; then unset RO_CUDA_LIB; else export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_$v.so; fi
  echo "## variant $v rep $rep"
  timeout 300 python tools/sweep.py --lo 16 --hi 16 --fmt int16 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
done
done
