#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
timeout 200 python tools/stress_determinism.py 2>&1 | tail -3
for rep in 1 2; do
for v in nobt bt; do
  if [ $v = bt ]; then unset RO_CUDA_LIB; else export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_$v.so; fi
  echo "## variant $v-float32 rep $rep"
  timeout 300 python tools/sweep.py --lo 16 --hi 16 --fmt float32 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
  echo "## variant $v-int16 rep $rep"
  timeout 300 python tools/sweep.py --lo 17 --hi 18 --fmt int16 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
done
done
