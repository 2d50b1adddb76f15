#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for v in base es; do
  if [ $v = es ]; then export RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_es.so; else unset RO_CUDA_LIB; fi
  echo "## variant $v rep $rep"
  timeout 300 python tools/sweep.py --lo 10 --hi 16 --fmt int16 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
done
done
unset RO_CUDA_LIB
RO_CUDA_LIB=$PWD/radio_observer_b200/csrc/_obj/libro_cuda_es.so python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "rows_match_oracle_int16" 2>&1 | tail -3
