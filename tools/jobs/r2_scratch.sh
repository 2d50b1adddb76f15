#!/bin/bash
mkdir -p gpurun_out
for mb in 0 19 37 56 74 111 148 512; do
  echo "## scratch_mb $mb"
  timeout 300 python tools/sweep.py --lo 16 --hi 17 --fmt int16 --scratch-mb $mb | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
done
