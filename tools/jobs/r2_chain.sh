#!/bin/bash
# A/B of the chain variants (stagger mask through RO_TUNE_STAGGER_SET), 16k and 32k, three overlaps each
mkdir -p gpurun_out
for rep in 1 2; do
for st in 14 46 38 78 110 102 126 30 62; do
  tun=$(( 0x10000 | (st << 8) ))
  echo "## stagger $st rep $rep"
  timeout 120 python tools/sweep.py --lo 14 --hi 15 --fmt int16 --tuning $tun | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print(j.get('bins'), j.get('overlap'), round(j.get('Mframes_per_s', 0), 3), round(j.get('hbm_frac', 0), 4))
"
done
done
