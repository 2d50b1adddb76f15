set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python tools/sweep.py --lo 10 --hi 16 --fmt int16 | cut -c1-110
timeout 100 python tools/stress_determinism.py 8 | tail -3
