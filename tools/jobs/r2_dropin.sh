set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_dropin_adapter.py -x -q -m gpu -s > gpurun_out/r2_dropin_tests.log 2>&1; echo rc=$? >> gpurun_out/r2_dropin_tests.log
grep -v "^met;" gpurun_out/r2_dropin_tests.log | tail -40
