set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_dropin_adapter.py -x -q -m gpu -s -k "lap or fp64 or silent" > gpurun_out/r2_dropin_tests2.log 2>&1; echo rc=$? >> gpurun_out/r2_dropin_tests2.log
grep -v "^met;" gpurun_out/r2_dropin_tests2.log | tail -12
timeout 900 python bench.py > gpurun_out/r2_bench2.json 2> gpurun_out/r2_bench2.err; echo bench rc=$?
tail -5 gpurun_out/r2_bench2.err
cat gpurun_out/r2_bench2.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/r2_bench2_ref.json 2>> gpurun_out/r2_bench2.err; cat gpurun_out/r2_bench2_ref.json
