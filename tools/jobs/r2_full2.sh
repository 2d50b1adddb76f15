set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo bench rc=$?
tail -12 gpurun_out/r2_bench.err | cut -c1-400
timeout 300 python bench.py --impl reference --steps 4 --warmup 3 > gpurun_out/r2_bench_ref.json 2>> gpurun_out/r2_bench.err; echo ref rc=$?
timeout 1700 python -m pytest tests -x -q -m gpu -s > gpurun_out/r2_gpu_tests.log 2>&1; echo rc=$? >> gpurun_out/r2_gpu_tests.log
grep -v "^met;" gpurun_out/r2_gpu_tests.log | grep "shipped\]\|dropin\|passed\|failed\|rc=" | tail -30 | cut -c1-300
