set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests.log 2>&1; echo rc=$? >> gpurun_out/r2_gpu_tests.log
tail -5 gpurun_out/r2_gpu_tests.log
timeout 600 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo bench rc=$?
cat gpurun_out/r2_bench.json
