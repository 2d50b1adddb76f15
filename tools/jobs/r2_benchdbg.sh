set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python bench.py --passes 2 --shards 1 --steps 3 --no-cpu-baseline > gpurun_out/r2_benchdbg.json 2> gpurun_out/r2_benchdbg.err; echo rc=$?
cat gpurun_out/r2_benchdbg.err | tail -20
timeout 900 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo bench rc=$?
tail -30 gpurun_out/r2_bench.err
