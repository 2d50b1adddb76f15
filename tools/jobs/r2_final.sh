#!/bin/bash
# final verification of the committed state: GPU tests (the no-alloc file last), that file alone in a
# fresh process, smoke, bench, sweep
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_final.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/r2_gpu_tests_final.log
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu -s > gpurun_out/r2_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -3 gpurun_out/r2_no_alloc_alone.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"
python bench.py --sweep > gpurun_out/r2_sweep_final.json 2> gpurun_out/r2_sweep_final.err; echo "sweep rc=$?"
