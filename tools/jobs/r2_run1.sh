set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or gain or int16 or float32" > gpurun_out/r2_t1.log 2>&1; echo rc=$? >> gpurun_out/r2_t1.log
tail -5 gpurun_out/r2_t1.log
timeout 300 python tools/sweep.py --lo 15 --hi 16 --fmt int16 > gpurun_out/r2_sweep_i16_long.jsonl 2> gpurun_out/r2_sweep1.err
timeout 300 python tools/sweep.py --lo 15 --hi 16 --fmt float32 > gpurun_out/r2_sweep_f32_long.jsonl 2>> gpurun_out/r2_sweep1.err
cat gpurun_out/r2_sweep_i16_long.jsonl gpurun_out/r2_sweep_f32_long.jsonl
