#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2_final.json 2> gpurun_out/r2_bench_n2_final.err; echo "bench n=2 rc=$?"
tail -2 gpurun_out/r2_bench_n2_final.err | cut -c1-200
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 tools/archive_day.py --hours 24 > gpurun_out/r2_archive_24h_n2_final.json 2> gpurun_out/r2_archive_n2_final.err; echo "archive n=2 rc=$?"
cut -c1-300 gpurun_out/r2_archive_24h_n2_final.json
rm -f /dev/shm/ro_archive_*.npy
