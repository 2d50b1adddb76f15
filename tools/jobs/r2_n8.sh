#!/bin/bash
mkdir -p gpurun_out
for n in 8 4; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench_n${n}_final.json 2> gpurun_out/r2_bench_n${n}_final.err; echo "bench n=$n rc=$?"
tail -1 gpurun_out/r2_bench_n${n}_final.err | cut -c1-200
done
