set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
nproc; free -g | head -2
for n in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench_n$n.json 2> gpurun_out/r2_bench_n$n.err; echo "bench n=$n rc=$?"
  tail -2 gpurun_out/r2_bench_n$n.err | cut -c1-200
done
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench n=1 rc=$?"
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n tools/archive_day.py --hours 24 --keep-archive > gpurun_out/r2_archive_24h_n$n.json 2> gpurun_out/r2_archive_n$n.err; echo "archive n=$n rc=$?"
  cat gpurun_out/r2_archive_24h_n$n.json | cut -c1-400
done
timeout 600 python tools/archive_day.py --hours 24 --keep-archive > gpurun_out/r2_archive_24h_n1.json 2> gpurun_out/r2_archive_n1.err; echo "archive n=1 rc=$?"
cat gpurun_out/r2_archive_24h_n1.json | cut -c1-400
rm -f /dev/shm/ro_archive_*.npy
