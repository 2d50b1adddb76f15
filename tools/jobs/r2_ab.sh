set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for i in 1 2; do
timeout 200 python bench.py --no-cpu-baseline --steps 10 --passes 20 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('default', j['value'], j['roofline']['frac'], j['roofline']['kernel_ms'])"
RO_CUDA_LIB=$GRAFT_REPO_ROOT/radio_observer_b200/csrc/_obj/libro_cuda_stcs.so timeout 200 python bench.py --no-cpu-baseline --steps 10 --passes 20 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('stcs_all', j['value'], j['roofline']['frac'], j['roofline']['kernel_ms'])"
done
timeout 200 python tools/sweep.py --lo 12 --hi 14 --fmt int16 | cut -c1-150
RO_CUDA_LIB=$GRAFT_REPO_ROOT/radio_observer_b200/csrc/_obj/libro_cuda_stcs.so timeout 200 python tools/sweep.py --lo 12 --hi 14 --fmt int16 | cut -c1-150
