set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 200 python tools/sweep.py --lo 15 --hi 15 --fmt int16 > gpurun_out/r2_stcs_off_i16.jsonl 2>/dev/null
timeout 200 python tools/sweep.py --lo 15 --hi 15 --fmt float32 > gpurun_out/r2_stcs_off_f32.jsonl 2>/dev/null
RO_CUDA_LIB=$GRAFT_REPO_ROOT/radio_observer_b200/csrc/_obj/libro_cuda_stcs.so timeout 200 python tools/sweep.py --lo 15 --hi 15 --fmt int16 > gpurun_out/r2_stcs_on_i16.jsonl 2>/dev/null
RO_CUDA_LIB=$GRAFT_REPO_ROOT/radio_observer_b200/csrc/_obj/libro_cuda_stcs.so timeout 200 python tools/sweep.py --lo 15 --hi 15 --fmt float32 > gpurun_out/r2_stcs_on_f32.jsonl 2>/dev/null
for f in gpurun_out/r2_stcs_*; do echo $f; cut -c1-140 $f; done
timeout 600 python bench.py --sweep > gpurun_out/r2_sweep.json 2> gpurun_out/r2_sweep.err; echo sweep rc=$?
tail -3 gpurun_out/r2_sweep.err | cut -c1-300
timeout 300 python tools/sweep.py --lo 10 --hi 16 --fmt float32 > gpurun_out/r2_sweep_f32.jsonl 2>/dev/null
