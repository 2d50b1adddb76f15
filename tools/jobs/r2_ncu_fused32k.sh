set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python tools/sweep.py --lo 15 --hi 15 --fmt int16 --reps 2"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wf_fft2 -s 7 -c 1 -o gpurun_out/r2_fused32k_i16 $CMD > gpurun_out/r2_ncu.log 2>&1
tail -3 gpurun_out/r2_ncu.log
