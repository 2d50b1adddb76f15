set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_archive.py -x -q -m gpu 2>&1 | tail -2
for t in 1 4 8 16; do
  timeout 600 python tools/archive_day.py --hours 24 --keep-archive --copy-threads $t > gpurun_out/r2_archive_24h_n1_t$t.json 2> gpurun_out/r2_archive_n1.err; echo "threads=$t rc=$?"
  python -c "import json;j=json.loads(open('gpurun_out/r2_archive_24h_n1_t$t.json').read());print(j['copy_threads'], j['seconds_compute_and_merge'], j['Msamples_per_s_end_to_end'], j['seams_bit_identical'])"
done
rm -f /dev/shm/ro_archive_*.npy
