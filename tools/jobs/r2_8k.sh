set -x
cd $GRAFT_REPO_ROOT
timeout 200 python tools/sweep.py --lo 11 --hi 14 --fmt int16 | cut -c1-110
timeout 200 python tools/sweep.py --lo 11 --hi 14 --fmt int16 --tuning 4 | cut -c1-110
