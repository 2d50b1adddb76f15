set -x
cd $GRAFT_REPO_ROOT
timeout 300 python tools/stress_determinism.py 20
