set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_workloads_gpu.py -q -m gpu --timeout 600 2>&1 | tail -40
