set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_mlp_gpu.py -q -m "gpu and not multigpu" --timeout 300 -x 2>&1 | tail -30
timeout 600 python scripts/profile_step.py --out gpurun_out/step_wide_v2.json 2>&1 | tail -32
