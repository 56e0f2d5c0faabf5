cd $GRAFT_REPO_ROOT
timeout 200 python -m pytest tests/test_gpu_kernels.py tests/test_mlp_gpu.py -q -m "gpu and not multigpu" --timeout 60 -x 2>&1 | tail -8
timeout 200 python scripts/profile_step.py --iters 3 2>&1 | grep -E "fwd1_tc|dw1|full_step|cublas"
CME_NO_PAIRS=1 timeout 200 python scripts/profile_step.py --iters 3 2>&1 | grep -E "fwd1_tc|dw1|full_step"
timeout 200 python scripts/profile_step.py --iters 3 --rows 8192 2>&1 | grep -E "fwd1_tc|dw1|full_step|cublas"
