cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 120 -x -k "dw1 or tensor_core_tail" 2>&1 | tail -3
python scripts/exp_dw1.py
