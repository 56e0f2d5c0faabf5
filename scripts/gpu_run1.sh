set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
python -c "import torch;print(torch.__version__, torch.cuda.get_device_name(0))"
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x --timeout 120 2>&1 | tail -40
