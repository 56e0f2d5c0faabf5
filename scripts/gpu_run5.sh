set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi topo -m 2>&1 | head -8
timeout 600 python -m pytest tests/test_mlp_gpu.py -q -m multigpu --timeout 400 -x 2>&1 | tail -30
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 10 --warmup 3 2>&1 | tail -4 | tee gpurun_out/bench_wide_2.log
