cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_mlp_gpu.py -q -m "gpu and not multigpu" --timeout 120 -x 2>&1 | tail -3
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29601 bench.py --gpus 8 --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench_wide_8.log
timeout 300 $TR --nproc-per-node 8 --master-port 29602 bench.py --gpus 8 --steps 20 --warmup 5 --comm nccl 2>&1 | tail -1 | tee gpurun_out/bench_wide_8_nccl.log
timeout 300 $TR --nproc-per-node 8 --master-port 29603 bench.py --gpus 8 --config small --steps 50 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench_small_8.log
timeout 600 $TR --nproc-per-node 8 --master-port 29604 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>&1 | tail -1 | tee gpurun_out/ref_wide_8.log
timeout 300 python bench.py --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench_wide_1b.log
