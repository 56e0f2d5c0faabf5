cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_mlp_gpu.py -q -m "gpu" --timeout 300 -x 2>&1 | tail -4
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for N in 8 4; do
timeout 300 $TR --nproc-per-node $N --master-port 2963$N bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_wide_${N}_d.log
timeout 300 $TR --nproc-per-node $N --master-port 2964$N bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --comm mc 2>&1 | tail -1 | tee gpurun_out/bench_wide_${N}_mc.log
done
timeout 300 $TR --nproc-per-node 8 --master-port 29651 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e --comm nccl 2>&1 | tail -1 | tee gpurun_out/bench_wide_8_nccl2.log
timeout 300 $TR --nproc-per-node 8 --master-port 29652 bench.py --gpus 8 --config small --steps 200 --warmup 10 --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_small_8_d.log
timeout 300 $TR --nproc-per-node 8 --master-port 29653 bench.py --gpus 8 --config small --steps 200 --warmup 10 --no-e2e --comm mc 2>&1 | tail -1 | tee gpurun_out/bench_small_8_mc.log
