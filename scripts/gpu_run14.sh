cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 2 --master-port 29621 bench.py --gpus 2 --steps 20 --warmup 5 --comm mc 2>&1 | tail -3 | tee gpurun_out/bench_wide_2_mc.log
timeout 300 $TR --nproc-per-node 2 --master-port 29622 bench.py --gpus 2 --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/bench_wide_2_p2p.log
