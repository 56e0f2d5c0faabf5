set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python bench.py --impl reference --config small --steps 10 --warmup 2 2>&1 | tail -3 | tee gpurun_out/ref_small_1.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -3 | tee gpurun_out/ref_wide_1.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --impl reference --config small --gpus 2 --steps 10 --warmup 2 2>&1 | tail -3 | tee gpurun_out/ref_small_2.log
