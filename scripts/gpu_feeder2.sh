#!/bin/bash
# 2-GPU check of the native runtime: test, then end-to-end numbers of both feeders.
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_mlp_gpu.py -q -m gpu --timeout 400 -x -k "two_gpu or native" 2>&1 | tail -15
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
for f in native python; do
  timeout 300 $TR --master-port 29721 bench.py --gpus 2 --steps 30 --warmup 5 --feeder $f 2>&1 | tail -1 | tee gpurun_out/feeder2_wide_$f.json
  timeout 300 $TR --master-port 29722 bench.py --gpus 2 --config small --steps 400 --warmup 20 --feeder $f 2>&1 | tail -1 | tee gpurun_out/feeder2_small_$f.json
done
