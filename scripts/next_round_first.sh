#!/bin/bash
# First 8-GPU call of the next round: the measurements this round could no longer afford.
#   gpurun --gpus 8 --timeout 600 -- 'bash scripts/next_round_first.sh'
# Everything lands in gpurun_out/next/ ; expected box time ~2 min (x8 charged).
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
O=gpurun_out/next
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
# 1. stencil 32768^2 with the TMA-fed slab kernel at 4 and 8 GPUs (r1: 114.5 k / 220.8 k GB/s without it)
for n in 4 8; do
  timeout 200 $TR --nproc-per-node $n --master-port $((29810+n)) -m cme213_sp18_b200.cli.heat -m --nx 32768 --ny 32768 \
      --iters 20 --no-cpu 2>&1 | grep -E "Slabs" | tee $O/heat_$n.log
done
CME_STENCIL_NO_TMA=1 timeout 200 $TR --nproc-per-node 8 --master-port 29820 -m cme213_sp18_b200.cli.heat -m --nx 32768 \
    --ny 32768 --iters 20 --no-cpu 2>&1 | grep -E "Slabs" | tee $O/heat_8_no_tma.log
# 2. headline at 8 GPUs with the native runtime feeding the end-to-end path, both configs
timeout 300 $TR --nproc-per-node 8 --master-port 29831 bench.py --gpus 8 --steps 30 --warmup 5 2>&1 | tail -1 | tee $O/wide_8.json | cut -c1-200
timeout 300 $TR --nproc-per-node 8 --master-port 29832 bench.py --gpus 8 --config small --steps 400 --warmup 20 2>&1 | tail -1 | tee $O/small_8.json | cut -c1-200
timeout 300 $TR --nproc-per-node 8 --master-port 29833 bench.py --gpus 8 --config small --steps 400 --warmup 20 --feeder python 2>&1 | tail -1 | tee $O/small_8_python_feeder.json | cut -c1-200
# 3. PageRank 64 M nodes at 8 GPUs (unchanged code; sanity)
timeout 200 $TR --nproc-per-node 8 --master-port 29841 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check 2>&1 | grep -E "^N=" | tee $O/pagerank_8.log
