#!/bin/bash
# Secondary metrics at N GPUs (BASELINE.json configs): heat stencil 32768^2 order 8 with the
# fused halo push, PageRank 64 M nodes with the fused rank all-gather.  usage: bench_secondary.sh N
N=${1:-8}
cd "$(dirname "$0")/.."
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
timeout 300 $TR --master-port 29801 -m cme213_sp18_b200.cli.heat -m --nx 32768 --ny 32768 --iters 20 --no-cpu 2>&1 | grep -E "Slabs|Order"
timeout 300 $TR --master-port 29802 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check 2>&1 | grep -E "^N="
