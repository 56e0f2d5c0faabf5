#!/bin/bash
# Secondary metrics at N GPUs (BASELINE.json configs): heat stencil 32768^2 order 8 with the
# fused halo push, PageRank 64 M nodes with the fused rank all-gather (column-blocked and plain).
# The big cases are timed without their host oracles; the same code paths are VERIFIED on cases
# the host can check in the same run (512 ULP stencil, 100 ULP PageRank).   usage: bench_secondary.sh N
N=${1:-8}
cd "$(dirname "$0")/.."
ROOT=$PWD
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
T=$(mktemp -d)
printf "4096 4096\n1 1\n20\n8\n" > $T/params.in
(cd $T && PYTHONPATH=$ROOT timeout 300 $TR --master-port 29800 -m cme213_sp18_b200.cli.heat -m params.in 2>&1 | grep -E "Slabs|ULP")
timeout 300 $TR --master-port 29801 -m cme213_sp18_b200.cli.heat -m --nx 32768 --ny 32768 --iters 20 --no-cpu 2>&1 | grep -E "Slabs|Order"
timeout 300 $TR --master-port 29803 -m cme213_sp18_b200.cli.pagerank --nodes 1048576 --avg 8 --block-nodes 262144 2>&1 | grep -E "^N="
timeout 300 $TR --master-port 29802 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check 2>&1 | grep -E "^N="
timeout 300 $TR --master-port 29804 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check --block-nodes 0 2>&1 | grep -E "^N="
rm -rf $T
