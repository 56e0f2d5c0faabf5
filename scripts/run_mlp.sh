#!/bin/bash
# Job script of the MLP trainer (ref fpcode/run.sh:38-39: mpirun -np 4 ./main -g 2).
# usage: scripts/run_mlp.sh [NPROC] [mlp flags...]   e.g. scripts/run_mlp.sh 4 -g 2
NP=${1:-4}; shift
cd "$(dirname "$0")/.."
mkdir -p Outputs/CPUmats data          # ref fpcode/init.sh creates these; MNIST itself cannot be fetched offline
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NP" --master-addr 127.0.0.1 \
     --master-port "${MASTER_PORT:-29500}" -m cme213_sp18_b200.cli.mlp "${@:--g 2}"
